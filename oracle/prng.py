"""xoshiro256** seeded by splitmix64 (oracle / harness PRNG).

The same generator is implemented in oracle/c/oracle_cpu.c so large inputs are
regenerated on the GPU box instead of shipped.
"""
M64 = (1 << 64) - 1
SEED_BASE = 0x6731364232303000


def _rotl(x, k):
    return ((x << k) | (x >> (64 - k))) & M64


class Xoshiro:
    def __init__(self, seed):
        s = seed & M64
        st = []
        for _ in range(4):
            s = (s + 0x9E3779B97F4A7C15) & M64
            z = s
            z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M64
            z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M64
            st.append(z ^ (z >> 31))
        self.s = st

    def next(self):
        s = self.s
        result = (_rotl((s[1] * 5) & M64, 7) * 9) & M64
        t = (s[1] << 17) & M64
        s[2] ^= s[0]
        s[3] ^= s[1]
        s[1] ^= s[2]
        s[0] ^= s[3]
        s[2] ^= t
        s[3] = _rotl(s[3], 45)
        return result

    def u256(self):
        v = 0
        for i in range(4):
            v |= self.next() << (64 * i)
        return v

    def field(self, modulus):
        """4 draws -> 256-bit integer mod modulus."""
        return self.u256() % modulus
