"""Regenerates tests/golden/*.json from the big-int Python oracle (oracle/*.py).

The reference (a Go tree) cannot run here and holds no MSM/FFT/proof vectors for this path, so
these fixtures pin the *restated* algorithms: they were produced by the Python oracle, which is independent
of the C oracle and of the CUDA code, and they are checked against both.  Run from the repo root:
    python tests/golden/make_golden.py
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import bn254 as C          # noqa: E402
from oracle import groth16 as og       # noqa: E402
from oracle import msm as omsm         # noqa: E402
from oracle import ntt as ontt         # noqa: E402
from oracle import r1cs as orcs        # noqa: E402
from oracle.prng import Xoshiro, SEED_BASE  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
hx = lambda b: b.hex()


def cubic():
    cs = orcs.cubic_circuit()
    w = orcs.cubic_witness(3, 35)
    rng = Xoshiro(SEED_BASE + 1)
    pk, vk, scal = og.setup(cs, rng)
    r, s = rng.field(C.R), rng.field(C.R)
    parts = og.prove_parts(cs, pk, w)
    proof = og.assemble(pk, parts, r, s)
    assert og.verify(proof, vk, [35])
    a, b, c = cs.abc(w)
    return dict(
        witness=hx(C.fr_vec_to_bytes(w)), a=hx(C.fr_vec_to_bytes(a)), b=hx(C.fr_vec_to_bytes(b)), c=hx(C.fr_vec_to_bytes(c)),
        r=hx(C.fr_to_bytes(r)), s=hx(C.fr_to_bytes(s)), domain_n=pk.domain.n,
        infinity_a=[int(x) for x in pk.infinity_a], infinity_b=[int(x) for x in pk.infinity_b], nb_public=pk.nb_public,
        g1_alpha=hx(C.g1_to_bytes(pk.g1_alpha)), g1_beta=hx(C.g1_to_bytes(pk.g1_beta)), g1_delta=hx(C.g1_to_bytes(pk.g1_delta)),
        g2_beta=hx(C.g2_to_bytes(pk.g2_beta)), g2_delta=hx(C.g2_to_bytes(pk.g2_delta)),
        g1_a=hx(C.g1_vec_to_bytes(pk.g1_a)), g1_b=hx(C.g1_vec_to_bytes(pk.g1_b)), g1_k=hx(C.g1_vec_to_bytes(pk.g1_k)),
        g1_z=hx(C.g1_vec_to_bytes(pk.g1_z)), g2_b=hx(C.g2_vec_to_bytes(pk.g2_b)),
        h=hx(C.fr_vec_to_bytes(parts["h"])),
        msm_points=hx(C.g1_to_bytes(parts["ar"]) + C.g1_to_bytes(parts["bs1"]) + C.g1_to_bytes(parts["krs_k"])
                      + C.g1_to_bytes(parts["krs_hz"]) + C.g2_to_bytes(parts["bs2"])),
        proof=hx(og.proof_bytes(proof)))


def msm_vectors():
    rng = Xoshiro(SEED_BASE + 2)
    n = 12
    ks = [rng.field(C.R) for _ in range(n)]
    sc = [rng.field(C.R) for _ in range(n)]
    sc[0], sc[1], sc[2] = 0, 1, C.R - 1
    p1 = [C.g1_mul(C.G1_GEN, k) for k in ks]
    p2 = [C.g2_mul(C.G2_GEN, k) for k in ks]
    p1[3] = None                                  # infinity base
    return dict(scalars=hx(C.fr_vec_to_bytes(sc)), g1_points=hx(C.g1_vec_to_bytes(p1)), g2_points=hx(C.g2_vec_to_bytes(p2)),
                g1_result=hx(C.g1_to_bytes(omsm.msm_naive(p1, sc))), g2_result=hx(C.g2_to_bytes(omsm.msm_naive(p2, sc, g2=True))))


def ntt_vectors():
    rng = Xoshiro(SEED_BASE + 3)
    n = 32
    d = ontt.Domain(n)
    x = [rng.field(C.R) for _ in range(n)]
    a, b = x[:20], [rng.field(C.R) for _ in range(20)]
    c = [u * v % C.R for u, v in zip(a, b)]
    return dict(n=n, x=hx(C.fr_vec_to_bytes(x)),
                inv_dif=hx(C.fr_vec_to_bytes(ontt.fft_inverse_dif(x, d))),
                dit_coset=hx(C.fr_vec_to_bytes(ontt.fft_dit(x, d, coset=True))),
                inv_dif_coset=hx(C.fr_vec_to_bytes(ontt.fft_inverse_dif(x, d, coset=True))),
                dit=hx(C.fr_vec_to_bytes(ontt.fft_dit(x, d))),
                h_a=hx(C.fr_vec_to_bytes(a)), h_b=hx(C.fr_vec_to_bytes(b)), h_c=hx(C.fr_vec_to_bytes(c)),
                h=hx(C.fr_vec_to_bytes(ontt.compute_h(a, b, c, d))))


if __name__ == "__main__":
    for name, fn in (("cubic_proof", cubic), ("msm_small", msm_vectors), ("ntt_small", ntt_vectors)):
        with open(os.path.join(OUT, name + ".json"), "w") as f:
            json.dump(fn(), f, indent=1)
        print("wrote", name)
