// Host-side check of Field::inv_binary (csrc/field.cuh) against the Fermat inverse and against a * a^-1 = 1, for the
// base and scalar fields of the curve selected with -DG16_CURVE (254 / 377).  field.cuh compiles as plain C++ for the
// host epilogue; tests/test_host_field.py builds this file with g++ and runs it (no GPU).
#include "field.cuh"
#include <cstdio>
#include <cstdlib>
#include <random>
using namespace g16;
template <class F, class PP> int run(const char* name, int iters) {
    std::mt19937_64 g(7);
    int bad = 0, maxit = 0;
    for (int k = 0; k < iters; k++) {
        typename F::T a;
        for (int i = 0; i < F::N; i++) a.l[i] = (uint32_t)g();
        if (k == 0) a = F::zero();
        if (k == 1) a = F::one();
        if (k == 2) { a = F::zero(); a.l[0] = 1; }          // image with integer value 1
        if (k == 3) { a = F::zero(); a.l[0] = 2; }
        if (k == 4) { for (int i = 0; i < F::N; i++) a.l[i] = PP::mod(i); a.l[0] -= 1; }   // p - 1
        a.l[F::N - 1] &= 0x0fffffffu >> (F::N == 12 ? 4 : 0);
        a = F::reduce_once(a);
        typename F::T x = F::inv(a), y = F::inv_binary(a);
        y = F::reduce_once(y);
        if (memcmp(&x, &y, sizeof x)) { bad++; if (bad < 4) printf("%s: mismatch at %d\n", name, k); }
        typename F::T pr = F::mul(a, y);
        typename F::T one = F::one();
        if (k && memcmp(&pr, &one, sizeof pr)) { bad++; if (bad < 4) printf("%s: a*inv != 1 at %d\n", name, k); }
    }
    printf("%s: %d cases, %d bad\n", name, iters, bad);
    return bad;
}
int main() {
    int bad = 0;
    bad += run<Field<FpParams>, FpParams>("Fp", 3000);
    bad += run<Field<FrParams>, FrParams>("Fr", 3000);
    return bad != 0;
}
