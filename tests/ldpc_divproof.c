/* ldpc_divproof.c -- exhaustive proof obligation for the LDPC kernel's division-free platanh
 * (msk144cudecoder_b200/csrc/kernels.cuh, platanh2).  Test infrastructure, not product code.
 *
 * The reference's platanh (ldpc_kernel.cuh:65-93) computes sign * (z - c_k) / d_k with an IEEE division.
 * The kernel computes, with r_k = RN(1 / d_k):  q0 = RN(a r),  rem = a - q0 d (one FMA),  q = RN(q0 + rem r).
 * For every float z of every segment this program compares q with the IEEE quotient, and also checks the
 * increment chains that produce c_k, d_k, r_k from the indicator values.  Prints, per segment, the number
 * of mismatches and the largest mismatching z; exit code 0 iff all mismatches lie below argv[1] (the
 * kernel's LDPC_TINY guard, default 1e-30). */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

int main(int argc, char** argv) {
    const float tiny = argc > 1 ? (float)atof(argv[1]) : 1e-30f;
    const float cs[4] = {0.0f, 0.4064f, 0.8378f, 0.9914f}, ds[4] = {0.83f, 0.322f, 0.0524f, 0.0012f};
    const float lo[4] = {0.0f, 0.664f, 0.9217f, 0.9951f}, hi[4] = {0.664f, 0.9217f, 0.9951f, 0.9998f};
    int fail = 0;
    for (int k = 0; k < 4; k++) {
        const float d = ds[k], c = cs[k];
        volatile float rv = 1.0f / d;
        const float r = rv;
        const uint32_t a0 = f2u(lo[k]), a1 = f2u(hi[k]);
        long bad = 0;
        uint32_t lastbad = 0;
#pragma omp parallel for reduction(+ : bad) reduction(max : lastbad)
        for (uint32_t zb = a0; zb <= a1; zb++) {
            const float z = u2f(zb);
            if (k > 0 && !(z > lo[k])) continue;
            volatile float av = z - c;
            const float a = av;
            volatile float ref = a / d;
            volatile float q0 = a * r;
            const float rem = fmaf(-q0, d, a);
            const float q = fmaf(rem, r, q0);
            if (f2u(q) != f2u(ref)) {
                bad++;
                if (zb > lastbad) lastbad = zb;
            }
        }
        printf("segment %d d=%g: mismatches=%ld largest_z=%g\n", k, d, bad, bad ? u2f(lastbad) : 0.0f);
        if (bad && !(u2f(lastbad) < tiny)) fail = 1;
    }
    return fail;
}
