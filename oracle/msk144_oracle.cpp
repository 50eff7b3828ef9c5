/*
 * msk144_oracle.cpp -- CPU restatement of the msk144cudecoder hot path.
 *
 * TEST INFRASTRUCTURE ONLY (see msk144_oracle.h).  Plain scalar C++/FP32 that follows the
 * reference's algorithm and, in scan_mode 0, its order of floating-point operations.  What it
 * cannot reproduce bit-for-bit: the GPU's FMA contraction choices and the device libm
 * (sincosf/tanhf/hypotf/atan2f differ from glibc in the last ulp).  Bit-exact anchoring is done
 * against golden outputs of the real reference run on a B200 (tests/golden/, oracle/ref_dump.cu).
 *
 * Build: see oracle/Makefile (g++ -O2 -ffp-contract=off -pthread).
 */
#include "msk144_oracle.h"
#include "ldpc_128_90_nm.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <set>
#include <string>
#include <vector>
#include <atomic>
#include <functional>
#include <thread>

namespace {

// Dynamic-schedule parallel loop on std::thread (no OpenMP runtime needed on the test boxes).
void parallel_for(long long n, int nthreads, const std::function<void(long long)>& fn) {
    if (nthreads <= 0) nthreads = (int)std::thread::hardware_concurrency();
    if (nthreads < 1) nthreads = 1;
    if ((long long)nthreads > n) nthreads = (int)(n > 0 ? n : 1);
    if (nthreads == 1) {
        for (long long i = 0; i < n; i++) fn(i);
        return;
    }
    std::atomic<long long> next{0};
    std::vector<std::thread> pool;
    for (int t = 0; t < nthreads; t++)
        pool.emplace_back([&]() {
            for (;;) {
                const long long i = next.fetch_add(1);
                if (i >= n) break;
                fn(i);
            }
        });
    for (auto& th : pool) th.join();
}

constexpr int N = ORC_NSAMP;       // common.h:15  Num6x864
constexpr int NFR = ORC_NFRAME;    // common.h:14
constexpr int N42 = 42;            // common.h:17
constexpr int SECOND_SYNC = (8 + 48) * 6;  // common.h:20  = 336
constexpr int NCAND = 8;           // common.h:34
constexpr int NITER = 10;          // common.h:29
constexpr float FS = 12000.0f;     // common.h:47
constexpr float PI_F = 3.141592654f;  // CUDART_PI_F

struct C {
    float re, im;
};
// smath_complex.h:48-71
inline C cadd(C x, C y) { return {x.re + y.re, x.im + y.im}; }
inline C cmul(C x, C y) { return {x.re * y.re - x.im * y.im, x.re * y.im + x.im * y.re}; }
inline C smul(float m, C z) { return {m * z.re, m * z.im}; }
inline C cconj(C z) { return {z.re, -z.im}; }
inline float cabs_(C z) { return hypotf(z.re, z.im); }

struct Tables {
    float pp[12];
    int s8[8];
    C cb42[42];
    int mask[8][6];
    int num_avg[8];
    Tables() {
        // msk_context.cuh:137-145
        for (size_t i = 0; i < 12; i++) {
            float angle = i * PI_F / 12.0f;
            pp[i] = sinf(angle);
        }
        // msk_context.cuh:147-154
        const int s8_org[8] = {0, 1, 1, 1, 0, 0, 1, 0};
        for (int i = 0; i < 8; i++) s8[i] = 2 * s8_org[i] - 1;
        // msk_context.cuh:176-197
        float cbi[42], cbq[42];
        for (int i = 0; i < 6; i++) cbq[0 + i] = pp[6 + i] * s8[0];
        for (int i = 0; i < 12; i++) cbq[6 + i] = pp[i] * s8[2];
        for (int i = 0; i < 12; i++) cbq[18 + i] = pp[i] * s8[4];
        for (int i = 0; i < 12; i++) cbq[30 + i] = pp[i] * s8[6];
        for (int i = 0; i < 12; i++) cbi[0 + i] = pp[i] * s8[1];
        for (int i = 0; i < 12; i++) cbi[12 + i] = pp[i] * s8[3];
        for (int i = 0; i < 12; i++) cbi[24 + i] = pp[i] * s8[5];
        for (int i = 0; i < 6; i++) cbi[36 + i] = pp[i] * s8[7];
        for (int i = 0; i < 42; i++) cb42[i] = {cbi[i], cbq[i]};
        // msk_context.cuh:229-240
        const int m[8][6] = {{1, 0, 0, 0, 0, 0}, {1, 1, 0, 0, 0, 0}, {1, 1, 1, 0, 0, 0}, {1, 1, 1, 1, 0, 0},
                             {1, 1, 1, 1, 1, 0}, {1, 1, 1, 1, 1, 1}, {1, 0, 0, 1, 0, 0}, {1, 0, 0, 1, 1, 0}};
        for (int p = 0; p < 8; p++) {
            int w = 0;
            for (int i = 0; i < 6; i++) {
                mask[p][i] = m[p][i];
                w += m[p][i];
            }
            num_avg[p] = w;  // pattern_item.h:33-50
        }
    }
};
const Tables T;

// ---------------------------------------------------------------- frequency grid
// msk_context.cuh:95-113
struct Grid {
    int nf;
    float if1, step, center;
};
Grid make_grid(float center, float width, float step) {
    float half_len_in_hz = width / 2;
    float half_len_cnt = half_len_in_hz / step;
    int half_len = static_cast<int>(half_len_cnt);
    Grid g;
    g.nf = half_len * 2 + 1;
    g.if1 = -1 * half_len * step;
    g.step = step;
    g.center = center;
    return g;
}
// msk_context.cuh:135  getFrequencyForThread
inline float grid_freq(const Grid& g, int idx) { return g.center + g.if1 + idx * g.step; }

int clamp_depth(int d) {  // msk_context.cuh:28-33
    if (d < 1) d = 1;
    if (d > 8) d = 8;
    return d;
}

// ---------------------------------------------------------------- mixing
// scan_kernel.cuh:45-69 == softbits_kernel.cuh:28-54
void mix_block(const C* x, float freq, C* cdat2) {
    const float f0 = -1 * freq;
    const float twopi = 2.0f * PI_F;
    for (int i = 0; i < N; i++) {
        const float phi = static_cast<float>(i) * twopi * f0 / FS;
        float s, c;
        sincosf(phi, &s, &c);  // smath_complex.h:81-87
        cdat2[i] = cmul(C{c, s}, x[i]);
    }
}

// ---------------------------------------------------------------- scan
struct Best {
    unsigned pos;
    float xb;
};

// scan_kernel.cuh:276-353: replace the minimum-xb slot (ties -> lowest slot) if strictly greater
inline void keep_best(Best* best, float xb, unsigned pos) {
    int worst = 0;
    for (int i = 1; i < NCAND; i++)
        if (best[i].xb < best[worst].xb) worst = i;
    if (xb > best[worst].xb) {
        best[worst].xb = xb;
        best[worst].pos = pos;
    }
}

// Brute force in the reference's summation order.  scan_kernel.cuh:71-364
void scan_freq_reforder(const C* cdat2, int depth, Best out[][NCAND]) {
    for (int p = 0; p < depth; p++) {
        Best* best = out[p];
        for (int i = 0; i < NCAND; i++) best[i] = {0u, 0.0f};
        for (int slice = 0; slice < 21; slice++) {
            float xb_best = -1.0f;
            unsigned pos_best = 0;
            for (int tid = 0; tid < 256; tid++) {
                const unsigned base = slice * 256 + tid;
                C s{0.0f, 0.0f};
                for (unsigned idx = 0; idx < N42; idx++) {
                    const unsigned long_idx = base + idx;
                    C y{0.0f, 0.0f};
                    for (unsigned m = 0; m < 6; m++) {
                        if (T.mask[p][m]) {
                            unsigned ia = long_idx + NFR * m;
                            if (ia >= (unsigned)N) ia -= N;
                            y = cadd(y, cdat2[ia]);
                            unsigned ib = long_idx + NFR * m + SECOND_SYNC;
                            if (ib >= (unsigned)N) ib -= N;
                            y = cadd(y, cdat2[ib]);
                        }
                    }
                    s = cadd(s, cmul(cconj(y), T.cb42[idx]));
                }
                const float xb = cabs_(s);
                // warp + inter-warp max with strict '>' (scan_kernel.cuh:143-268): ties keep lowest pos
                if (tid == 0 || xb > xb_best) {
                    xb_best = xb;
                    pos_best = base;
                }
            }
            keep_best(best, xb_best, pos_best);
        }
    }
}

// Linear reformulation (same metric, different FP32 summation order): one 42-tap correlation per
// position, sync pair, then per-pattern frame sums.  Used for whole streams.
void scan_freq_fast(const C* cdat2, int depth, Best out[][NCAND]) {
    std::vector<C> r(N), q(N);
    for (int p = 0; p < N; p++) {
        C s{0.0f, 0.0f};
        if (p + N42 <= N) {
            for (int k = 0; k < N42; k++) s = cadd(s, cmul(cconj(cdat2[p + k]), T.cb42[k]));
        } else {
            for (int k = 0; k < N42; k++) s = cadd(s, cmul(cconj(cdat2[(p + k) % N]), T.cb42[k]));
        }
        r[p] = s;
    }
    for (int p = 0; p < N; p++) q[p] = cadd(r[p], r[(p + SECOND_SYNC) % N]);
    for (int pt = 0; pt < depth; pt++) {
        Best* best = out[pt];
        for (int i = 0; i < NCAND; i++) best[i] = {0u, 0.0f};
        for (int slice = 0; slice < 21; slice++) {
            float xb_best = -1.0f;
            unsigned pos_best = 0;
            for (int tid = 0; tid < 256; tid++) {
                const unsigned base = slice * 256 + tid;
                const unsigned p0 = base >= (unsigned)N ? base - N : base;
                C s{0.0f, 0.0f};
                for (int m = 0; m < 6; m++)
                    if (T.mask[pt][m]) s = cadd(s, q[(p0 + NFR * m) % N]);
                const float xb = cabs_(s);
                if (tid == 0 || xb > xb_best) {
                    xb_best = xb;
                    pos_best = base;
                }
            }
            keep_best(best, xb_best, pos_best);
        }
    }
}

// ---------------------------------------------------------------- reductions (sum_reduction.cuh:13-44)
// Literal emulation of the two-round shuffle tree for nthr threads (values beyond nthr do not exist).
template <typename TT>
TT sum_reduction_two_cycles(const std::vector<TT>& vals) {
    const int nthr = (int)vals.size();
    const int nwarps = (nthr + 31) / 32;
    TT shared[32];
    for (int i = 0; i < 32; i++) shared[i] = 0;
    for (int w = 0; w < nwarps; w++) {
        TT v[32];
        for (int l = 0; l < 32; l++) v[l] = (w * 32 + l < nthr) ? vals[w * 32 + l] : 0;
        for (int off = 1; off < 32; off <<= 1) {
            TT nv[32];
            for (int l = 0; l < 32; l++) nv[l] = v[l] + (l + off < 32 ? v[l + off] : v[l]);
            std::memcpy(v, nv, sizeof(v));
        }
        shared[w] = v[0];
    }
    TT v[32];
    for (int l = 0; l < 32; l++) v[l] = shared[l];
    for (int off = 1; off < 32; off <<= 1) {
        TT nv[32];
        for (int l = 0; l < 32; l++) nv[l] = v[l] + (l + off < 32 ? v[l + off] : v[l]);
        std::memcpy(v, nv, sizeof(v));
    }
    return v[0];
}

// ---------------------------------------------------------------- softbits (softbits_kernel.cuh:56-241)
void softbits_from_mixed(const C* cdat2, int pattern_idx, unsigned pos, float* llr128, int* nbadsync_out,
                         float* sb_out) {
    C cdat3[NFR];
    // :59-83 frame average with single wrap.  For pos in [5185,5375] and the 6th frame the reference
    // indexes past the end of its shared array (undefined); we wrap a second time instead.
    for (int i = 0; i < NFR; i++) {
        C s{0.0f, 0.0f};
        for (unsigned m = 0; m < 6; m++) {
            if (T.mask[pattern_idx][m]) {
                unsigned k = pos + i + NFR * m;
                if (k >= (unsigned)N) k -= N;
                if (k >= (unsigned)N) k -= N;
                s = cadd(s, cdat2[k]);
            }
        }
        cdat3[i] = s;
    }
    // :88-137 carrier phase from both sync words, reduction order 42 -> +42 -> 32 -> tree
    C red[N42];
    for (int t = 0; t < N42; t++) red[t] = cmul(cdat3[t], cconj(T.cb42[t]));
    for (int t = 0; t < N42; t++) red[t] = cadd(red[t], cmul(cdat3[SECOND_SYNC + t], cconj(T.cb42[t])));
    for (int t = 0; t < 10; t++) red[t] = cadd(red[t], red[32 + t]);
    for (int size = 16; size > 0; size /= 2)
        for (int t = 0; t < size; t++) red[t] = cadd(red[t], red[t + size]);
    const C s = red[0];
    // :143-154
    const float phase0 = atan2f(s.im, s.re);
    float sn, cs;
    sincosf(phase0, &sn, &cs);
    const C cfac = cconj(C{cs, sn});
    for (int i = 0; i < NFR; i++) cdat3[i] = cmul(cdat3[i], cfac);
    // :157-181 matched filter, QIQI...
    float sb[144];
    for (int t = 0; t < 144; t++) {
        const int pos_iq = t % 72;
        const int iq = t / 72;
        const int base1 = iq ? 0 : (NFR - 6);
        const int d = 12 * pos_iq;
        float acc = 0.0f;
        for (int j = 0; j < 12; j++) {
            const int k = (base1 + d + j) % NFR;
            const float v = iq ? cdat3[k].re : cdat3[k].im;
            acc = acc + v * T.pp[j];
        }
        sb[pos_iq * 2 + iq] = acc;
    }
    if (sb_out) std::memcpy(sb_out, sb, sizeof(sb));
    // :186-201 normalisation (160 threads, 144 carry data)
    std::vector<float> a(160, 0.0f), a2(160, 0.0f);
    for (int t = 0; t < 144; t++) {
        a[t] = sb[t];
        a2[t] = sb[t] * sb[t];
    }
    const float sum_sav = sum_reduction_two_cycles(a);
    const float sum_s2av = sum_reduction_two_cycles(a2);
    const float sav = (float)(sum_sav / 144.0);
    const float s2av = (float)(sum_s2av / 144.0);
    const float ssig = sqrtf(s2av - sav * sav);
    const float sigma = 0.60f;
    const float scale = 2.0f / (ssig * sigma * sigma);
    // :204-212
    for (int t = 0; t < 48; t++) llr128[t] = scale * sb[8 + t];
    for (int t = 0; t < 80; t++) llr128[48 + t] = scale * sb[8 + 48 + 8 + t];
    // :214-241
    int nbad = 0;
    for (int grp = 0; grp < 2; grp++) {
        const int base = grp ? 0 : 56;  // thread bit 3 set -> first sync block
        int v = 0;
        for (int b = 0; b < 8; b++) {
            const int hard = (sb[base + b] < 0.0f) ? -1 : 1;
            v += hard * T.s8[b];
        }
        nbad += (8 - v) / 2;
    }
    *nbadsync_out = nbad;
}

// ---------------------------------------------------------------- LDPC
struct LdpcTables {
    // bit -> (slot, check) for its three edges in reference order (ldpc_context.cuh:10-139)
    int slot[128][3], chk[128][3];
    int nrw[38];
    uint16_t crc_table[256];
    uint8_t gen[38][90];  // parity = gen * msg90 over GF(2)
    LdpcTables() {
        int cnt[128] = {0};
        for (int c = 0; c < 38; c++) {
            nrw[c] = 0;
            for (int s = 0; s < 11; s++) {
                const int b = ORC_NM[c][s];
                if (b < 0) continue;
                nrw[c]++;
                slot[b][cnt[b]] = s;
                chk[b][cnt[b]] = c;
                cnt[b]++;
            }
        }
        // ldpc_context.cuh:185-213 (table for the augmented CRC-13, poly 0x15D7)
        for (int i = 0; i < 256; i++) {
            uint16_t dividend = (uint16_t)i, rem = 0;
            for (int bit = 0; bit < 8; bit++) {
                if (dividend & 0x80) rem ^= (1 << 12);
                const bool q = rem & (1 << 12);
                rem <<= 1;
                if (q) rem ^= 0x15D7;
                dividend <<= 1;
            }
            crc_table[i] = rem;
        }
        // systematic generator: H = [A | B], parity = B^-1 A m  (Gauss-Jordan over GF(2))
        uint8_t M[38][128];
        std::memset(M, 0, sizeof(M));
        for (int c = 0; c < 38; c++)
            for (int s = 0; s < 11; s++)
                if (ORC_NM[c][s] >= 0) M[c][ORC_NM[c][s]] = 1;
        for (int col = 0; col < 38; col++) {
            int piv = -1;
            for (int r = col; r < 38; r++)
                if (M[r][90 + col]) {
                    piv = r;
                    break;
                }
            if (piv < 0) {
                std::fprintf(stderr, "oracle: parity part of H singular\n");
                std::abort();
            }
            if (piv != col)
                for (int j = 0; j < 128; j++) std::swap(M[piv][j], M[col][j]);
            for (int r = 0; r < 38; r++)
                if (r != col && M[r][90 + col])
                    for (int j = 0; j < 128; j++) M[r][j] ^= M[col][j];
        }
        for (int r = 0; r < 38; r++)
            for (int j = 0; j < 90; j++) gen[r][j] = M[r][j];
    }
};
const LdpcTables L;

// ldpc_kernel.cuh:32-43
uint16_t calc_crc13(const unsigned char* buf, unsigned length) {
    uint16_t rem = 0;
    for (unsigned i = 0; i < length; i++) {
        const int index = (rem >> (13 - 8)) & 0xff;
        rem <<= 8;
        rem |= buf[i];
        rem ^= L.crc_table[index];
    }
    return rem & 0x1fff;
}

void bits_to_bytes(const uint8_t* cw, int nbits, unsigned char* bytes) {
    for (int i = 0; i < nbits / 8; i++) {
        int v = 0;
        for (int b = 0; b < 8; b++) v = (v << 1) | (cw[i * 8 + b] & 1);
        bytes[i] = (unsigned char)v;
    }
}

// ldpc_kernel.cuh:45-63
bool check_crc(unsigned char* b) {
    const uint32_t un = ((uint32_t)(b[9] & 0x7) << 16) | ((uint32_t)b[10] << 8) | (b[11] & 0xc0);
    const uint16_t crc_msg = un >> 6;
    b[9] &= 0xf8;
    b[10] = 0;
    b[11] = 0;
    return crc_msg == calc_crc13(b, 12);
}

// ldpc_kernel.cuh:65-93
float platanh(float x) {
    float isign = 1.0f;
    float z = x;
    if (x < 0.0) {
        isign = -1.0f;
        z = fabsf(x);
    }
    if (z <= 0.664f) return x / 0.83f;
    if (z <= 0.9217f) return isign * (z - 0.4064f) / 0.322f;
    if (z <= 0.9951f) return isign * (z - 0.8378f) / 0.0524f;
    if (z <= 0.9998f) return isign * (z - 0.9914f) / 0.0012f;
    return isign * 7.0f;
}

// ldpc_kernel.cuh:144-247
int ldpc_decode(const float* llr, uint8_t* msg77, int* iters_out, int* nhard_out) {
    float toc[11][38], tov[3][128], zn[128];
    uint8_t cw[128];
    for (int k = 0; k < 3; k++)
        for (int b = 0; b < 128; b++) tov[k][b] = 0.0f;
    for (int c = 0; c < 38; c++) toc[10][c] = 0.0f;
    for (int iter = 0; iter < NITER; iter++) {
        for (int b = 0; b < 128; b++) {
            float sum = 0.0f;
            for (int k = 0; k < 3; k++) sum += tov[k][b];
            zn[b] = llr[b] + sum;
            cw[b] = (zn[b] > 0.0f) ? 1 : 0;
        }
        int ncheck = 0;
        for (int c = 0; c < 38; c++) {
            int s = 0;
            for (int j = 0; j < L.nrw[c]; j++) s += cw[ORC_NM[c][j]];
            ncheck += s % 2;
        }
        bool crc_ok = false;
        if (ncheck == 0) {
            unsigned char bytes[16];
            bits_to_bytes(cw, 128, bytes);
            crc_ok = check_crc(bytes);
        }
        int nhard = 0;
        for (int b = 0; b < 128; b++) {
            const bool good = (cw[b] == 1 && llr[b] > 0.0f) || (cw[b] == 0 && llr[b] <= 0.0f);
            nhard += good ? 0 : 1;
        }
        if (crc_ok && nhard < 18) {
            std::memcpy(msg77, cw, 77);
            *iters_out = iter;
            *nhard_out = nhard;
            return 1;
        }
        for (int b = 0; b < 128; b++)
            for (int k = 0; k < 3; k++) toc[L.slot[b][k]][L.chk[b][k]] = zn[b] - tov[k][b];
        float th[11][38];
        for (int c = 0; c < 38; c++)
            for (int j = 0; j < L.nrw[c]; j++) th[j][c] = tanhf(-0.5f * toc[j][c]);
        for (int b = 0; b < 128; b++)
            for (int k = 0; k < 3; k++) {
                const int c = L.chk[b][k], ex = L.slot[b][k];
                float product = 1.0f;
                for (int j = 0; j < L.nrw[c]; j++)
                    if (j != ex) product *= th[j][c];
                tov[k][b] = 2.0f * platanh(-product);
            }
    }
    *iters_out = NITER;
    *nhard_out = -1;
    return 0;
}

// the same sum-product decoder with the control flow of WSJT-X's
// bpdecode128_90 as used by its MSK144 decoder, where the reference kernel deviates from it:
//   * iter = 0 .. maxiterations (10): 11 parity checks and 10 message updates (the reference: 10 and 9);
//   * the first hard decision that satisfies all parity checks AND the CRC ends the decode; the caller accepts it only
//     if nharderror < 18 (the reference keeps iterating when nhard >= 18); nharderror counts the bits with
//     (2 cw - 1) * llr < 0, so an LLR of exactly zero never counts (the reference counts it against a 1 bit);
//   * early stop: ncnt counts consecutive checks at which the number of unsatisfied parity checks did not decrease;
//     the decode is abandoned when ncnt >= 3 and iter >= 5 and ncheck > 10;
//   * apmask: bits marked a-priori keep zn = llr (no extrinsic information is added to them).
// The arithmetic (tanhf, platanh, product order) is the reference's.  bpdecode128_90.f90 itself is not on this box
// (deps/wsjtx is an empty submodule): this mode is restated from the published algorithm and its parity with WSJT-X is
// UNPINNED; what the tests pin is kernel == this function, bit for bit.
int ldpc_decode_wsjtx(const float* llr, const uint8_t* apmask, uint8_t* msg77, int* iters_out, int* nhard_out) {
    constexpr int MAXIT = 10;
    float toc[11][38], tov[3][128], zn[128];
    uint8_t cw[128];
    for (int k = 0; k < 3; k++)
        for (int b = 0; b < 128; b++) tov[k][b] = 0.0f;
    for (int c = 0; c < 38; c++) toc[10][c] = 0.0f;
    int ncnt = 0, nclast = 0;
    for (int iter = 0; iter <= MAXIT; iter++) {
        for (int b = 0; b < 128; b++) {
            float sum = 0.0f;
            for (int k = 0; k < 3; k++) sum += tov[k][b];
            zn[b] = (apmask && apmask[b]) ? llr[b] : llr[b] + sum;
            cw[b] = (zn[b] > 0.0f) ? 1 : 0;
        }
        int ncheck = 0;
        for (int c = 0; c < 38; c++) {
            int s = 0;
            for (int j = 0; j < L.nrw[c]; j++) s += cw[ORC_NM[c][j]];
            ncheck += s % 2;
        }
        if (ncheck == 0) {
            unsigned char bytes[16];
            bits_to_bytes(cw, 128, bytes);
            if (check_crc(bytes)) {
                int nhard = 0;
                for (int b = 0; b < 128; b++) nhard += ((cw[b] ? 1.0f : -1.0f) * llr[b] < 0.0f) ? 1 : 0;
                std::memcpy(msg77, cw, 77);
                *iters_out = iter;
                *nhard_out = nhard;
                return nhard < 18 ? 1 : 0;
            }
        }
        if (iter > 0) {
            if (ncheck - nclast < 0) ncnt = 0;
            else ncnt++;
            if (ncnt >= 3 && iter >= 5 && ncheck > 10) {
                *iters_out = iter;
                *nhard_out = -1;
                return 0;
            }
        }
        nclast = ncheck;
        if (iter == MAXIT) break;
        for (int b = 0; b < 128; b++)
            for (int k = 0; k < 3; k++) toc[L.slot[b][k]][L.chk[b][k]] = zn[b] - tov[k][b];
        float th[11][38];
        for (int c = 0; c < 38; c++)
            for (int j = 0; j < L.nrw[c]; j++) th[j][c] = tanhf(-0.5f * toc[j][c]);
        for (int b = 0; b < 128; b++)
            for (int k = 0; k < 3; k++) {
                const int c = L.chk[b][k], ex = L.slot[b][k];
                float product = 1.0f;
                for (int j = 0; j < L.nrw[c]; j++)
                    if (j != ex) product *= th[j][c];
                tov[k][b] = 2.0f * platanh(-product);
            }
    }
    *iters_out = MAXIT + 1;
    *nhard_out = -1;
    return 0;
}

// ---------------------------------------------------------------- front end
// analytic2.cuh:93-233.  The per-slice read-all/sync/write-all structure of the reference makes
// both in-place passes equal to out-of-place FIRs; we emulate the slices literally anyway.
const float H15[16] = {0.f,         -0.04225694f, -0.03046893f, 0.f,         0.04570339f, 0.09859952f,
                       0.14789927f, 0.18281356f,  0.19542026f,  0.18281356f, 0.14789927f, 0.09859952f,
                       0.04570339f, 0.f,          -0.03046893f, -0.04225694f};
const int TAPS[13] = {1, 2, 4, 5, 6, 7, 8, 9, 10, 11, 12, 14, 15};
constexpr float SC45 = 0.707106781f;

void lpf_pair(C* cdat /*5248*/) {
    constexpr int TOT = 5248, NSL = TOT / 32;
    for (int slice = 0; slice < NSL - 1; slice++) {
        C tmp[32];
        for (int t = 0; t < 32; t++) {
            const int i = slice * 32 + t;
            C s{0.0f, 0.0f};
            for (int k : TAPS) s = cadd(s, smul(H15[k], cdat[i + (16 - k)]));
            tmp[t] = s;
        }
        for (int t = 0; t < 32; t++) cdat[slice * 32 + t] = tmp[t];
    }
    for (int slice = 0; slice < NSL - 1; slice++) {
        C tmp[32];
        for (int t = 0; t < 32; t++) {
            const int i = TOT - slice * 32 - t - 1;
            C s{0.0f, 0.0f};
            for (int k : TAPS) s = cadd(s, smul(H15[k], cdat[i - (16 - k)]));
            tmp[t] = s;
        }
        for (int t = 0; t < 32; t++) cdat[TOT - slice * 32 - t - 1] = tmp[t];
    }
}

void shift_filter_shift(const C* in, C* out, bool shifts) {
    constexpr int TOT = 5248;
    static const C WL[8] = {{SC45, -SC45}, {0.f, -1.f}, {-SC45, -SC45}, {-1.f, 0.f},
                            {-SC45, SC45}, {0.f, 1.f},  {SC45, SC45},   {1.f, 0.f}};
    static const C WR[8] = {{1.f, 0.f},  {SC45, SC45},   {0.f, 1.f},  {-SC45, SC45},
                            {-1.f, 0.f}, {-SC45, -SC45}, {0.f, -1.f}, {SC45, -SC45}};
    std::vector<C> cdat(TOT, C{0.f, 0.f});
    for (int i = 0; i < N; i++) cdat[32 + i] = in[i];
    if (shifts)
        for (int i = 0; i < TOT; i++) cdat[i] = cmul(cdat[i], WL[i & 7]);
    lpf_pair(cdat.data());
    if (shifts)
        for (int i = 0; i < TOT; i++) cdat[i] = cmul(cdat[i], WR[i & 7]);
    for (int i = 0; i < N; i++) out[i] = cdat[32 + i];
}

// radix-2 FP32 FFT (stand-in for cuFFT C2C, analytic_fft.cu:23,102,134); sign=-1 forward
void fft8192(std::vector<C>& a, int sign) {
    const int n = (int)a.size();
    for (int i = 1, j = 0; i < n; i++) {
        int bit = n >> 1;
        for (; j & bit; bit >>= 1) j ^= bit;
        j ^= bit;
        if (i < j) std::swap(a[i], a[j]);
    }
    for (int len = 2; len <= n; len <<= 1) {
        const double ang = sign * 2.0 * M_PI / len;
        for (int i = 0; i < n; i += len)
            for (int k = 0; k < len / 2; k++) {
                const C w{(float)std::cos(ang * k), (float)std::sin(ang * k)};
                const C u = a[i + k], v = cmul(a[i + k + len / 2], w);
                a[i + k] = cadd(u, v);
                a[i + k + len / 2] = C{u.re - v.re, u.im - v.im};
            }
    }
}

// analytic_fft.cu:18-58,84-157
void analytic_fft(const C* in, C* out) {
    constexpr int NFFT = 8192;
    static std::vector<float> h;
    if (h.empty()) {
        std::vector<float> hh(NFFT, 0.0f);
        const float df = 12000.0f / NFFT;
        const int nh = NFFT / 2;
        const float t = 1.0f / 2000.0f, beta = 0.1f;
        for (int i = 0; i < nh; i++) {
            const float ff = i * df;
            const float f = ff - 1500.0f;
            hh[i] = 1.0f;
            if (fabsf(f) > (1 - beta) / (2 * t) && fabsf(f) <= (1 + beta) / (2 * t))
                hh[i] = hh[i] * 0.5f * (1.0f + cosf((PI_F * t / beta) * (fabsf(f) - (1 - beta) / (2 * t))));
            else if (fabsf(f) > (1 + beta) / (2 * t))
                hh[i] = 0;
        }
        h.swap(hh);
    }
    const float fac = 2.0f / NFFT;
    std::vector<C> a(NFFT, C{0.f, 0.f});
    for (int i = 0; i < N; i++) a[i] = smul(fac, in[i]);
    fft8192(a, -1);
    for (int i = 0; i < NFFT / 2; i++) a[i] = smul(h[i], a[i]);
    a[0] = smul(0.5f, a[0]);
    for (int i = NFFT / 2; i < NFFT; i++) a[i] = C{0.f, 0.f};
    fft8192(a, +1);
    for (int i = 0; i < N; i++) out[i] = a[i];
}

// ---------------------------------------------------------------- host post-processing
extern "C" void __packjt77_MOD_unpack77(char c77[], const int* nrx, char msg[], int* ok, size_t, size_t);

// decode_softbits.cpp:23-55
bool decode_message(const uint8_t* m, std::string* text) {
    const int n3 = (m[71] << 2) | (m[72] << 1) | m[73];
    const int i3 = (m[74] << 2) | (m[75] << 1) | m[76];
    if ((i3 == 0 && (n3 == 1 || n3 == 3 || n3 == 4 || n3 > 5)) || i3 == 3 || i3 > 5) return false;
    char c77[77];
    for (int i = 0; i < 77; i++) c77[i] = '0' + m[i];
    const int nrx = 1;
    char msg[37];
    std::memset(msg, ' ', 37);
    int ok = 0;
    __packjt77_MOD_unpack77(c77, &nrx, msg, &ok, 77, 37);
    if (!ok) return false;
    int len = 37;
    while (len > 0 && msg[len - 1] == ' ') len--;
    text->assign(msg, msg + len);
    return true;
}

struct FilterItem {  // result_filter.h:17-29
    int snr;
    float f0;
    int num_avg, nbadsync, pattern_idx;
    std::string message;
    uint8_t payload[80];
};

// result_filter.cpp:43-74
std::vector<FilterItem> block_end(const std::vector<FilterItem>& scope) {
    std::vector<FilterItem> result;
    std::set<std::string> only;
    for (auto const& e : scope) only.insert(e.message);
    for (auto const& msg : only) {
        std::vector<FilterItem> filtered;
        std::copy_if(scope.begin(), scope.end(), std::back_inserter(filtered),
                     [&msg](FilterItem const& x) { return x.message == msg; });
        std::sort(filtered.begin(), filtered.end(), [](FilterItem const& a, FilterItem const& b) {
            if (a.num_avg == b.num_avg) return a.nbadsync < b.nbadsync;
            return a.num_avg < b.num_avg;
        });
        result.push_back(*filtered.begin());
    }
    return result;
}

struct PresentItem {
    uint32_t index;
    float f0;
    int num_avg, nbadsync, pattern_idx;
    uint8_t message[80];
};

// main.cu:480-525 host loop over items in index order
std::vector<FilterItem> host_block(const std::vector<PresentItem>& items, int compat_cache, int snr) {
    std::vector<FilterItem> scope;
    std::map<std::string, std::pair<bool, std::string>> cache;
    bool have_first = false;
    std::pair<bool, std::string> first;
    for (auto const& it : items) {
        std::pair<bool, std::string> res;
        if (compat_cache) {
            // MessageItem::operator< never iterates (main.cu:437-445): all keys are equivalent, the
            // map holds one entry and every later candidate reuses the FIRST candidate's result.
            if (!have_first) {
                first.first = decode_message(it.message, &first.second);
                have_first = true;
            }
            res = first;
        } else {
            std::string key((const char*)it.message, 77);
            auto f = cache.find(key);
            if (f == cache.end()) {
                std::pair<bool, std::string> r;
                r.first = decode_message(it.message, &r.second);
                f = cache.emplace(key, r).first;
            }
            res = f->second;
        }
        if (res.first) {
            FilterItem fi;
            fi.snr = snr;
            fi.f0 = it.f0;
            fi.num_avg = it.num_avg;
            fi.nbadsync = it.nbadsync;
            fi.pattern_idx = it.pattern_idx;
            fi.message = res.second;
            std::memcpy(fi.payload, it.message, 80);
            scope.push_back(fi);
        }
    }
    return block_end(scope);
}

void decode_block_impl(const C* x, const orc_config* cfg, orc_item* items, bool parallel) {
    const Grid g = make_grid(cfg->center_hz, cfg->width_hz, cfg->step_hz);
    const int depth = clamp_depth(cfg->scan_depth);
    const int thr = cfg->nbadsync_threshold;
    const bool cfg_wsjtx_bp = cfg->wsjtx_bp != 0;
    parallel_for(g.nf, parallel ? cfg->nthreads : 1, [&](long long fi_) {
        const int fi = (int)fi_;
        std::vector<C> cdat2(N);
        const float freq = grid_freq(g, fi);
        mix_block(x, freq, cdat2.data());
        Best best[8][NCAND];
        if (cfg->scan_mode == 0)
            scan_freq_reforder(cdat2.data(), depth, best);
        else
            scan_freq_fast(cdat2.data(), depth, best);
        for (int p = 0; p < depth; p++)
            for (int c = 0; c < NCAND; c++) {
                orc_item& it = items[(fi * depth + p) * NCAND + c];
                std::memset(&it, 0, sizeof(it));  // clear_result, result_keeper.cuh:61-73
                it.freq_idx = fi;
                it.pattern_idx = p;
                it.pos = best[p][c].pos;
                it.xb = best[p][c].xb;
                it.f0 = -(-1 * freq);
                it.num_avg = T.num_avg[p];
                softbits_from_mixed(cdat2.data(), p, it.pos, it.llr, &it.nbadsync, nullptr);
                if (it.nbadsync <= thr) {  // index_kernel.cuh:17-75
                    int iters = 0, nhard = 0;
                    if (cfg_wsjtx_bp ? ldpc_decode_wsjtx(it.llr, nullptr, it.message, &iters, &nhard)
                                     : ldpc_decode(it.llr, it.message, &iters, &nhard)) {
                        it.is_message_present = 1;
                        it.ldpc_iters = iters;
                        it.ldpc_nhard = nhard;
                    }
                }
            }
    });
}

void frontend(const void* raw, const orc_config* cfg, C* out) {
    if (cfg->read_mode == 2)
        orc_frontend_iq((const int8_t*)raw, (float*)out);
    else
        orc_frontend_audio((const int16_t*)raw, cfg->analytic_method, (float*)out);
}

}  // namespace

// ================================================================ C API
extern "C" {

int orc_num_freqs(float width_hz, float step_hz) { return make_grid(0.f, width_hz, step_hz).nf; }
float orc_freq(float center_hz, float width_hz, float step_hz, int idx) {
    return grid_freq(make_grid(center_hz, width_hz, step_hz), idx);
}
int orc_effective_depth(int d) { return clamp_depth(d); }
void orc_sync_waveform(float* cb, float* pp) {
    for (int i = 0; i < 42; i++) {
        cb[2 * i] = T.cb42[i].re;
        cb[2 * i + 1] = T.cb42[i].im;
    }
    for (int i = 0; i < 12; i++) pp[i] = T.pp[i];
}
void orc_pattern(int p, int* num_avg, int* mask6) {
    *num_avg = T.num_avg[p];
    for (int i = 0; i < 6; i++) mask6[i] = T.mask[p][i];
}

void orc_frontend_audio(const int16_t* s16, int method, float* out) {
    // main.cu:300-307: sequential FP32 sum of squares (thrust::reduce on the host, init 0.0f)
    float sum = 0.0f;
    for (int i = 0; i < N; i++) {
        const float b = static_cast<float>(s16[i]);
        sum = sum + b * b;
    }
    const float rms = sqrtf(sum / N);
    const float fac = 1.0f / rms;
    std::vector<C> in(N);
    for (int i = 0; i < N; i++) in[i] = C{fac * s16[i], 0.0f};  // main.cu:313,323
    if (method == 1)
        analytic_fft(in.data(), (C*)out);
    else
        shift_filter_shift(in.data(), (C*)out, true);
}

void orc_frontend_iq(const int8_t* iq, float* out) {
    std::vector<C> in(N);
    for (int i = 0; i < N; i++) {  // main.cu:365-371
        const float a = static_cast<float>(iq[2 * i]);
        const float b = static_cast<float>(iq[2 * i + 1]);
        in[i] = C{a / 128.0f, b / 128.0f};
    }
    shift_filter_shift(in.data(), (C*)out, false);
}

void orc_snr_powers(const float* blk, float* powers8) {
    const C* d = (const C*)blk;
    const int bs = N / 8;
    for (int i = 0; i < 8; i++) powers8[i] = 0.0f;
    for (int i = 0; i < bs * 8; i++) {
        const C y = cmul(cconj(d[i]), d[i]);
        powers8[i / bs] += y.re;
    }
}

float orc_snr_update(float* state, const float* arr, int* snri) {
    float summ = 0.0f;
    for (int i = 0; i < 8; i++) summ += arr[i];
    const float avg = summ / 8;
    float peak = arr[0];
    for (int i = 1; i < 8; i++)
        if (peak < arr[i]) peak = arr[i];
    float np = *state;
    if (np <= 0.0f)
        np = avg;
    else if (avg > np)
        np = 0.9f * np + 0.1f * avg;
    else
        np = avg;
    *state = np;
    float snr;
    if (np > 0.0f)
        snr = 10.0f * std::log10(peak / np - 1.0f);
    else
        snr = 0.0f;
    if (snr > 24.0f) snr = 24.0f;
    if (snr < -8.0f) snr = -8.0f;
    if (snri) *snri = static_cast<int>(snr);
    return snr;
}

void orc_scan(const float* blk, const orc_config* cfg, orc_item* items) {
    const C* x = (const C*)blk;
    const Grid g = make_grid(cfg->center_hz, cfg->width_hz, cfg->step_hz);
    const int depth = clamp_depth(cfg->scan_depth);
    parallel_for(g.nf, cfg->nthreads, [&](long long fi_) {
        const int fi = (int)fi_;
        std::vector<C> cdat2(N);
        const float freq = grid_freq(g, fi);
        mix_block(x, freq, cdat2.data());
        Best best[8][NCAND];
        if (cfg->scan_mode == 0)
            scan_freq_reforder(cdat2.data(), depth, best);
        else
            scan_freq_fast(cdat2.data(), depth, best);
        for (int p = 0; p < depth; p++)
            for (int c = 0; c < NCAND; c++) {
                orc_item& it = items[(fi * depth + p) * NCAND + c];
                std::memset(&it, 0, sizeof(it));
                it.freq_idx = fi;
                it.pattern_idx = p;
                it.pos = best[p][c].pos;
                it.xb = best[p][c].xb;
                it.f0 = -(-1 * freq);
                it.num_avg = T.num_avg[p];
            }
    });
}

void orc_softbits(const float* blk, float freq_hz, int pattern_idx, uint32_t pos, float* llr128, int* nbadsync,
                  float* sb144) {
    std::vector<C> cdat2(N);
    mix_block((const C*)blk, freq_hz, cdat2.data());
    softbits_from_mixed(cdat2.data(), pattern_idx, pos, llr128, nbadsync, sb144);
}

int orc_ldpc_decode(const float* llr, uint8_t* msg77, int* iters, int* nhard) {
    return ldpc_decode(llr, msg77, iters, nhard);
}

void orc_ldpc_decode_batch(const float* llr, size_t ncw, uint8_t* found, uint8_t* msg77, int32_t* iters,
                           int32_t* nhard, int nthreads) {
    const long long chunk = 1024, nchunks = ((long long)ncw + chunk - 1) / chunk;
    parallel_for(nchunks, nthreads, [&](long long ci) {
        const long long hi = std::min<long long>((long long)ncw, (ci + 1) * chunk);
        for (long long i = ci * chunk; i < hi; i++) {
            uint8_t m[80] = {0};
            int it = 0, nh = 0;
            const int f = ldpc_decode(llr + 128 * i, m, &it, &nh);
            found[i] = (uint8_t)f;
            if (msg77) std::memcpy(msg77 + 77 * i, m, 77);
            if (iters) iters[i] = it;
            if (nhard) nhard[i] = nh;
        }
    });
}

void orc_ldpc_decode_wsjtx_batch(const float* llr, size_t ncw, const uint8_t* apmask128, uint8_t* found, uint8_t* msg77,
                                 int32_t* iters, int32_t* nhard, int nthreads) {
    const long long chunk = 1024, nchunks = ((long long)ncw + chunk - 1) / chunk;
    parallel_for(nchunks, nthreads, [&](long long ci) {
        const long long hi = std::min<long long>((long long)ncw, (ci + 1) * chunk);
        for (long long i = ci * chunk; i < hi; i++) {
            uint8_t m[80] = {0};
            int it = 0, nh = 0;
            const int f = ldpc_decode_wsjtx(llr + 128 * i, apmask128, m, &it, &nh);
            found[i] = (uint8_t)f;
            if (msg77) std::memcpy(msg77 + 77 * i, m, 77);
            if (iters) iters[i] = it;
            if (nhard) nhard[i] = nh;
        }
    });
}

uint32_t orc_crc13(const uint8_t* bits77) {
    uint8_t cw[96] = {0};
    std::memcpy(cw, bits77, 77);
    unsigned char bytes[12];
    bits_to_bytes(cw, 96, bytes);
    return calc_crc13(bytes, 12);
}

void orc_ldpc_encode(const uint8_t* msg77, uint8_t* cw) {
    std::memcpy(cw, msg77, 77);
    const uint32_t crc = orc_crc13(msg77);
    for (int i = 0; i < 13; i++) cw[77 + i] = (crc >> (12 - i)) & 1;
    for (int r = 0; r < 38; r++) {
        int s = 0;
        for (int j = 0; j < 90; j++) s ^= (L.gen[r][j] & cw[j]);
        cw[90 + r] = (uint8_t)s;
    }
}

void orc_ldpc_parity_matrix(uint8_t* h) {
    std::memset(h, 0, 38 * 128);
    for (int c = 0; c < 38; c++)
        for (int s = 0; s < 11; s++)
            if (ORC_NM[c][s] >= 0) h[c * 128 + ORC_NM[c][s]] = 1;
}

void orc_decode_block(const float* blk, const orc_config* cfg, orc_item* items) {
    decode_block_impl((const C*)blk, cfg, items, true);
}

int orc_decode_message(const uint8_t* msg77, char* text) {
    std::string s;
    const bool ok = decode_message(msg77, &s);
    std::snprintf(text, 38, "%s", s.c_str());
    return ok ? 1 : 0;
}

int orc_block_lines(const orc_item* items, int nitems, const orc_config* cfg, int snr, int64_t block_index,
                    orc_line* lines, int cap) {
    std::vector<PresentItem> present;
    for (int i = 0; i < nitems; i++)
        if (items[i].is_message_present) {
            PresentItem p;
            p.index = i;
            p.f0 = items[i].f0;
            p.num_avg = items[i].num_avg;
            p.nbadsync = items[i].nbadsync;
            p.pattern_idx = items[i].pattern_idx;
            std::memcpy(p.message, items[i].message, 80);
            present.push_back(p);
        }
    const auto res = host_block(present, cfg->compat_cache, snr);
    int n = 0;
    for (auto const& r : res) {
        if (n >= cap) break;
        orc_line& l = lines[n++];
        std::memset(&l, 0, sizeof(l));
        l.block_index = block_index;
        l.snr = r.snr;
        l.f0 = r.f0;
        l.num_avg = r.num_avg;
        l.nbadsync = r.nbadsync;
        l.pattern_idx = r.pattern_idx;
        std::memcpy(l.payload, r.payload, 80);
        std::snprintf(l.text, sizeof(l.text), "%s", r.message.c_str());
    }
    return n;
}

int64_t orc_decode_stream(const void* bytes, size_t nbytes, const orc_config* cfg, int64_t first_block,
                          int64_t max_blocks, orc_line* lines, int64_t cap, int64_t* nblocks_out) {
    // main.cu:261-386: first read = one whole block, then half-block hops; a short read ends the loop.
    const size_t unit = (cfg->read_mode == 2) ? 2 : 2;  // bytes per sample: s16 mono, or s8 I + s8 Q
    const size_t nsamp = nbytes / unit;
    int64_t nblocks = nsamp >= (size_t)N ? (int64_t)((nsamp - N) / (N / 2)) + 1 : 0;
    if (first_block < 0) first_block = 0;
    int64_t end_block = nblocks;
    if (max_blocks > 0 && first_block + max_blocks < end_block) end_block = first_block + max_blocks;
    const Grid g = make_grid(cfg->center_hz, cfg->width_hz, cfg->step_hz);
    const int depth = clamp_depth(cfg->scan_depth);
    const int nitems = g.nf * depth * NCAND;
    const int64_t nb = end_block > first_block ? end_block - first_block : 0;
    std::vector<std::vector<PresentItem>> present(nb);
    std::vector<float> powers(nb * 8);
    {
        parallel_for(nb, cfg->nthreads, [&](long long k) {
            thread_local std::vector<orc_item> items;
            thread_local std::vector<C> blk;
            items.resize(nitems);
            blk.resize(N);
            const int64_t b = first_block + k;
            const char* raw = (const char*)bytes + (size_t)b * (N / 2) * unit;
            frontend(raw, cfg, blk.data());
            orc_snr_powers((const float*)blk.data(), &powers[k * 8]);
            decode_block_impl(blk.data(), cfg, items.data(), false);
            for (int i = 0; i < nitems; i++)
                if (items[i].is_message_present) {
                    PresentItem p;
                    p.index = i;
                    p.f0 = items[i].f0;
                    p.num_avg = items[i].num_avg;
                    p.nbadsync = items[i].nbadsync;
                    p.pattern_idx = items[i].pattern_idx;
                    std::memcpy(p.message, items[i].message, 80);
                    present[k].push_back(p);
                }
        });
    }
    // The SNR noise floor is a recurrence over ALL blocks from the start of the stream
    // (snr_tracker.cu:39-50); when first_block > 0 the skipped blocks still feed it.
    float state = 0.0f;
    if (first_block > 0) {
        std::vector<C> blk(N);
        float pw[8];
        for (int64_t b = 0; b < first_block && b < nblocks; b++) {
            const char* raw = (const char*)bytes + (size_t)b * (N / 2) * unit;
            frontend(raw, cfg, blk.data());
            orc_snr_powers((const float*)blk.data(), pw);
            orc_snr_update(&state, pw, nullptr);
        }
    }
    int64_t nlines = 0;
    for (int64_t k = 0; k < nb; k++) {
        int snri = 0;
        orc_snr_update(&state, &powers[k * 8], &snri);
        const auto res = host_block(present[k], cfg->compat_cache, snri);
        for (auto const& r : res) {
            if (nlines < cap) {
                orc_line& l = lines[nlines];
                std::memset(&l, 0, sizeof(l));
                l.block_index = first_block + k;
                l.snr = r.snr;
                l.f0 = r.f0;
                l.num_avg = r.num_avg;
                l.nbadsync = r.nbadsync;
                l.pattern_idx = r.pattern_idx;
                std::memcpy(l.payload, r.payload, 80);
                std::snprintf(l.text, sizeof(l.text), "%s", r.message.c_str());
            }
            nlines++;
        }
    }
    if (nblocks_out) *nblocks_out = nb;
    return nlines;
}

int orc_format_line(const orc_line* l, const char* date14, char* out, int cap) {
    // main.cu:406-419.  f0 goes through operator<<(float) with setw(6): %g-style, 6 significant digits.
    char f0buf[32];
    std::snprintf(f0buf, sizeof(f0buf), "%g", (double)l->f0);
    return std::snprintf(out, cap, "***  snr=%2d; f0=%6s; num_avg=%d; nbadsync=%d; pattern_idx=%d; date=%s; msg='%s'; ",
                         l->snr, f0buf, l->num_avg, l->nbadsync, l->pattern_idx, date14 ? date14 : "xxxxxxxxxxxxxx",
                         l->text);
}

}  // extern "C"
