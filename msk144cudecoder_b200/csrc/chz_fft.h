// chz_fft.h -- mixed-radix Stockham stages for the channeliser's M-point DFT across polyphase branches (channelizer.cu).
//
// Plain functions of (thread index, thread count) with no CUDA-only construct, so the same source runs on the device and
// in tests/test_channelizer.py's host harness (tests/chz_fft_host.cpp, which walks the thread index in a loop).
//
// One stage of radix R over `nfr` frames of M points each (frame-major, float2):  butterfly j < M/R of a frame takes
// inputs j + t M/R, multiplies input t by e^{-2 pi i t k / (Ns R)} (k = j mod Ns, Ns = product of the radices already
// done), does an R-point DFT and writes outputs (j - k) R + k + t Ns.  After all stages the bins are in natural order.
// The twiddle e^{-2 pi i a / M} is tw[a]; every angle a stage needs is a multiple of 2 pi / M because Ns R divides M.
#pragma once
#include <vector_types.h>

#include <initializer_list>

#if defined(__CUDACC__)
#define CHZ_HD __host__ __device__ __forceinline__
#else
#define CHZ_HD inline
#endif

namespace chzfft {

CHZ_HD float2 cadd(float2 a, float2 b) { return float2{a.x + b.x, a.y + b.y}; }
CHZ_HD float2 csub(float2 a, float2 b) { return float2{a.x - b.x, a.y - b.y}; }
CHZ_HD float2 cmul(float2 a, float2 b) { return float2{a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x}; }
CHZ_HD float2 mul_mi(float2 a) { return float2{a.y, -a.x}; }  // a x (-i)
CHZ_HD float2 scale(float2 a, float s) { return float2{a.x * s, a.y * s}; }

template <int R>
struct Dft {
    // dense R-point DFT, W_R^a = tw[a x step] (R prime: 7, 11, 13)
    static CHZ_HD void run(float2 (&v)[R], const float2* tw, int step) {
        float2 y[R];
#pragma unroll
        for (int o = 0; o < R; o++) {
            float2 acc = v[0];
#pragma unroll
            for (int t = 1; t < R; t++) acc = cadd(acc, cmul(v[t], tw[((t * o) % R) * step]));
            y[o] = acc;
        }
#pragma unroll
        for (int o = 0; o < R; o++) v[o] = y[o];
    }
};
template <>
struct Dft<2> {
    static CHZ_HD void run(float2 (&v)[2], const float2*, int) {
        const float2 a = v[0], b = v[1];
        v[0] = cadd(a, b), v[1] = csub(a, b);
    }
};
template <>
struct Dft<3> {
    static CHZ_HD void run(float2 (&v)[3], const float2*, int) {
        const float2 t = cadd(v[1], v[2]);
        const float2 m = float2{v[0].x - 0.5f * t.x, v[0].y - 0.5f * t.y};
        const float2 s = mul_mi(scale(csub(v[1], v[2]), 0.86602540378443865f));
        v[0] = cadd(v[0], t), v[1] = cadd(m, s), v[2] = csub(m, s);
    }
};
template <>
struct Dft<4> {
    static CHZ_HD void run(float2 (&v)[4], const float2*, int) {
        const float2 t0 = cadd(v[0], v[2]), t1 = csub(v[0], v[2]), t2 = cadd(v[1], v[3]), t3 = mul_mi(csub(v[1], v[3]));
        v[0] = cadd(t0, t2), v[1] = cadd(t1, t3), v[2] = csub(t0, t2), v[3] = csub(t1, t3);
    }
};
template <>
struct Dft<5> {
    static CHZ_HD void run(float2 (&v)[5], const float2*, int) {
        constexpr float c1 = 0.30901699437494742f, c2 = -0.80901699437494742f, s1 = 0.95105651629515357f, s2 = 0.58778525229247313f;
        const float2 t1 = cadd(v[1], v[4]), t2 = cadd(v[2], v[3]), t3 = csub(v[1], v[4]), t4 = csub(v[2], v[3]);
        const float2 m1 = float2{v[0].x + c1 * t1.x + c2 * t2.x, v[0].y + c1 * t1.y + c2 * t2.y};
        const float2 m2 = float2{v[0].x + c2 * t1.x + c1 * t2.x, v[0].y + c2 * t1.y + c1 * t2.y};
        const float2 n1 = mul_mi(float2{s1 * t3.x + s2 * t4.x, s1 * t3.y + s2 * t4.y});
        const float2 n2 = mul_mi(float2{s2 * t3.x - s1 * t4.x, s2 * t3.y - s1 * t4.y});
        v[0] = cadd(v[0], cadd(t1, t2));
        v[1] = cadd(m1, n1), v[4] = csub(m1, n1), v[2] = cadd(m2, n2), v[3] = csub(m2, n2);
    }
};
template <>
struct Dft<8> {
    static CHZ_HD void run(float2 (&v)[8], const float2*, int) {
        constexpr float h = 0.70710678118654752f;
        float2 e[4] = {v[0], v[2], v[4], v[6]}, o[4] = {v[1], v[3], v[5], v[7]};
        Dft<4>::run(e, nullptr, 0), Dft<4>::run(o, nullptr, 0);
        o[1] = float2{h * (o[1].x + o[1].y), h * (o[1].y - o[1].x)};    // x (1 - i) / sqrt 2
        o[2] = mul_mi(o[2]);
        o[3] = float2{h * (o[3].y - o[3].x), -h * (o[3].x + o[3].y)};   // x (-1 - i) / sqrt 2
#pragma unroll
        for (int k = 0; k < 4; k++) v[k] = cadd(e[k], o[k]), v[k + 4] = csub(e[k], o[k]);
    }
};

// magic = ceil(2^32 / d): n / d == umulhi(n, magic) for n < 2^20, 2 <= d <= 256 (error term n (magic d - 2^32) < 2^32)
CHZ_HD unsigned fast_div(unsigned n, unsigned magic) {
#if defined(__CUDA_ARCH__)
    return __umulhi(n, magic);
#else
    return static_cast<unsigned>((static_cast<unsigned long long>(n) * magic) >> 32);
#endif
}

// shared-memory index of point i of a frame: one slot of padding per 8, so that the stride-R writes of the early
// stages (R = 8: every lane 64 bytes apart) spread over all banks; a frame occupies row_len(M) slots
CHZ_HD int pad(int i) { return i + (i >> 3); }
CHZ_HD int row_len(int M) { return pad(M - 1) + 1; }

struct StagePlan {
    int radix, Ns, per_frame;            // per_frame = M / radix butterflies
    unsigned magic_per_frame, magic_Ns;  // fast_div constants (per_frame or Ns == 1: unused)
    int tw_step;                         // M / (Ns radix)
};

template <int R>
CHZ_HD void stage(const float2* __restrict__ in, float2* __restrict__ out, int M, int nfr, const StagePlan& sp,
                  const float2* __restrict__ tw, int tid, int nthreads) {
    const int total = nfr * sp.per_frame;
    for (int b = tid; b < total; b += nthreads) {
        const int f = sp.per_frame == 1 ? b : static_cast<int>(fast_div(static_cast<unsigned>(b), sp.magic_per_frame));
        const int j = b - f * sp.per_frame;
        const int k = sp.Ns == 1 ? 0 : j - static_cast<int>(fast_div(static_cast<unsigned>(j), sp.magic_Ns)) * sp.Ns;
        const float2* x = in + f * row_len(M);
        float2* y = out + f * row_len(M);
        float2 v[R];
#pragma unroll
        for (int t = 0; t < R; t++) v[t] = x[pad(j + t * sp.per_frame)];
        if (sp.Ns > 1) {
            const int ts = k * sp.tw_step;
#pragma unroll
            for (int t = 1; t < R; t++) v[t] = cmul(v[t], tw[t * ts]);
        }
        Dft<R>::run(v, tw, M / R);
        const int j0 = (j - k) * R + k;
#pragma unroll
        for (int t = 0; t < R; t++) y[pad(j0 + t * sp.Ns)] = v[t];
    }
}

// BIG = false leaves the dense radices (7, 11, 13: 2 R registers of butterfly state) out of the instantiation
template <bool BIG>
CHZ_HD void run_stage(const float2* in, float2* out, int M, int nfr, const StagePlan& sp, const float2* tw, int tid, int nthreads) {
    switch (sp.radix) {
        case 2: stage<2>(in, out, M, nfr, sp, tw, tid, nthreads); break;
        case 3: stage<3>(in, out, M, nfr, sp, tw, tid, nthreads); break;
        case 4: stage<4>(in, out, M, nfr, sp, tw, tid, nthreads); break;
        case 5: stage<5>(in, out, M, nfr, sp, tw, tid, nthreads); break;
        case 8: stage<8>(in, out, M, nfr, sp, tw, tid, nthreads); break;
        default:
            if (BIG) {
                if (sp.radix == 7) stage<7>(in, out, M, nfr, sp, tw, tid, nthreads);
                else if (sp.radix == 11) stage<11>(in, out, M, nfr, sp, tw, tid, nthreads);
                else stage<13>(in, out, M, nfr, sp, tw, tid, nthreads);
            }
            break;
    }
}

constexpr int MAX_STAGES = 8;

// M -> radices (8s first, then 4, 2, 5, 3, 7, 11, 13); returns the number of stages, 0 if M has a larger prime factor
inline int plan(int M, StagePlan (&sp)[MAX_STAGES]) {
    int rad[16], n = 0, rest = M;
    while (rest % 8 == 0) rad[n++] = 8, rest /= 8;
    while (rest % 4 == 0) rad[n++] = 4, rest /= 4;
    while (rest % 2 == 0) rad[n++] = 2, rest /= 2;
    for (int r : {5, 3, 7, 11, 13})
        while (rest % r == 0) {
            if (n >= MAX_STAGES) return 0;
            rad[n++] = r, rest /= r;
        }
    if (rest != 1 || n > MAX_STAGES) return 0;
    int Ns = 1;
    for (int s = 0; s < n; s++) {
        sp[s].radix = rad[s], sp[s].Ns = Ns, sp[s].per_frame = M / rad[s];
        sp[s].magic_per_frame = sp[s].per_frame > 1 ? static_cast<unsigned>((0x100000000ull + sp[s].per_frame - 1) / sp[s].per_frame) : 0u;
        sp[s].magic_Ns = Ns > 1 ? static_cast<unsigned>((0x100000000ull + Ns - 1) / Ns) : 0u;
        sp[s].tw_step = M / (Ns * rad[s]);
        Ns *= rad[s];
    }
    return n;
}

}  // namespace chzfft
