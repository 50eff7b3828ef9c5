// channelizer.cu -- wideband I/Q capture -> N channels of 12 ksps s8 I/Q, the input format of --read-mode=2
// (include/msk144_b200.h, msk144_channelizer_*; SURVEY.md section 8f, ingest side).
//
// The reference has no channeliser: its README documents an external chain for ONE channel,
//   rtl_sdr -s 1920000 | csdr convert_u8_f | fir_decimate_cc 8 | 5 | 4 | gain_ff 100 | convert_f_s8   (README.md:75-78)
// i.e. 1.92 Msps u8 I/Q -> low-pass -> decimate by 160 -> gain -> s8 I/Q at 12 ksps.  This file does that step on the
// GPU for any number of channels of one capture at once (csdr is not on this box: the filter is this file's own
// Kaiser design, stated below, not csdr's).
//
// Uniform polyphase filter bank, decimation M = input rate / 12000, channel k centred k x 12 kHz from the capture's
// centre, prototype low-pass h of L = 8 M taps:
//     y_k[f] = sum_n h[n] x[fM + M-1 - n] e^{-2 pi j k (fM + M-1 - n) / M}
//            = sum_{q<M} e^{-2 pi j k q / M} u_f[q],     u_f[q] = sum_{t<8} h[tM + M-1-q] x[(f-t) M + q]
// (branch filters u, then an M-point DFT across the branches of which only the listed channels are kept).
//
// One THREAD owns one branch q and marches along a run of consecutive frames: the 8-frame history of its branch stays
// in registers (the frame loop is unrolled by 16, window slots are compile-time constants), so u_f[q] costs one
// shared-memory load, two integer->float conversions (PRMT into the mantissa of 2^23, one FADD) and 16 FMAs.  A CTA =
// ceil32(M) threads owns a run of `fpr` frames and streams the raw capture bytes through a 3-stage cp.async ring of
// 16-frame tiles (16-byte copies, zero-filled past the end), so HBM requests stay in flight while the FMAs run; each
// input byte is read from HBM once (plus 7 frames of history per run, read directly).
//
// Across the branches:
//   * direct_kernel (<= 4 channels per group): 4 FMAs per (branch, channel) against the thread's own twiddle, the
//     2 x channels x frames partial sums of a batch (32 values) are added over the 32 lanes by a transposing butterfly
//     (31 shuffles for 32 values instead of 5 each), over the warps through 1 KB of shared memory, then quantised.
//   * fft_kernel (more channels, M = 2^a 3^b 5^c 7^d 11^e 13^f): u goes to shared memory [frame][branch], a mixed-radix
//     Stockham FFT (chz_fft.h: radix 8/4/2/5/3 butterflies, dense 7/11/13) runs over the 16 frames of the tile, and the
//     listed bins leave as 128-bit stores (channel-major) or byte pairs (time-major).  Cost independent of the channel
//     count: all 160 channels of an rtl_sdr capture cost what 1 costs plus the FFT.
// unsigned 8-bit input is taken as (v - 128) + 0.5: the 0.5 goes into the accumulator's start value (0.5 x the branch's
// coefficient sum), which keeps the conversion at one FADD and the data exact.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/msk144_b200.h"
#include "chz_fft.h"

extern "C" int msk144_fail_(int code, const char* what);  // decoder.cu: sets msk144_last_error() of this thread

namespace chz {

constexpr int T = 8;          // taps per branch
constexpr int MAX_M = 256;    // one thread per branch
constexpr int TF = 16;        // frames per staged tile
constexpr int NSTAGE = 3;     // cp.async ring depth

struct Params {
    const unsigned char* in;  // first sample of frame 0; `hist` whole frames before it are valid too
    long long nframes;
    int M, fmt, hist;
    const float* g;           // [T][NT] branch coefficients, 0 for q >= M
    const float* cq;          // [NT] accumulator start (u8: 0.5 x coefficient sum of the branch, else 0)
    const float2* w;          // direct: [groups * NCH][NT] e^{-2 pi j k q / M}, 0 for q >= M and padded channels
    const float2* tw;         // fft: [M] e^{-2 pi j a / M}
    const int* bins;          // fft: [nchan] bin mod M
    float scale;
    signed char* out;
    long long chan_stride, frame_stride;  // in complex samples
    int nchan, fpr;           // frames per run (per CTA), a multiple of TF
    int nstages;
    chzfft::StagePlan sp[chzfft::MAX_STAGES];
};

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc, int src_bytes) {
    const unsigned d = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(gsrc), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

typedef unsigned long long u64;
__device__ __forceinline__ u64 pk2(float a, float b) {
    u64 r;
    asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ float2 up2(u64 v) {
    float2 r;
    asm("mov.b64 {%0,%1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
    return r;
}
// packed FP32 (FFMA2 / FADD2): same FMA-pipe time as two scalar instructions, one issue slot
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
    u64 r;
    asm("fma.rn.f32x2 %0,%1,%2,%3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ u64 add2(u64 a, u64 b) {
    u64 r;
    asm("add.rn.f32x2 %0,%1,%2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}

// integer sample -> float through the mantissa of 2^23: PRMT builds 0x4B00'00bb (or 0x4B00'bbbb), one FADD2 removes the
// bias from both components; signed formats are made unsigned by flipping their sign bits first (the bias carries the
// 128 / 32768).  `missing`: what a sample before the start of the capture counts as (true value 0).
template <int FMT>
struct Conv {
    static constexpr bool WIDE = FMT == MSK144_IQ_S16;
    static constexpr unsigned flip = FMT == MSK144_IQ_S16 ? 0x80008000u : FMT == MSK144_IQ_S8 ? 0x8080u : 0u;
    static constexpr unsigned sel_re = WIDE ? 0x7410u : 0x7540u, sel_im = WIDE ? 0x7432u : 0x7541u;
    static constexpr float bias = WIDE ? 8421376.0f : 8388736.0f;
    static constexpr float missing = FMT == MSK144_IQ_U8 ? -0.5f : 0.0f;
    static __device__ __forceinline__ u64 run(unsigned w) {
        if (flip) w ^= flip;
        return add2(pk2(__uint_as_float(__byte_perm(w, 0x4B000000u, sel_re)), __uint_as_float(__byte_perm(w, 0x4B000000u, sel_im))),
                    pk2(-bias, -bias));
    }
};

// The run of a CTA, its tiles, and the ring that brings their bytes in.
template <int FMT>
struct Feed {
    static constexpr bool WIDE = FMT == MSK144_IQ_S16;
    static constexpr int SB = WIDE ? 4 : 2;  // bytes per complex sample
    const Params& p;
    unsigned char* ring;
    int stage_bytes, ntiles;
    long long f_begin, f_end;
    // 32-bit byte offsets from g_al, the 16-byte aligned address at or below the run's first byte
    const unsigned char* g_al;
    int head, frame_bytes, tile_bytes, run_bytes, avail;
    bool ragged_head;  // the first 16-byte piece of the run starts before the caller's buffer (unaligned pointer, no history)
    __device__ Feed(const Params& p_, unsigned char* ring_) : p(p_), ring(ring_) {
        frame_bytes = p.M * SB;
        tile_bytes = TF * frame_bytes;
        stage_bytes = stage_size(p.M);
        f_begin = static_cast<long long>(blockIdx.x) * p.fpr;
        f_end = min(p.nframes, f_begin + p.fpr);
        ntiles = f_end > f_begin ? static_cast<int>((f_end - f_begin + TF - 1) / TF) : 0;
        const unsigned char* a = p.in + f_begin * frame_bytes;
        head = static_cast<int>(reinterpret_cast<uintptr_t>(a) & 15);
        g_al = a - head;
        run_bytes = head + static_cast<int>(f_end - f_begin) * frame_bytes;
        // bytes that may be read from g_al on: to the end of the capture (a piece may run into the next run's frames)
        avail = head + static_cast<int>(min(p.nframes - f_begin, static_cast<long long>(1 << 20))) * frame_bytes;
        ragged_head = head != 0 && f_begin == 0 && p.hist == 0;
    }
    __host__ __device__ static int stage_size(int M) { return ((TF * M * SB + 15) & ~15) + 16; }
    // every thread: its 16-byte pieces of tile `it` (nothing if the run has no such tile), then one commit
    __device__ __forceinline__ void issue(int it) const {
        if (it < ntiles) {
            const int first = head + it * tile_bytes;            // offset of the tile's first byte
            const int start = first & ~15;
            const int end = min(first + tile_bytes, run_bytes);
            unsigned char* dst = ring + (it % NSTAGE) * stage_bytes - start;
            for (int c = start + threadIdx.x * 16; c < end; c += blockDim.x * 16) {
                if (c == 0 && ragged_head) {
                    for (int b = head; b < 16; b++) dst[b] = g_al[b];
                } else {
                    cp_async16(dst + c, g_al + c, min(16, avail - c));
                }
            }
        }
        cp_async_commit();
    }
    __device__ __forceinline__ const unsigned char* tile_smem(int it) const {
        return ring + (it % NSTAGE) * stage_bytes + ((head + it * tile_bytes) & 15);
    }
    __device__ __forceinline__ static unsigned raw(const unsigned char* base, int idx) {
        if (WIDE) return *reinterpret_cast<const unsigned*>(base + 4 * idx);
        return *reinterpret_cast<const unsigned short*>(base + 2 * idx);
    }
    // frames f_begin - 7 .. f_begin - 1 of branch qc into window slots 1..7 (frame f_begin - t -> slot 8 - t)
    __device__ __forceinline__ void history(int qc, u64 (&win)[T]) const {
        win[0] = pk2(0.0f, 0.0f);
#pragma unroll
        for (int t = 1; t < T; t++) {
            const long long f = f_begin - t;
            u64 v = pk2(Conv<FMT>::missing, Conv<FMT>::missing);
            if (f >= -static_cast<long long>(p.hist)) v = Conv<FMT>::run(raw(p.in + f * static_cast<long long>(frame_bytes), qc));
            win[T - t] = v;
        }
    }
};

// frame j of the tile enters the window (slot j mod 8) and the branch output comes back: 8 FFMA2
template <int FMT>
__device__ __forceinline__ float2 branch_step(const unsigned char* tile, int M, int qc, int j, const u64 (&g)[T], u64 cq, u64 (&win)[T]) {
    win[j % T] = Conv<FMT>::run(Feed<FMT>::raw(tile, j * M + qc));
    u64 a = cq;
#pragma unroll
    for (int t = 0; t < T; t++) a = fma2(g[t], win[(j - t + 2 * T) % T], a);
    return up2(a);
}

// Sum V values (a power of two <= 32) over the 32 lanes.  Each stage trades half of a lane's values with the lane
// whose index differs in one bit, so V - 1 + (5 - log2 V) shuffles do what 5 V would.  Every lane returns the total of
// value number bitrev_{log2 V}(lane mod V); the order of additions is fixed.
template <int V>
__device__ __forceinline__ float reduce_transposed(float (&v)[V], int lane) {
    int bit = 0;
#pragma unroll
    for (int n = V; n > 1; n >>= 1, bit++) {
        const bool upper = (lane >> bit) & 1;
#pragma unroll
        for (int i = 0; i < n / 2; i++) {
            const float send = upper ? v[i] : v[i + n / 2];
            const float keep = upper ? v[i + n / 2] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 1 << bit);
        }
    }
    float s = v[0];
#pragma unroll
    for (; bit < 5; bit++) s += __shfl_xor_sync(0xffffffffu, s, 1 << bit);
    return s;
}

template <int V>
__device__ __forceinline__ int value_of_lane(int lane) {  // bitrev over log2 V bits of lane mod V
    int idx = 0, n = V;
#pragma unroll
    for (int bit = 0; n > 1; n >>= 1, bit++) idx += ((lane >> bit) & 1) * (n >> 1);
    return idx;
}

__device__ __forceinline__ int quantize(float v, float scale) { return max(-127, min(127, __float2int_rn(v * scale))); }

template <int FMT, int NCH>
__global__ void __launch_bounds__(MAX_M, 3) direct_kernel(const __grid_constant__ Params p) {
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int FB = TF / NCH;  // frames per reduction batch: FB x NCH x 2 = 32 values
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int group = blockIdx.y, NT = blockDim.x;
    Feed<FMT> feed(p, smem);
    if (feed.ntiles == 0) return;
    float* red = reinterpret_cast<float*>(smem + NSTAGE * feed.stage_bytes);  // [2][8 warps][32]
    const int qc = min(tid, p.M - 1);  // threads past the last branch read a valid sample against zero coefficients

    u64 g[T];
#pragma unroll
    for (int t = 0; t < T; t++) {
        const float c = __ldg(p.g + t * NT + tid);
        g[t] = pk2(c, c);
    }
    const float cq1 = __ldg(p.cq + tid);
    const u64 cq = pk2(cq1, cq1);
    float2 wc[NCH];
#pragma unroll
    for (int c = 0; c < NCH; c++) wc[c] = __ldg(p.w + static_cast<size_t>(group * NCH + c) * NT + tid);
    u64 win[T];
    feed.history(qc, win);

    // which of a batch's 32 sums this lane holds after the butterfly, and where it goes
    const int vi = value_of_lane<32>(lane);
    const int my_frame = vi / (2 * NCH), my_chan = group * NCH + (vi >> 1) % NCH, my_part = vi & 1;
    signed char* const my_out = p.out + 2 * (static_cast<long long>(my_chan) * p.chan_stride) + my_part;
    auto finish = [&](int nb, long long f) {  // one warp (they take turns), after the barrier that follows batch nb (first frame f)
        float s = 0.0f;
        for (int w = 0; w < nwarps; w++) s += red[((nb & 1) * 8 + w) * 32 + lane];
        f += my_frame;
        if (my_chan < p.nchan && f < feed.f_end) my_out[2 * f * p.frame_stride] = static_cast<signed char>(quantize(s, p.scale));
    };

    for (int s = 0; s < NSTAGE - 1; s++) feed.issue(s);
    int nb = 0;
    long long pending = -1;  // first frame of the batch whose sums sit in `red`, not yet written out
    for (int it = 0; it < feed.ntiles; it++) {
        cp_async_wait<NSTAGE - 2>();
        __syncthreads();
        if (pending >= 0 && warp == (nb - 1) % nwarps) finish(nb - 1, pending);
        feed.issue(it + NSTAGE - 1);
        const unsigned char* tile = feed.tile_smem(it);
        const long long f0 = feed.f_begin + static_cast<long long>(it) * TF;
#pragma unroll
        for (int b = 0; b < NCH; b++) {
            float pv[32];
#pragma unroll
            for (int jj = 0; jj < FB; jj++) {
                const float2 u = branch_step<FMT>(tile, p.M, qc, b * FB + jj, g, cq, win);
#pragma unroll
                for (int c = 0; c < NCH; c++) {
                    pv[(jj * NCH + c) * 2] = fmaf(-u.y, wc[c].y, u.x * wc[c].x);
                    pv[(jj * NCH + c) * 2 + 1] = fmaf(u.y, wc[c].x, u.x * wc[c].y);
                }
            }
            red[((nb & 1) * 8 + warp) * 32 + lane] = reduce_transposed<32>(pv, lane);
            nb++;
            if (b < NCH - 1) {
                __syncthreads();
                if (warp == (nb - 1) % nwarps) finish(nb - 1, f0 + b * FB);
            } else {
                pending = f0 + b * FB;
            }
        }
    }
    __syncthreads();
    if (warp == (nb - 1) % nwarps) finish(nb - 1, pending);
}

template <int FMT, bool BIG>
__global__ void __launch_bounds__(MAX_M, BIG ? 1 : 3) fft_kernel(const __grid_constant__ Params p) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x, NT = blockDim.x, M = p.M;
    Feed<FMT> feed(p, smem);
    if (feed.ntiles == 0) return;
    const int RL = chzfft::row_len(M);  // padded frame (chz_fft.h)
    float2* U0 = reinterpret_cast<float2*>(smem + NSTAGE * feed.stage_bytes);
    float2* U1 = U0 + TF * RL;
    float2* tw = U1 + TF * RL;
    int* bins = reinterpret_cast<int*>(tw + M);
    for (int i = tid; i < M; i += NT) tw[i] = __ldg(p.tw + i);
    for (int i = tid; i < p.nchan; i += NT) bins[i] = chzfft::pad(__ldg(p.bins + i));
    const int qc = min(tid, M - 1);
    u64 g[T];
#pragma unroll
    for (int t = 0; t < T; t++) {
        const float c = __ldg(p.g + t * NT + tid);
        g[t] = pk2(c, c);
    }
    const float cq1 = __ldg(p.cq + tid);
    const u64 cq = pk2(cq1, cq1);
    u64 win[T];
    feed.history(qc, win);
    const int my_slot = chzfft::pad(tid);
    const bool vec_ok = p.frame_stride == 1 && (p.chan_stride & 7) == 0 && (reinterpret_cast<uintptr_t>(p.out) & 15) == 0;

    for (int s = 0; s < NSTAGE - 1; s++) feed.issue(s);
    for (int it = 0; it < feed.ntiles; it++) {
        cp_async_wait<NSTAGE - 2>();
        __syncthreads();  // tile `it` has landed; everybody is past the output pass of tile it - 1
        feed.issue(it + NSTAGE - 1);
        const unsigned char* tile = feed.tile_smem(it);
#pragma unroll
        for (int j = 0; j < TF; j++) {
            const float2 u = branch_step<FMT>(tile, M, qc, j, g, cq, win);
            if (tid < M) U0[j * RL + my_slot] = u;
        }
        __syncthreads();
        float2 *src = U0, *dst = U1;
        for (int s = 0; s < p.nstages; s++) {
            chzfft::run_stage<BIG>(src, dst, M, TF, p.sp[s], tw, tid, NT);
            __syncthreads();
            float2* t = src;
            src = dst, dst = t;
        }
        const long long f0 = feed.f_begin + static_cast<long long>(it) * TF;
        const int nf = static_cast<int>(min(static_cast<long long>(TF), feed.f_end - f0));
        if (vec_ok) {  // channel-major: 8 frames of one channel = one 128-bit store
            for (int u = tid; u < p.nchan * (TF / 8); u += NT) {
                const int c = u >> 1, j0 = (u & 1) * 8;
                const float2* y = src + j0 * RL + bins[c];
                unsigned pk[4];
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const float2 a = y[(2 * k) * RL], b = y[(2 * k + 1) * RL];
                    pk[k] = (quantize(a.x, p.scale) & 0xff) | (quantize(a.y, p.scale) & 0xff) << 8 |
                            (quantize(b.x, p.scale) & 0xff) << 16 | (quantize(b.y, p.scale) & 0xff) << 24;
                }
                signed char* o = p.out + 2 * (static_cast<long long>(c) * p.chan_stride + f0 + j0);
                if (j0 + 8 <= nf) {
                    *reinterpret_cast<uint4*>(o) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                } else {
                    for (int k = 0; j0 + k < nf; k++) reinterpret_cast<unsigned short*>(o)[k] = static_cast<unsigned short>(pk[k >> 1] >> (16 * (k & 1)));
                }
            }
        } else {       // any strides: one byte pair per (frame, channel), channels fastest
            for (int u = tid; u < nf * p.nchan; u += NT) {
                const int j = u / p.nchan, c = u - j * p.nchan;
                const float2 a = src[j * RL + bins[c]];
                signed char* o = p.out + 2 * (static_cast<long long>(c) * p.chan_stride + (f0 + j) * p.frame_stride);
                o[0] = static_cast<signed char>(quantize(a.x, p.scale));
                o[1] = static_cast<signed char>(quantize(a.y, p.scale));
            }
        }
    }
}

using KernelFn = void (*)(Params);
template <int FMT>
KernelFn pick_direct_fmt(int nch) { return nch == 1 ? direct_kernel<FMT, 1> : nch == 2 ? direct_kernel<FMT, 2> : direct_kernel<FMT, 4>; }
KernelFn pick_direct(int fmt, int nch) {
    return fmt == MSK144_IQ_S16 ? pick_direct_fmt<MSK144_IQ_S16>(nch) : fmt == MSK144_IQ_S8 ? pick_direct_fmt<MSK144_IQ_S8>(nch) : pick_direct_fmt<MSK144_IQ_U8>(nch);
}
template <int FMT>
KernelFn pick_fft_fmt(bool big) { return big ? fft_kernel<FMT, true> : fft_kernel<FMT, false>; }
KernelFn pick_fft(int fmt, bool big) {
    return fmt == MSK144_IQ_S16 ? pick_fft_fmt<MSK144_IQ_S16>(big) : fmt == MSK144_IQ_S8 ? pick_fft_fmt<MSK144_IQ_S8>(big) : pick_fft_fmt<MSK144_IQ_U8>(big);
}
int stage_size(int fmt, int M) { return fmt == MSK144_IQ_S16 ? Feed<MSK144_IQ_S16>::stage_size(M) : Feed<MSK144_IQ_U8>::stage_size(M); }

double bessel_i0(double x) {
    double sum = 1.0, term = 1.0;
    for (int k = 1; k < 200; k++) {
        term *= (x / (2.0 * k)) * (x / (2.0 * k));
        sum += term;
        if (term < 1e-18 * sum) break;
    }
    return sum;
}

// Prototype low-pass: L = 8 M taps, Kaiser window beta = 7.857 (80 dB), -6 dB point at 6 kHz (half the channel
// spacing), unit DC gain.  Flat within 0.2 dB to 3.4 kHz; everything that decimation folds onto |f| < 2 kHz (input
// beyond 10 kHz from the channel centre) is >= 79 dB down.  MSK144 occupies ~ +-1.3 kHz around the centre.
std::vector<double> prototype(int M) {
    const int L = T * M;
    const double beta = 7.857, fc = 0.5 / M, mid = 0.5 * (L - 1);
    std::vector<double> h(L);
    double sum = 0.0;
    for (int n = 0; n < L; n++) {
        const double x = n - mid, r = x / mid;
        const double wdw = bessel_i0(beta * std::sqrt(std::max(0.0, 1.0 - r * r))) / bessel_i0(beta);
        const double arg = 2.0 * fc * x;
        const double sinc = arg == 0.0 ? 1.0 : std::sin(M_PI * arg) / (M_PI * arg);
        h[n] = 2.0 * fc * sinc * wdw;
        sum += h[n];
    }
    for (auto& v : h) v /= sum;
    return h;
}

}  // namespace chz

struct msk144_channelizer {
    msk144_channelizer_config cfg;
    std::vector<int32_t> bins;
    std::vector<float> taps;
    int NT = 32, nch = 1, groups = 1, sample_bytes = 2, sm_count = 148;
    bool use_fft = false;
    int nstages = 0;
    chzfft::StagePlan sp[chzfft::MAX_STAGES];
    size_t smem = 0;
    int ctas_per_sm = 1;
    float* d_g = nullptr;
    float* d_cq = nullptr;
    float2* d_w = nullptr;
    float2* d_tw = nullptr;
    int* d_bins = nullptr;
    cudaStream_t stream = nullptr;
    chz::KernelFn kernel = nullptr;
    // one-shot / streaming staging
    unsigned char* d_in = nullptr;
    signed char* d_out = nullptr;
    size_t d_in_cap = 0, d_out_cap = 0;
    std::vector<unsigned char> pending;  // streaming: up to 7 frames of history, then bytes not yet turned into frames
    int hist_frames = 0;
    int64_t launches = 0;
};

namespace {
#define CKZ(call)                                                                                          \
    do {                                                                                                   \
        cudaError_t e_ = (call);                                                                           \
        if (e_ != cudaSuccess) return msk144_fail_(MSK144_ECUDA, (std::string(#call) + ": " + cudaGetErrorString(e_)).c_str()); \
    } while (0)

int reserve(msk144_channelizer* z, size_t in_bytes, size_t out_bytes) {
    CKZ(cudaSetDevice(z->cfg.device));
    if (in_bytes > z->d_in_cap) {
        cudaFree(z->d_in), z->d_in = nullptr, z->d_in_cap = 0;
        CKZ(cudaMalloc(&z->d_in, in_bytes + 16));
        z->d_in_cap = in_bytes;
    }
    if (out_bytes > z->d_out_cap) {
        cudaFree(z->d_out), z->d_out = nullptr, z->d_out_cap = 0;
        CKZ(cudaMalloc(&z->d_out, out_bytes + 16));
        z->d_out_cap = out_bytes;
    }
    return MSK144_OK;
}

int launch(msk144_channelizer* z, const void* d_in, size_t nframes, int hist, void* d_out, size_t chan_stride,
           size_t frame_stride, cudaStream_t st) {
    if (nframes == 0) return MSK144_OK;
    chz::Params p;
    p.in = static_cast<const unsigned char*>(d_in);
    p.nframes = static_cast<long long>(nframes);
    p.M = z->cfg.decimation, p.fmt = z->cfg.input_format, p.hist = hist;
    p.g = z->d_g, p.cq = z->d_cq, p.w = z->d_w, p.tw = z->d_tw, p.bins = z->d_bins;
    const float full = z->cfg.input_format == MSK144_IQ_U8 ? 127.5f : z->cfg.input_format == MSK144_IQ_S8 ? 128.0f : 32768.0f;
    p.scale = 127.0f * z->cfg.gain / full;
    p.out = static_cast<signed char*>(d_out);
    p.chan_stride = static_cast<long long>(chan_stride), p.frame_stride = static_cast<long long>(frame_stride);
    p.nchan = static_cast<int>(z->bins.size());
    p.nstages = z->nstages;
    for (int i = 0; i < z->nstages; i++) p.sp[i] = z->sp[i];
    // equal runs, a whole number of waves (re-reading 7 frames of history and filling the ring cost little per run)
    const int groups = z->use_fft ? 1 : z->groups;
    const long long want_ctas = std::max<long long>(1, 3LL * z->sm_count * z->ctas_per_sm / groups);  // 3 waves: measured best of 1..8
    long long fpr = (static_cast<long long>(nframes) + want_ctas - 1) / want_ctas;
    fpr = std::max<long long>(4 * chz::TF, (fpr + chz::TF - 1) / chz::TF * chz::TF);
    p.fpr = static_cast<int>(std::min<long long>(fpr, 1 << 20));
    const long long runs = (static_cast<long long>(nframes) + p.fpr - 1) / p.fpr;
    dim3 grid(static_cast<unsigned>(runs), static_cast<unsigned>(groups));
    z->kernel<<<grid, z->NT, z->smem, st>>>(p);
    CKZ(cudaGetLastError());
    z->launches += 1;
    return MSK144_OK;
}
}  // namespace

extern "C" {

int msk144_channelizer_create(const msk144_channelizer_config* cfg, const int32_t* bins, int32_t nbins, msk144_channelizer** out) {
    if (!cfg || !out || !bins || nbins < 1) return msk144_fail_(MSK144_EINVAL, "bad argument");
    *out = nullptr;
    const int M = cfg->decimation;
    if (M < 2 || M > chz::MAX_M) return msk144_fail_(MSK144_EINVAL, "decimation must be 2..256 (input rate = decimation x 12000)");
    if (cfg->input_format < MSK144_IQ_U8 || cfg->input_format > MSK144_IQ_S16) return msk144_fail_(MSK144_EINVAL, "unknown input_format");
    if (!(cfg->gain > 0.0f)) return msk144_fail_(MSK144_EINVAL, "gain must be positive");
    for (int i = 0; i < nbins; i++)
        if (bins[i] <= -M || bins[i] >= M) return msk144_fail_(MSK144_EINVAL, "channel bin out of range (|bin| < decimation)");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || cfg->device < 0 || cfg->device >= ndev)
        return msk144_fail_(MSK144_ENODEV, "no usable GPU (there is no CPU fallback)");
    cudaDeviceProp prop;
    CKZ(cudaGetDeviceProperties(&prop, cfg->device));
    if (prop.major < 10) return msk144_fail_(MSK144_ENODEV, "an sm_100-class GPU is required");
    CKZ(cudaSetDevice(cfg->device));
    auto* z = new msk144_channelizer();
    z->cfg = *cfg;
    z->bins.assign(bins, bins + nbins);
    z->sm_count = prop.multiProcessorCount;
    z->NT = (M + 31) / 32 * 32;
    z->sample_bytes = cfg->input_format == MSK144_IQ_S16 ? 4 : 2;
    z->nstages = chzfft::plan(M, z->sp);
    // more than 4 channels: the FFT across the branches costs less than 4 FMAs per (branch, channel) -- if M factors
    z->use_fft = nbins > 4 && z->nstages > 0;
    const size_t ring = static_cast<size_t>(chz::NSTAGE) * chz::stage_size(cfg->input_format, M);
    if (z->use_fft) {
        bool big = false;
        for (int i = 0; i < z->nstages; i++) big = big || z->sp[i].radix > 8 || z->sp[i].radix == 7;
        z->nch = 0, z->groups = 1;
        z->kernel = chz::pick_fft(cfg->input_format, big);
        z->smem = ring + (2 * static_cast<size_t>(chz::TF) * chzfft::row_len(M) + M) * sizeof(float2) + nbins * sizeof(int);
    } else {
        z->nch = nbins > 2 ? 4 : nbins;
        z->groups = (nbins + z->nch - 1) / z->nch;
        z->kernel = chz::pick_direct(cfg->input_format, z->nch);
        z->smem = ring + 2 * 8 * 32 * sizeof(float);
    }
    const std::vector<double> h = chz::prototype(M);
    z->taps.assign(h.begin(), h.end());
    const int row = z->NT;
    std::vector<float> g(static_cast<size_t>(chz::T) * row, 0.0f), cq(row, 0.0f);
    for (int t = 0; t < chz::T; t++)
        for (int q = 0; q < M; q++) g[static_cast<size_t>(t) * row + q] = static_cast<float>(h[t * M + M - 1 - q]);
    if (cfg->input_format == MSK144_IQ_U8)  // (v - 128) + 0.5: the half goes in with the accumulator's start value
        for (int q = 0; q < M; q++) {
            double sum = 0.0;
            for (int t = 0; t < chz::T; t++) sum += static_cast<double>(g[static_cast<size_t>(t) * row + q]);
            cq[q] = static_cast<float>(0.5 * sum);
        }
    auto root = [&](long long a) {  // e^{-2 pi j a / M}, a reduced exactly
        const double ph = -2.0 * M_PI * static_cast<double>(((a % M) + M) % M) / M;
        return make_float2(static_cast<float>(std::cos(ph)), static_cast<float>(std::sin(ph)));
    };
    std::vector<float2> w(static_cast<size_t>(std::max(1, z->groups * z->nch)) * row, make_float2(0.0f, 0.0f)), tw(M);
    std::vector<int> binmod(nbins);
    for (int c = 0; c < nbins; c++) {
        binmod[c] = ((bins[c] % M) + M) % M;
        if (!z->use_fft)
            for (int q = 0; q < M; q++) w[static_cast<size_t>(c) * row + q] = root(static_cast<long long>(binmod[c]) * q);
    }
    for (int a = 0; a < M; a++) tw[a] = root(a);
    auto bail = [&](int rc) {
        msk144_channelizer_destroy(z);
        return rc;
    };
#define CKB(call)                                                                                                           \
    do {                                                                                                                    \
        cudaError_t e_ = (call);                                                                                            \
        if (e_ != cudaSuccess) return bail(msk144_fail_(MSK144_ECUDA, (std::string(#call) + ": " + cudaGetErrorString(e_)).c_str())); \
    } while (0)
    CKB(cudaStreamCreateWithFlags(&z->stream, cudaStreamNonBlocking));
    CKB(cudaMalloc(&z->d_g, g.size() * sizeof(float)));
    CKB(cudaMalloc(&z->d_cq, cq.size() * sizeof(float)));
    CKB(cudaMalloc(&z->d_w, w.size() * sizeof(float2)));
    CKB(cudaMalloc(&z->d_tw, tw.size() * sizeof(float2)));
    CKB(cudaMalloc(&z->d_bins, binmod.size() * sizeof(int)));
    CKB(cudaMemcpy(z->d_g, g.data(), g.size() * sizeof(float), cudaMemcpyHostToDevice));
    CKB(cudaMemcpy(z->d_cq, cq.data(), cq.size() * sizeof(float), cudaMemcpyHostToDevice));
    CKB(cudaMemcpy(z->d_w, w.data(), w.size() * sizeof(float2), cudaMemcpyHostToDevice));
    CKB(cudaMemcpy(z->d_tw, tw.data(), tw.size() * sizeof(float2), cudaMemcpyHostToDevice));
    CKB(cudaMemcpy(z->d_bins, binmod.data(), binmod.size() * sizeof(int), cudaMemcpyHostToDevice));
    CKB(cudaFuncSetAttribute(reinterpret_cast<const void*>(z->kernel), cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(z->smem)));
    CKB(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&z->ctas_per_sm, reinterpret_cast<const void*>(z->kernel), z->NT, z->smem));
    z->ctas_per_sm = std::max(1, z->ctas_per_sm);
#undef CKB
    *out = z;
    return MSK144_OK;
}

void msk144_channelizer_destroy(msk144_channelizer* z) {
    if (!z) return;
    cudaSetDevice(z->cfg.device);
    if (z->stream) cudaStreamSynchronize(z->stream), cudaStreamDestroy(z->stream);
    cudaFree(z->d_g), cudaFree(z->d_cq), cudaFree(z->d_w), cudaFree(z->d_tw), cudaFree(z->d_bins), cudaFree(z->d_in), cudaFree(z->d_out);
    delete z;
}

int msk144_channelizer_prototype(const msk144_channelizer* z, float* taps, size_t cap, size_t* ntaps) {
    if (!z || !ntaps) return msk144_fail_(MSK144_EINVAL, "null argument");
    *ntaps = z->taps.size();
    if (taps) std::memcpy(taps, z->taps.data(), std::min(cap, z->taps.size()) * sizeof(float));
    return (taps && cap < z->taps.size()) ? msk144_fail_(MSK144_EOVERFLOW, "output array too small") : MSK144_OK;
}

int64_t msk144_channelizer_launches(const msk144_channelizer* z) { return z ? z->launches : 0; }

int msk144_channelize_device(msk144_channelizer* z, const void* d_in, size_t nframes, int32_t history_frames, void* d_out,
                             size_t out_chan_stride, size_t out_frame_stride, void* stream) {
    if (!z || !d_in || !d_out) return msk144_fail_(MSK144_EINVAL, "null argument");
    if (history_frames < 0 || history_frames > chz::T - 1) return msk144_fail_(MSK144_EINVAL, "history_frames must be 0..7");
    if (reinterpret_cast<uintptr_t>(d_in) % z->sample_bytes) return msk144_fail_(MSK144_EINVAL, "input not aligned to its sample size");
    CKZ(cudaSetDevice(z->cfg.device));
    return launch(z, d_in, nframes, history_frames, d_out, out_chan_stride, out_frame_stride,
                  stream ? static_cast<cudaStream_t>(stream) : z->stream);
}

int msk144_channelize_host(msk144_channelizer* z, const void* in, size_t nbytes, int8_t* out, size_t out_cap_frames, size_t* nframes) {
    if (!z || (!in && nbytes) || !nframes) return msk144_fail_(MSK144_EINVAL, "null argument");
    const size_t frame_bytes = static_cast<size_t>(z->cfg.decimation) * z->sample_bytes;
    const size_t nf = nbytes / frame_bytes, nchan = z->bins.size();
    *nframes = nf;
    if (nf == 0) return MSK144_OK;
    if (!out || out_cap_frames < nf) return msk144_fail_(MSK144_EOVERFLOW, "output array too small");
    const size_t stride = (nf + 7) / 8 * 8;  // keeps every channel row 16-byte aligned on the device
    if (int rc = reserve(z, nf * frame_bytes, nchan * stride * 2)) return rc;
    CKZ(cudaMemcpyAsync(z->d_in, in, nf * frame_bytes, cudaMemcpyHostToDevice, z->stream));
    if (int rc = launch(z, z->d_in, nf, 0, z->d_out, stride, 1, z->stream)) return rc;
    CKZ(cudaMemcpy2DAsync(out, nf * 2, z->d_out, stride * 2, nf * 2, nchan, cudaMemcpyDeviceToHost, z->stream));
    CKZ(cudaStreamSynchronize(z->stream));
    return MSK144_OK;
}

int msk144_channelizer_push(msk144_channelizer* z, const void* bytes, size_t nbytes, int8_t* out, size_t out_cap_frames, size_t* nframes) {
    if (!z || (!bytes && nbytes) || !nframes) return msk144_fail_(MSK144_EINVAL, "null argument");
    *nframes = 0;
    const size_t frame_bytes = static_cast<size_t>(z->cfg.decimation) * z->sample_bytes;
    const size_t hist_bytes = static_cast<size_t>(z->hist_frames) * frame_bytes, nchan = z->bins.size();
    const size_t nf = (z->pending.size() - hist_bytes + nbytes) / frame_bytes;
    if (nf > 0 && (!out || out_cap_frames < nf)) {
        *nframes = nf;
        return msk144_fail_(MSK144_EOVERFLOW, "output array too small (nothing consumed)");
    }
    const unsigned char* b = static_cast<const unsigned char*>(bytes);
    z->pending.insert(z->pending.end(), b, b + nbytes);
    if (nf == 0) return MSK144_OK;
    const size_t used = hist_bytes + nf * frame_bytes;
    if (int rc = reserve(z, used, nf * nchan * 2)) return rc;
    CKZ(cudaMemcpyAsync(z->d_in, z->pending.data(), used, cudaMemcpyHostToDevice, z->stream));
    // one sample of every channel per time step: what msk144_push_interleaved (or msk144_push for one channel) takes
    if (int rc = launch(z, z->d_in + hist_bytes, nf, z->hist_frames, z->d_out, 1, nchan, z->stream)) return rc;
    CKZ(cudaMemcpyAsync(out, z->d_out, nf * nchan * 2, cudaMemcpyDeviceToHost, z->stream));
    CKZ(cudaStreamSynchronize(z->stream));
    const int keep = static_cast<int>(std::min<size_t>(chz::T - 1, z->hist_frames + nf));
    z->pending.erase(z->pending.begin(), z->pending.begin() + (used - static_cast<size_t>(keep) * frame_bytes));
    z->hist_frames = keep;
    *nframes = nf;
    return MSK144_OK;
}

int msk144_channelizer_reset(msk144_channelizer* z) {
    if (!z) return msk144_fail_(MSK144_EINVAL, "null argument");
    z->pending.clear(), z->hist_frames = 0;
    return MSK144_OK;
}

}  // extern "C"
