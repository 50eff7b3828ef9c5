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
// (branch filters u, then one DFT row per wanted channel; only the listed channels are evaluated).
//
// One WARP owns a run of consecutive output frames.  Lane l owns branches q = l + 32 r (r < R = ceil(M/32)), so every
// load instruction of a warp reads 32 consecutive input samples; the 8-frame history of its branches stays in
// registers (the frame loop is unrolled by 8, the window index is a compile-time constant), u_f[q] costs 16 FMAs per
// branch, the DFT rows 4 FMAs per (branch, channel) against twiddles in shared memory, and the 2 x channels partial
// sums of the 32 lanes are added by a transposing butterfly (2 channels - 1 + 5 - log2 shuffles instead of 5 per
// value).  Results are packed 8 frames at a time and leave as one 128-bit store per channel.  No barrier, no shared
// memory for samples; each input byte is read from HBM once (plus 7 frames of history per run).
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/msk144_b200.h"

extern "C" int msk144_fail_(int code, const char* what);  // decoder.cu: sets msk144_last_error() of this thread

namespace chz {

constexpr int T = 8;          // taps per branch
constexpr int MAX_M = 256;    // 8 branches per lane
constexpr int WARPS = 4;      // per CTA

struct Params {
    const unsigned char* in;  // first sample of frame 0; `hist` whole frames before it are valid too
    long long nframes;
    int M, fmt, hist;
    const float* g;           // [T][32 R] branch coefficients, 0 for q >= M
    const float2* w;          // [groups * NCH][32 R] e^{-2 pi j k q / M}, 0 for q >= M and padded channels
    float scale, offset;
    signed char* out;
    long long chan_stride, frame_stride;  // in complex samples
    int nchan, fpw;           // frames per warp, a multiple of 8
};

template <int R>
__device__ __forceinline__ void load_row(const Params& p, long long f, int lane, float2 (&row)[R]) {
    const bool inside = f >= -static_cast<long long>(p.hist) && f < p.nframes;
#pragma unroll
    for (int r = 0; r < R; r++) {
        const int q = lane + 32 * r;
        float2 v = make_float2(0.0f, 0.0f);
        if (inside && q < p.M) {
            const long long idx = f * p.M + q;
            if (p.fmt == MSK144_IQ_S16) {
                const unsigned w = __ldg(reinterpret_cast<const unsigned*>(p.in) + idx);
                v = make_float2(static_cast<float>(static_cast<short>(w & 0xffff)), static_cast<float>(static_cast<short>(w >> 16)));
            } else {
                const unsigned short w = __ldg(reinterpret_cast<const unsigned short*>(p.in) + idx);
                if (p.fmt == MSK144_IQ_U8)
                    v = make_float2(static_cast<float>(w & 0xff) - p.offset, static_cast<float>(w >> 8) - p.offset);
                else
                    v = make_float2(static_cast<float>(static_cast<signed char>(w & 0xff)), static_cast<float>(static_cast<signed char>(w >> 8)));
            }
        }
        row[r] = v;
    }
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

template <int R, int NCH>
__global__ void __launch_bounds__(32 * WARPS) channelize_kernel(Params p) {
    extern __shared__ float2 s_w[];  // [NCH][32 R]
    constexpr int ROW = 32 * R, V = 2 * NCH;
    const int lane = threadIdx.x & 31;
    const int group = blockIdx.y;
    for (int i = threadIdx.x; i < NCH * ROW; i += blockDim.x) s_w[i] = p.w[static_cast<size_t>(group) * NCH * ROW + i];
    __syncthreads();
    const long long warp = static_cast<long long>(blockIdx.x) * WARPS + (threadIdx.x >> 5);
    const long long f_begin = warp * p.fpw;
    if (f_begin >= p.nframes) return;
    const long long f_end = min(p.nframes, f_begin + p.fpw);

    float g[T][R];
#pragma unroll
    for (int t = 0; t < T; t++)
#pragma unroll
        for (int r = 0; r < R; r++) g[t][r] = __ldg(p.g + t * ROW + lane + 32 * r);
    // window: at step j of the unrolled loop slot j holds frame f0 + j, so frame f - t sits in slot (j - t) mod 8
    float2 win[T][R];
#pragma unroll
    for (int t = 1; t < T; t++) load_row<R>(p, f_begin - t, lane, win[T - t]);

    const int vi = value_of_lane<V>(lane);           // which (channel, re/im) this lane ends up with
    const int chan = group * NCH + (vi >> 1), part = vi & 1;
    const bool writer = lane < V && chan < p.nchan;
    signed char* const out_c = p.out + 2 * (static_cast<long long>(chan) * p.chan_stride);
    const bool vec_ok = p.frame_stride == 1 && (p.chan_stride & 7) == 0 && (reinterpret_cast<uintptr_t>(p.out) & 15) == 0;

    for (long long f0 = f_begin; f0 < f_end; f0 += T) {
        unsigned pk[2] = {0u, 0u};
#pragma unroll
        for (int j = 0; j < T; j++) {
            load_row<R>(p, f0 + j, lane, win[j]);
            float2 u[R];
#pragma unroll
            for (int r = 0; r < R; r++) {
                float2 a = make_float2(0.0f, 0.0f);
#pragma unroll
                for (int t = 0; t < T; t++) {
                    const float2 x = win[(j - t + T) % T][r];
                    a.x = fmaf(g[t][r], x.x, a.x);
                    a.y = fmaf(g[t][r], x.y, a.y);
                }
                u[r] = a;
            }
            float acc[V];
#pragma unroll
            for (int c = 0; c < NCH; c++) {
                float re = 0.0f, im = 0.0f;
#pragma unroll
                for (int r = 0; r < R; r++) {
                    const float2 w = s_w[c * ROW + lane + 32 * r];
                    re = fmaf(u[r].x, w.x, re);
                    re = fmaf(-u[r].y, w.y, re);
                    im = fmaf(u[r].x, w.y, im);
                    im = fmaf(u[r].y, w.x, im);
                }
                acc[2 * c] = re, acc[2 * c + 1] = im;
            }
            const float total = reduce_transposed<V>(acc, lane);
            const int b = max(-127, min(127, __float2int_rn(total * p.scale)));
            pk[j >> 2] |= static_cast<unsigned>(b & 0xff) << (8 * (j & 3));
        }
        // lanes holding re and im of one channel differ in bit log2(V) - 1
        const unsigned other0 = __shfl_xor_sync(0xffffffffu, pk[0], V / 2);
        const unsigned other1 = __shfl_xor_sync(0xffffffffu, pk[1], V / 2);
        if (!writer) continue;
        if (vec_ok && f0 + T <= f_end) {
            if (part == 0) {
                uint4 o;
                o.x = __byte_perm(pk[0], other0, 0x5140), o.y = __byte_perm(pk[0], other0, 0x7362);
                o.z = __byte_perm(pk[1], other1, 0x5140), o.w = __byte_perm(pk[1], other1, 0x7362);
                *reinterpret_cast<uint4*>(out_c + 2 * f0) = o;
            }
        } else {
#pragma unroll
            for (int j = 0; j < T; j++)
                if (f0 + j < f_end) out_c[2 * (f0 + j) * p.frame_stride + part] = static_cast<signed char>((pk[j >> 2] >> (8 * (j & 3))) & 0xff);
        }
    }
}

using KernelFn = void (*)(Params);
template <int R>
KernelFn pick_nch(int nch) {
    switch (nch) {
        case 1: return channelize_kernel<R, 1>;
        case 2: return channelize_kernel<R, 2>;
        case 4: return channelize_kernel<R, 4>;
        default: return channelize_kernel<R, 8>;
    }
}
KernelFn pick(int R, int nch) {
    switch (R) {
        case 1: return pick_nch<1>(nch);
        case 2: return pick_nch<2>(nch);
        case 3: return pick_nch<3>(nch);
        case 4: return pick_nch<4>(nch);
        case 5: return pick_nch<5>(nch);
        case 6: return pick_nch<6>(nch);
        case 7: return pick_nch<7>(nch);
        default: return pick_nch<8>(nch);
    }
}

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
    int R = 1, nch = 1, groups = 1, sample_bytes = 2, sm_count = 148;
    float* d_g = nullptr;
    float2* d_w = nullptr;
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
    p.g = z->d_g, p.w = z->d_w;
    const float full = z->cfg.input_format == MSK144_IQ_U8 ? 127.5f : z->cfg.input_format == MSK144_IQ_S8 ? 128.0f : 32768.0f;
    p.scale = 127.0f * z->cfg.gain / full;
    p.offset = 127.5f;
    p.out = static_cast<signed char*>(d_out);
    p.chan_stride = static_cast<long long>(chan_stride), p.frame_stride = static_cast<long long>(frame_stride);
    p.nchan = static_cast<int>(z->bins.size());
    // runs long enough that re-reading 7 frames of history costs little, short enough to fill the GPU
    const long long want_warps = static_cast<long long>(z->sm_count) * 16 / z->groups + 1;
    long long fpw = (static_cast<long long>(nframes) + want_warps - 1) / want_warps;
    fpw = std::max<long long>(64, std::min<long long>(512, (fpw + 7) / 8 * 8));
    p.fpw = static_cast<int>(fpw);
    const long long warps = (static_cast<long long>(nframes) + fpw - 1) / fpw;
    dim3 grid(static_cast<unsigned>((warps + chz::WARPS - 1) / chz::WARPS), static_cast<unsigned>(z->groups));
    const size_t smem = static_cast<size_t>(z->nch) * 32 * z->R * sizeof(float2);
    z->kernel<<<grid, 32 * chz::WARPS, smem, st>>>(p);
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
    z->R = (M + 31) / 32;
    z->nch = nbins >= 8 ? 8 : nbins > 2 ? 4 : nbins;
    z->groups = (nbins + z->nch - 1) / z->nch;
    z->sample_bytes = cfg->input_format == MSK144_IQ_S16 ? 4 : 2;
    z->kernel = chz::pick(z->R, z->nch);
    const std::vector<double> h = chz::prototype(M);
    z->taps.assign(h.begin(), h.end());
    const int row = 32 * z->R;
    std::vector<float> g(static_cast<size_t>(chz::T) * row, 0.0f);
    for (int t = 0; t < chz::T; t++)
        for (int q = 0; q < M; q++) g[static_cast<size_t>(t) * row + q] = static_cast<float>(h[t * M + M - 1 - q]);
    std::vector<float2> w(static_cast<size_t>(z->groups) * z->nch * row, make_float2(0.0f, 0.0f));
    for (int c = 0; c < nbins; c++)
        for (int q = 0; q < M; q++) {
            const long long kq = (static_cast<long long>(((bins[c] % M) + M) % M) * q) % M;  // exact phase index
            const double ph = -2.0 * M_PI * static_cast<double>(kq) / M;
            w[static_cast<size_t>(c) * row + q] = make_float2(static_cast<float>(std::cos(ph)), static_cast<float>(std::sin(ph)));
        }
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
    CKB(cudaMalloc(&z->d_w, w.size() * sizeof(float2)));
    CKB(cudaMemcpy(z->d_g, g.data(), g.size() * sizeof(float), cudaMemcpyHostToDevice));
    CKB(cudaMemcpy(z->d_w, w.data(), w.size() * sizeof(float2), cudaMemcpyHostToDevice));
#undef CKB
    *out = z;
    return MSK144_OK;
}

void msk144_channelizer_destroy(msk144_channelizer* z) {
    if (!z) return;
    cudaSetDevice(z->cfg.device);
    if (z->stream) cudaStreamSynchronize(z->stream), cudaStreamDestroy(z->stream);
    cudaFree(z->d_g), cudaFree(z->d_w), cudaFree(z->d_in), cudaFree(z->d_out);
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
