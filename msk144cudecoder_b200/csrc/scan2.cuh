// scan2.cuh -- scan + soft bits + sync gate, the form with TWO resident CTAs per SM (included by kernels.cuh).
//
// Same stages and the same arithmetic, operation for operation, as the first form of this kernel (S1 mix, S2a half-pulse
// sums, S2b pulse filter + 42-tap sync correlation, S3 frame sums and chunk maxima, S4 top 8 per pattern, S5 soft bits /
// sync gate / LLR queue: scan_kernel.cuh:27-393, softbits_kernel.cuh:56-241, index_kernel.cuh:17-44).  What changed is
// the shape.  S4 and S5 are chains of shared-memory and shuffle round trips on a few warps (profiles/r02_scan_notes.md:
// 10 k of the 19 k cycles per (block, frequency), issue slots 47 % used), so the cure is a second CTA on the SM whose
// throughput stages run meanwhile, and that needs the working set of one (block, frequency) in <= 113 KB instead of 197 KB:
//   * the half-pulse sums HH (85 KB for a block) exist only for one tile of 1296 positions at a time (22 KB), in the
//     buffer the chunk maxima of S3 use later; S2a / S2b alternate over four tiles, r' overwrites the mixed samples of
//     the tile it was computed from (nothing later reads them: tile k + 1 starts 1296 samples further on, and the last
//     tile's wrap-around reads a 48-entry copy of the block's start kept behind the block);
//   * soft bit 0 of the kept candidates, the one bit that needs Hlo and Hhi separately (softbits_kernel.cuh:157-181),
//     is recomputed for those candidates from the block and the mixer row in L2: 12 samples per frame;
//   * 448 threads (14 warps) per CTA, 72 registers, 110.5 KB of shared memory: two CTAs per SM.
#pragma once

constexpr int SCAN2_THREADS = 448;
constexpr int SCAN2_WARPS = SCAN2_THREADS / 32;
constexpr int SCAN2_COLS = NFR / 32;         // 27 columns of 32 frame offsets (S3)
constexpr int A2_PAD = 48;                   // c[0 .. 47] again at A[N ..]: the last tile reads up to c[N + 41]
constexpr int S2_TILES = 4;
constexpr int S2_TILE = N / S2_TILES;        // 1296 positions
constexpr int S2_TK = S2_TILE / 6;           // 216 cells per residue and tile
constexpr int S2_HCELLS = (S2_TILE + 36) / 6;  // 222: outputs of a tile read half-pulse sums up to 36 entries past it
constexpr int S2_HK = 225;                   // residue stride of the tile buffer, = 1 mod 8 like KH
constexpr size_t SCAN2_REGION = (6 * S2_HK > 8 * NCHUNK ? 6 * S2_HK : 8 * NCHUNK) * sizeof(float4);
constexpr size_t SCAN2_SMEM = (N + A2_PAD) * sizeof(float2) + 6 * KM * sizeof(float2) + SCAN2_REGION +
                              8 * NCAND * sizeof(float4) + 8 * NCAND * sizeof(float);
static_assert(2 * (SCAN2_SMEM + 1024) <= 228 * 1024, "two CTAs per SM");

// S1 with the longer circular copy
__device__ __forceinline__ void scan2_mix(float2* A, const float2* __restrict__ x, const float2* __restrict__ w, int t0, int nthr) {
    const float4* x4 = reinterpret_cast<const float4*>(x);
    const float4* w4 = reinterpret_cast<const float4*>(w);
    float4* A4 = reinterpret_cast<float4*>(A);
#pragma unroll 3
    for (int i = t0; i < N / 2; i += nthr) {
        const float4 xv = __ldg(x4 + i), wv = __ldg(w4 + i);
        const float2 a = cmul_ref(make_float2(wv.x, wv.y), make_float2(xv.x, xv.y));
        const float2 b = cmul_ref(make_float2(wv.z, wv.w), make_float2(xv.z, xv.w));
        const float4 o = make_float4(a.x, a.y, b.x, b.y);
        A4[i] = o;
        if (i < A2_PAD / 2) A4[N / 2 + i] = o;
    }
}

// S2a for one tile: (Hlo, Hhi) of the local entries e = 0 .. 1331 (global u = 1296 tile + e), three per thread.
// Thread t = 222 h + k takes e = 6 k + 3 h + i, i < 3, from c[u .. u + 7]: consecutive lanes write consecutive cells
// of one residue row and read 16-byte cells three apart (no bank conflicts; h = 1 starts on an odd sample and reads
// its first and last sample as 8-byte cells).
__device__ __forceinline__ void scan2_pulse_tile(const float2* A, float4* HT, const float (&pp)[12], int tile, int t) {
    if (t >= 2 * S2_HCELLS) return;
    const int h = (t >= S2_HCELLS) ? 1 : 0, k = t - S2_HCELLS * h;
    const float2* c0 = A + S2_TILE * tile + 6 * k + 3 * h;
    float2 c[8];
    if (h == 0) {
        const float4* c4 = reinterpret_cast<const float4*>(c0);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const float4 v = c4[i];
            c[2 * i] = make_float2(v.x, v.y);
            c[2 * i + 1] = make_float2(v.z, v.w);
        }
    } else {
        const float4* c4 = reinterpret_cast<const float4*>(c0 + 1);
        c[0] = c0[0];
#pragma unroll
        for (int i = 0; i < 3; i++) {
            const float4 v = c4[i];
            c[2 * i + 1] = make_float2(v.x, v.y);
            c[2 * i + 2] = make_float2(v.z, v.w);
        }
        c[7] = c0[7];
    }
#pragma unroll
    for (int i = 0; i < 3; i++) {
        float2 lo = make_float2(0.0f, 0.0f), hi = make_float2(0.0f, 0.0f);
#pragma unroll
        for (int j = 0; j < 6; j++) {
            lo = fma2(c[i + j], bc2(pp[j]), lo);
            hi = fma2(c[i + j], bc2(pp[6 + j]), hi);
        }
        HT[(3 * h + i) * S2_HK + k] = make_float4(lo.x, lo.y, hi.x, hi.y);
    }
}

// S2b for one tile: Mf and r' of the outputs u = 1296 tile + 6 (3 bj + j) + br, j < 3, from the tile's cells
// 3 bj .. 3 bj + 8 of residue br.  Threads of a warp share br and walk bj: 16-byte cells three apart.
__device__ __forceinline__ void scan2_corr_tile(float2* A, const float4* HT, float2* MF, int tile, int tid) {
    if (tid >= 6 * (S2_TK / 3)) return;
    const int br = tid / (S2_TK / 3), bj = tid - (S2_TK / 3) * br;
    float4 h[9];
#pragma unroll
    for (int m = 0; m < 9; m++) h[m] = HT[br * S2_HK + 3 * bj + m];
    float2 mf[8];
#pragma unroll
    for (int m = 0; m < 8; m++) mf[m] = add2(make_float2(h[m].x, h[m].y), make_float2(h[m + 1].z, h[m + 1].w));
#pragma unroll
    for (int j = 0; j < 3; j++) {
        // sync word signs s8 = -+++--+- (msk_context.cuh:147-154), I rail odd bits, Q rail even bits
        const float2 I = sub2(sub2(add2(mf[j], mf[j + 2]), mf[j + 4]), make_float2(h[j + 6].x, h[j + 6].y));
        const float2 Q = add2(sub2(sub2(mf[j + 1], make_float2(h[j].z, h[j].w)), mf[j + 3]), mf[j + 5]);
        const int k = S2_TK * tile + 3 * bj + j;
        A[6 * k + br] = make_float2(__fadd_rn(I.x, Q.y), __fsub_rn(I.y, Q.x));  // I - jQ, natural order
        MF[br * KM + k] = mf[j];
        if (k < MF_PAD / 6) MF[br * KM + N / 6 + k] = mf[j];  // circular copy of entries u < 864 at N + u
    }
}

// (Hlo[il] + Hhi[ih]) of the mixed block, straight from the block and the mixer row (both L2 hits): the terms of soft
// bit 0 for one frame, with the operations of scan2_mix and scan2_pulse_tile
__device__ __forceinline__ float2 scan2_bit0_term(const float2* __restrict__ x, const float2* __restrict__ w, int il, int ih,
                                                  const float (&pp)[12]) {
    float2 cl[6], ch[6];
#pragma unroll
    for (int j = 0; j < 6; j++) {
        int a = il + j, b = ih + j;
        if (a >= N) a -= N;
        if (b >= N) b -= N;
        cl[j] = cmul_ref(__ldg(w + a), __ldg(x + a));
        ch[j] = cmul_ref(__ldg(w + b), __ldg(x + b));
    }
    float2 lo = make_float2(0.0f, 0.0f), hi = make_float2(0.0f, 0.0f);
#pragma unroll
    for (int j = 0; j < 6; j++) {
        lo = fma2(cl[j], bc2(pp[j]), lo);
        hi = fma2(ch[j], bc2(pp[6 + j]), hi);
    }
    return add2(lo, hi);
}

template <int DEPTH>
__global__ void __launch_bounds__(SCAN2_THREADS, 2) scan2_kernel(ScanParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* A = reinterpret_cast<float2*>(smem_raw);                 // mixed block c[], tile by tile r'[]
    float2* MF = A + N + A2_PAD;                                     // 12-tap pulse filter output
    float4* WM = reinterpret_cast<float4*>(MF + 6 * KM);             // [8][162] chunk maxima (|S|^2, pos, S) ...
    float4* HT = WM;                                                 // ... and, before S3, the tile's half-pulse sums
    float4* CT = reinterpret_cast<float4*>(reinterpret_cast<unsigned char*>(WM) + SCAN2_REGION);  // [8][8] kept candidates
    float* SB0 = reinterpret_cast<float*>(CT + 8 * NCAND);           // [8][8] their soft bit 0

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int blk = blockIdx.y;
    const float2* xblk = p.blocks + static_cast<size_t>(blk) * N;
    const int f_begin = blockIdx.x * p.fpc;
    const int f_end = min(p.nf, f_begin + p.fpc);
    float pp[12];
#pragma unroll
    for (int j = 0; j < 12; j++) pp[j] = c_tab.pp[j];

#ifdef MSK_SCAN_PROF
    long long prof_[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, last_ = clock64();
#endif
    scan2_mix(A, xblk, p.mix + static_cast<size_t>(f_begin) * N, tid, SCAN2_THREADS);
    __syncthreads();
    scan2_pulse_tile(A, HT, pp, 0, tid);
    __syncthreads();
    SCAN_TICK(0)

    for (int fi = f_begin; fi < f_end; ++fi) {
        const float freq = fmaf(static_cast<float>(fi), p.step, p.center + p.if1);  // msk_context.cuh:135
        const float2* wrow = p.mix + static_cast<size_t>(fi) * N;
        // ---- S2b / S2a over the tiles ------------------------------------------------------------------------
#pragma unroll 1
        for (int tile = 0; tile < S2_TILES; tile++) {
            scan2_corr_tile(A, HT, MF, tile, tid);
            __syncthreads();
            if (tile + 1 < S2_TILES) {
                scan2_pulse_tile(A, HT, pp, tile + 1, tid);
                __syncthreads();
            }
        }
        SCAN_TICK(1)
        // ---- S3: frame sums for every pattern and start frame; chunk maxima ----------------------------------
        {
            unsigned nm = 0u, want = 0u;
#pragma unroll 1
            for (int col = warp; col < SCAN2_COLS; col += SCAN2_WARPS) {
                float2 q[6];
                scan_load_q(A, 32 * col + lane, q);
                nm += scan_frame_sums<DEPTH, false>(q, WM, 32 * col + lane, col, lt_mask);
                want += 6u * DEPTH;
            }
            // one maximum per chunk and pattern expected; anything else means a tie
            if (__reduce_add_sync(0xffffffffu, nm) != want) {
#pragma unroll 1
                for (int col = warp; col < SCAN2_COLS; col += SCAN2_WARPS)
                    scan_frame_sums_exact<DEPTH>(A, WM, 32 * col + lane, col, lt_mask);
            }
        }
        __syncthreads();
        SCAN_TICK(2)
        // ---- S4 (warps < DEPTH) || S1 of the next frequency (the other warps) --------------------------------
        if (warp < DEPTH) {
            const int pt = warp;
            float4 best;
            float xb;
            int sidx;
            scan_slice_max(WM + pt * NCHUNK, lane, best);
            SCAN_TICK(7)
            xb = hypotf(best.z, best.w);  // abs(S), scan_kernel.cuh:124
            const unsigned xkey = (lane < NSLICE && xb == xb) ? __float_as_uint(xb) : 0u;
            SCAN_TICK(8)
            scan_top8(xkey, lane, lt_mask, sidx);
            SCAN_TICK(9)
            const int src = (sidx < 0) ? 0 : sidx;
            const float cpos = __shfl_sync(0xffffffffu, best.y, src);
            const float cxb = __shfl_sync(0xffffffffu, xb, src);
            const float cre = __shfl_sync(0xffffffffu, best.z, src);
            const float cim = __shfl_sync(0xffffffffu, best.w, src);
            // soft bit 0 of the kept candidates: its 12-tap window wraps inside the 864-sample frame
            // (softbits_kernel.cuh:157-181), so it needs Hlo[ih + 858] and Hhi[ih] separately, which no buffer holds
            // here.  Lane l recomputes the terms of candidate l mod 8 for the frames l / 8 and l / 8 + 4 from the
            // block and the mixer row; lanes 0..7 then add them up in frame order.
            const unsigned posbits = (sidx >= 0) ? __float_as_uint(cpos) : 0u;
            const unsigned msk = c_tab.mask[pt];
            float2 t0 = make_float2(0.0f, 0.0f), t1 = make_float2(0.0f, 0.0f);
            {
                const unsigned cp = __shfl_sync(0xffffffffu, posbits, lane & 7);
                const int pw = (cp >= N) ? cp - N : cp;
                const int part = lane >> 3;
                if (msk & (1u << part)) {
                    int ih = pw + NFR * part;
                    if (ih >= N) ih -= N;
                    int il = ih + NFR - 6;
                    if (il >= N) il -= N;
                    t0 = scan2_bit0_term(xblk, wrow, il, ih, pp);
                }
                if (part < 2 && (msk & (16u << part))) {
                    int ih = pw + NFR * (part + 4);
                    if (ih >= N) ih -= N;
                    int il = ih + NFR - 6;
                    if (il >= N) il -= N;
                    t1 = scan2_bit0_term(xblk, wrow, il, ih, pp);
                }
            }
            float2 b0 = make_float2(0.0f, 0.0f);
#pragma unroll
            for (int m = 0; m < 6; m++) {
                const int src_lane = (lane & 7) + 8 * (m & 3);
                const float tx = __shfl_sync(0xffffffffu, (m < 4) ? t0.x : t1.x, src_lane);
                const float ty = __shfl_sync(0xffffffffu, (m < 4) ? t0.y : t1.y, src_lane);
                if (msk & (1u << m)) b0 = add2(b0, make_float2(tx, ty));
            }
            if (lane < NCAND) {
                // carrier phase from both sync words (== conj of the scan sum), softbits_kernel.cuh:88-147
                float4 ct = make_float4(__uint_as_float(0u), 0.0f, 0.0f, 0.0f);
                float sre = 0.0f, sim = 0.0f;
                if (sidx >= 0) ct.x = cpos, ct.y = cxb, sre = cre, sim = cim;
                const float phase0 = atan2f(sim, sre);
                sincosf(phase0, &ct.w, &ct.z);
                CT[pt * NCAND + lane] = ct;
                SB0[pt * NCAND + lane] = fmaf(b0.y, ct.z, -__fmul_rn(b0.x, ct.w));  // even bit: Im(v e^{-j phase0})
            }
            SCAN_TICK(11)
        } else if (fi + 1 < f_end) {
            scan2_mix(A, xblk, p.mix + static_cast<size_t>(fi + 1) * N, tid - 32 * DEPTH, SCAN2_THREADS - 32 * DEPTH);
        }
        __syncthreads();
        SCAN_TICK(3)
        // ---- S2a of the next frequency's first tile (reads A, writes HT) || S5 of this one (reads MF, CT, SB0) --
        if (fi + 1 < f_end) scan2_pulse_tile(A, HT, pp, 0, tid);
        SCAN_TICK(4)
        scan_softbits<DEPTH, SCAN2_WARPS>(MF, CT, SB0, p, blk, fi, freq, lane, warp);
        SCAN_TICK(5)
        __syncthreads();
        SCAN_TICK(6)
    }
#ifdef MSK_SCAN_PROF
    if (tid == 0)
        for (int k = 0; k < 16; k++) atomicAdd(&g_scan_prof[k], static_cast<unsigned long long>(prof_[k]));
#endif
}
