// Throughput of the warp-collective and shared-memory instructions scan_kernel's S3 / S4 lean on (sm_100a):
// warp instructions per cycle per SM sub-partition, independent chains, 1..8 warps per sub-partition.
#include <cstdio>
#include <cuda_runtime.h>
#define ITER 256
#define NCH 8
template <int KIND>
__global__ void k(unsigned* out, unsigned seed, long long* cyc) {
    unsigned v[NCH];
    float4 f[NCH];
    __shared__ float4 sm[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = make_float4(i, 0, 0, 0);
    for (int i = 0; i < NCH; i++) v[i] = seed * (threadIdx.x + 1) + i, f[i] = make_float4(i, 1, 2, 3);
    const int lane = threadIdx.x & 31;
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int i = 0; i < NCH; i++) {
                if (KIND == 0) v[i] = __reduce_max_sync(0xffffffffu, v[i]) + lane;          // REDUX.MAX
                if (KIND == 1) v[i] = __ballot_sync(0xffffffffu, v[i] & 1u) + lane;          // VOTE.BALLOT (+ ISETP)
                if (KIND == 2) v[i] = __shfl_sync(0xffffffffu, v[i], (lane + 1) & 31);       // SHFL.IDX
                if (KIND == 3) v[i] = __reduce_add_sync(0xffffffffu, v[i]) ^ lane;           // REDUX.ADD
                if (KIND == 4) { const float2 t = *reinterpret_cast<float2*>(&sm[(v[i] + lane) & 1023]); v[i] += __float_as_uint(t.x); }  // LDS.64 dependent
                if (KIND == 5) { const float4 t = sm[(lane + 32 * i + r) & 1023]; f[i].x += t.x; f[i].y += t.w; }   // LDS.128 independent, conflict-free
                if (KIND == 6) { sm[(lane + 32 * i) & 1023] = f[i]; f[i].x += 1.0f; }                                 // STS.128
                if (KIND == 7) {   // S3's inner step: FMUL, FFMA, REDUX.MAX, compare, predicated STS.128
                    const float s = __uint_as_float(v[i]);
                    const unsigned key = __float_as_uint(fmaf(s, s, s * 0.5f));
                    const unsigned kmax = __reduce_max_sync(0xffffffffu, key);
                    if (key == kmax) sm[(i * 32 + r) & 1023] = f[i];
                    v[i] = key + 1u;
                }
            }
    }
    const long long t1 = clock64();
    unsigned s = 0;
    for (int i = 0; i < NCH; i++) s += v[i] + __float_as_uint(f[i].x + f[i].y);
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int KIND>
void run(const char* name, double per, unsigned* d, long long* dc) {
    for (int wps : {1, 2, 4, 7}) {
        const int threads = wps * 4 * 32;
        k<KIND><<<148, threads>>>(d, 3u, dc);
        k<KIND><<<148, threads>>>(d, 3u, dc);
        long long c = 0;
        cudaMemcpy(&c, dc, 8, cudaMemcpyDeviceToHost);
        printf("%-26s warps/SMSP=%d  cycles=%8lld  ops/cycle/SMSP=%.4f  cycles/op/SMSP=%.2f\n", name, wps, c,
               double(ITER) * 4 * NCH * per * wps / c, c / (double(ITER) * 4 * NCH * per * wps));
    }
}
int main() {
    unsigned* d;
    long long* dc;
    cudaMalloc(&d, 148 * 1024 * 4);
    cudaMalloc(&dc, 8);
    run<0>("REDUX.MAX", 1, d, dc);
    run<3>("REDUX.ADD", 1, d, dc);
    run<1>("VOTE.BALLOT", 1, d, dc);
    run<2>("SHFL.IDX", 1, d, dc);
    run<4>("LDS.64 (dependent)", 1, d, dc);
    run<5>("LDS.128", 1, d, dc);
    run<6>("STS.128", 1, d, dc);
    run<7>("S3 step (key+REDUX+STS)", 1, d, dc);
    printf("err=%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
