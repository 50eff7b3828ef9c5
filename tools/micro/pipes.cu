// Pipe-throughput micro-benchmark for sm_100a: warp instructions per cycle per SM sub-partition (SMSP)
// for the instruction kinds the LDPC / scan kernels are made of.  nvcc -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
#define ITER 256
#define NCH 8
#ifndef LDPC_MIX_13  // measured loop sizes / 32; rebuild with -D after changing the mix
#define LDPC_MIX_13 10.875
#define LDPC_MIX_14 7.78
#define LDPC_MIX_15 7.72
#endif
template <int KIND>
__global__ void k(float* out, float seed, int warps_note, long long* cyc) {
    float a[NCH], b[NCH];
    u64 pa[NCH], pb[NCH];
    unsigned ia[NCH], ib[NCH];
    __shared__ u64 sm[NCH * 32 * 16];
    const unsigned saddr = static_cast<unsigned>(__cvta_generic_to_shared(sm + (threadIdx.x & 511)));
    for (int i = threadIdx.x; i < NCH * 512; i += blockDim.x) sm[i] = 0;
    for (int i = 0; i < NCH; i++) {
        a[i] = seed + i + threadIdx.x * 1e-3f, b[i] = seed * 0.5f + i;
        ia[i] = threadIdx.x * 7 + i;
        asm("mov.b64 %0, {%1,%2};" : "=l"(pa[i]) : "f"(a[i]), "f"(b[i]));
        pb[i] = pa[i], ib[i] = ia[i] * 3u + 1u;
    }
    u64 pm, pc;
    asm("mov.b64 %0, {%1,%2};" : "=l"(pm) : "f"(0.999f), "f"(1.001f));
    asm("mov.b64 %0, {%1,%2};" : "=l"(pc) : "f"(seed), "f"(-seed));
    const float m = 0.999f + seed * 1e-9f, c = seed;
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int i = 0; i < NCH; i++) {
                if (KIND == 0) a[i] = fmaf(a[i], m, c);                                     // FFMA 3-reg
                if (KIND == 1) asm("fma.rn.f32x2 %0,%0,%1,%2;" : "+l"(pa[i]) : "l"(pm), "l"(pc));  // FFMA2
                if (KIND == 2) asm("mul.rn.f32x2 %0,%0,%1;" : "+l"(pa[i]) : "l"(pm));       // FMUL2
                if (KIND == 3) asm("add.rn.f32x2 %0,%0,%1;" : "+l"(pa[i]) : "l"(pc));       // FADD2
                if (KIND == 4) ia[i] = (ia[i] ^ (ia[i] >> 3)) & 0x7fffffffu | it;            // LOP3/SHF
                if (KIND == 5) a[i] = (a[i] > b[i]) ? a[i] * 0.5f : a[i] + 1.0f;             // mixed
                if (KIND == 6) { asm("fma.rn.f32x2 %0,%0,%1,%2;" : "+l"(pa[i]) : "l"(pm), "l"(pc)); ia[i] = (ia[i] ^ 0x55u) + it; }  // FFMA2 + ALU
                if (KIND == 7) asm("ex2.approx.ftz.f32 %0,%0;" : "+f"(a[i]));               // MUFU
                if (KIND == 8) { asm("fma.rn.f32x2 %0,%0,%1,%2;" : "+l"(pa[i]) : "l"(pm), "l"(pc)); a[i] = fmaf(a[i], m, c); }  // FFMA2 + FFMA
                if (KIND == 9) a[i] = a[i] * m;                                              // FMUL
                if (KIND == 10) { asm("fma.rn.f32x2 %0,%0,%1,%2;" : "+l"(pa[i]) : "l"(pm), "l"(pc)); ia[i] = (ia[i] ^ 0x55u) + it; ia[i] = (ia[i] >> 1) | 0x40000000u; }  // FFMA2 + 2 ALU
                if (KIND == 11) a[i] = fmaf(a[i], 0.999f, 0.5f);                             // FFMA imm
                if (KIND == 12) { a[i] = fmaf(a[i], m, c); ia[i] = (ia[i] ^ 0x55u) + it; }   // FFMA + ALU
                if (KIND == 13 || KIND == 14 || KIND == 15) {
                    // the LDPC kernel's dynamic mix (profiles/r01_v6_ncu_full.txt source page): per two steps
                    // 8 packed FP32, 7 ALU (LOP3 / SHF / FMNMX / FSETP+FSEL), 1 MUFU, 1 LDS.64 + 1 STS.64 = 18;
                    // eight independent chains; KIND 14 drops the shared-memory pair, KIND 15 the MUFU as well
                    asm("fma.rn.f32x2 %0,%0,%1,%2;" : "+l"(pa[i]) : "l"(pm), "l"(pc));
                    asm("mul.rn.f32x2 %0,%0,%1;" : "+l"(pb[i]) : "l"(pm));
                    asm("fma.rn.f32x2 %0,%0,%1,%2;" : "+l"(pa[i]) : "l"(pm), "l"(pc));
                    asm("add.rn.f32x2 %0,%0,%1;" : "+l"(pb[i]) : "l"(pc));
                    asm("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(ia[i]) : "r"(ib[i]), "r"(it));
                    asm("max.f32 %0,%0,%1;" : "+f"(b[i]) : "f"(a[i]));
                    if (r & 1) {
                        asm("{.reg .pred p; setp.gt.f32 p,%0,%1; selp.f32 %0,%0,%2,p;}" : "+f"(a[i]) : "f"(b[i]), "f"(c));
                        if (KIND != 15) asm("ex2.approx.ftz.f32 %0,%0;" : "+f"(a[i]));
                        else asm("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(ib[i]) : "r"(ia[i]), "r"(it));
                    } else if (KIND == 13) {
                        u64 t;
                        asm volatile("ld.shared.b64 %0,[%1];" : "=l"(t) : "r"(saddr + 4096u * i));
                        asm("add.rn.f32x2 %0,%0,%1;" : "+l"(t) : "l"(pc));
                        asm volatile("st.shared.b64 [%0],%1;" ::"r"(saddr + 4096u * i), "l"(t) : "memory");
                    } else {
                        asm("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(ia[i]) : "r"(ib[i]), "r"(it));
                        asm("shf.l.wrap.b32 %0,%0,%0,%1;" : "+r"(ib[i]) : "r"(5));
                    }
                }
            }
        }
    }
    long long t1 = clock64();
    float s = 0;
    for (int i = 0; i < NCH; i++) {
        float x, y;
        asm("mov.b64 {%0,%1}, %2;" : "=f"(x), "=f"(y) : "l"(pa[i]));
        s += a[i] + x + y + __uint_as_float(ia[i]);
        asm("mov.b64 {%0,%1}, %2;" : "=f"(x), "=f"(y) : "l"(pb[i]));
        s += b[i] + x + y + __uint_as_float(ib[i]);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int KIND>
void run(const char* name, double per_iter_instr, float* d, long long* dc) {
    for (int wps : {1, 2, 4, 8}) {  // warps per SMSP
        const int threads = wps * 4 * 32;
        k<KIND><<<148, threads>>>(d, 1.0f, wps, dc);
        k<KIND><<<148, threads>>>(d, 1.0f, wps, dc);
        long long c = 0;
        cudaMemcpy(&c, dc, 8, cudaMemcpyDeviceToHost);
        const double instr_per_warp = double(ITER) * 4 * NCH * per_iter_instr;
        printf("%-14s warps/SMSP=%d  cycles=%lld  warp-instr/cycle/SMSP=%.3f\n", name, wps, c, instr_per_warp * wps / c);
    }
}
int main() {
    float* d;
    long long* dc;
    cudaMalloc(&d, 148 * 1024 * 4);
    cudaMalloc(&dc, 8);
    run<0>("FFMA", 1, d, dc);
    run<11>("FFMA.imm", 1, d, dc);
    run<9>("FMUL", 1, d, dc);
    run<1>("FFMA2", 1, d, dc);
    run<2>("FMUL2", 1, d, dc);
    run<3>("FADD2", 1, d, dc);
    run<4>("ALU(3)", 3, d, dc);
    run<7>("MUFU.EX2", 1, d, dc);
    run<12>("FFMA+2ALU", 3, d, dc);
    run<6>("FFMA2+2ALU", 3, d, dc);
    run<10>("FFMA2+4ALU", 5, d, dc);
    run<8>("FFMA2+FFMA", 2, d, dc);
    // instructions per step = SASS instructions of the loop body / 32 steps (cuobjdump, counted after compilation)
    run<13>("LDPC-mix", LDPC_MIX_13, d, dc);
    run<14>("LDPC-mix-noLSU", LDPC_MIX_14, d, dc);
    run<15>("LDPC-mix-noXU", LDPC_MIX_15, d, dc);
    printf("err=%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
