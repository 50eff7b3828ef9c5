// Host harness for csrc/chz_fft.h (TEST INFRASTRUCTURE): walks the thread index of every Stockham stage in a loop and
// compares the result with a direct double-precision DFT.  Prints "M stages max_err" per size; exit 1 on a miss.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../msk144cudecoder_b200/csrc/chz_fft.h"

int main() {
    int bad = 0;
    for (int M = 2; M <= 256; M++) {
        chzfft::StagePlan sp[chzfft::MAX_STAGES];
        const int ns = chzfft::plan(M, sp);
        if (ns == 0) continue;
        const int nfr = 5, nthreads = 96;
        const int RL = chzfft::row_len(M);
        std::vector<float2> a(nfr * RL), b(nfr * RL), tw(M);
        for (int i = 0; i < M; i++) tw[i] = float2{(float)std::cos(-2.0 * M_PI * i / M), (float)std::sin(-2.0 * M_PI * i / M)};
        srand(M);
        for (auto& v : a) v = float2{(float)(rand() % 511 - 255), (float)(rand() % 511 - 255)};
        std::vector<float2> x(nfr * M);
        for (int f = 0; f < nfr; f++)
            for (int q = 0; q < M; q++) x[f * M + q] = a[f * RL + chzfft::pad(q)];
        float2 *src = a.data(), *dst = b.data();
        for (int s = 0; s < ns; s++) {
            for (int tid = 0; tid < nthreads; tid++) chzfft::run_stage<true>(src, dst, M, nfr, sp[s], tw.data(), tid, nthreads);
            std::swap(src, dst);
        }
        double err = 0, mag = 0;
        for (int f = 0; f < nfr; f++)
            for (int k = 0; k < M; k++) {
                double re = 0, im = 0;
                for (int q = 0; q < M; q++) {
                    const double ph = -2.0 * M_PI * ((long long)k * q % M) / M;
                    re += x[f * M + q].x * std::cos(ph) - x[f * M + q].y * std::sin(ph);
                    im += x[f * M + q].x * std::sin(ph) + x[f * M + q].y * std::cos(ph);
                }
                const float2 got = src[f * RL + chzfft::pad(k)];
                err = std::fmax(err, std::hypot(got.x - re, got.y - im));
                mag = std::fmax(mag, std::hypot(re, im));
            }
        std::printf("%d %d %.3g\n", M, ns, err / mag);
        if (!(err / mag < 2e-6)) bad++;
    }
    return bad ? 1 : 0;
}
