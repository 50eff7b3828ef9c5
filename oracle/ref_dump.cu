// ref_dump.cu -- TEST INFRASTRUCTURE.  Runs the REFERENCE's own kernels (compiled from
// /root/reference/src where they lie; nothing is copied into this repository) and dumps their
// stage outputs so that tests/golden/ can pin the oracle and the product bit-for-bit:
//
//   ref_dump stream <read_mode> <method> <center> <width> <step> <depth> <thr> <in.raw> <out_prefix> [max_blocks] [compact]
//       per block: filtered block (complex64 x 5184)        -> <out_prefix>.blocks.c64
//                  every ResultItem after the LDPC kernel    -> <out_prefix>.items.bin  (632 B each)
//                  SNRTracker::getSNRI()                     -> <out_prefix>.snr.i32
//       with a 12th argument "compact" (long streams: 15 MB of items per deep block would not fit a fixture):
//                  only the items with is_message_present, 120 B each {block, item, pos, f0, nbadsync, xb,
//                  num_avg, iters, nhard, pad, message[80]}  -> <out_prefix>.present.bin
//                  number of items with nbadsync <= thr       -> <out_prefix>.ngated.i32
//                  the block's output lines from the reference's own do_decode + ResultFilter (main.cu:390-419),
//                  "<block>\t<line, date masked>"            -> <out_prefix>.lines.txt
//                  SNR as above; no filtered blocks
//   ref_dump ldpc <in_llr.f32> <out.bin>
//       feeds [ncw][128] LLRs straight into ldpc_kernel; per codeword 80+4+4+4 bytes
//       {message[80], found, iters, nhard}
//
// The per-block host steps between fread and do_decode (main.cu:300-332,365-377) are restated here
// in a dozen lines because they live inside the reference's main(); everything on the GPU is the
// reference's code, included below.
#define main reference_main
#include "main.cu"
#undef main

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iomanip>
#include <sstream>
#include <vector>

static std::vector<char> read_file(const char* path) {
    FILE* f = fopen(path, "rb");
    if (!f) {
        perror(path);
        exit(2);
    }
    fseek(f, 0, SEEK_END);
    const long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    std::vector<char> b(n);
    if (n && fread(b.data(), 1, n, f) != (size_t)n) {
        perror("fread");
        exit(2);
    }
    fclose(f);
    return b;
}

__global__ void fill_llr_kernel(MSK144SearchContext ctx, const float* llr, int n, int depth) {
    const int i = blockIdx.x;
    if (i >= n || threadIdx.x != 0) return;
    const unsigned blk = i / (depth * NumCandidatesPerPattern);
    const unsigned rem = i % (depth * NumCandidatesPerPattern);
    ctx.resultKeeper().put_softbits(blk, rem / NumCandidatesPerPattern, rem % NumCandidatesPerPattern, 0, llr + 128 * (size_t)i);
}

static int run_ldpc(const char* in, const char* out) {
    const std::vector<char> raw = read_file(in);
    const size_t ncw = raw.size() / (128 * sizeof(float));
    const float* llr = reinterpret_cast<const float*>(raw.data());
    const int depth = 8, nf = 1025;  // 1025*8*8 = 65600 items per pass
    MSK144SearchContext ctx(1500.0f, 1024.0f, 1.0f, depth, 1000);
    const size_t cap = (size_t)nf * depth * NumCandidatesPerPattern;
    thrust::device_vector<float> d_llr(cap * 128);
    FILE* fo = fopen(out, "wb");
    for (size_t base = 0; base < ncw; base += cap) {
        const int n = (int)std::min(cap, ncw - base);
        ctx.resultKeeper().clear_result();
        thrust::copy(llr + base * 128, llr + (base + n) * 128, d_llr.begin());
        fill_llr_kernel<<<n, 32>>>(ctx, thrust::raw_pointer_cast(d_llr.data()), n, depth);
        // items beyond n keep nbadsync == 0 after the memset and would be decoded too; harmless (all-zero LLRs)
        index_kernel<<<1, NumIndexThreads>>>(ctx);
        const auto blocks = ctx.resultKeeper().getIndexedCandidatesBlocks();
        ldpc_kernel<<<blocks, 128>>>(ctx);
        if (cudaDeviceSynchronize() != cudaSuccess) {
            fprintf(stderr, "cuda error\n");
            return 3;
        }
        thrust::host_vector<ResultKeeper::ResultItem> all = ctx.resultKeeper().get_all_results();
        for (int i = 0; i < n; i++) {
            const auto& it = all[i];
            char msg[80] = {0};
            memcpy(msg, it.message, 77);
            const int found = it.is_message_present ? 1 : 0, iters = it.ldpc_num_iterations, nhard = it.ldpc_num_hard_errors;
            fwrite(msg, 1, 80, fo);
            fwrite(&found, 4, 1, fo);
            fwrite(&iters, 4, 1, fo);
            fwrite(&nhard, 4, 1, fo);
        }
    }
    fclose(fo);
    return 0;
}

static int run_stream(int argc, char** argv) {
    const int read_mode = atoi(argv[2]), method = atoi(argv[3]);
    const float center = atof(argv[4]), width = atof(argv[5]), step = atof(argv[6]);
    const int depth = atoi(argv[7]), thr = atoi(argv[8]);
    const std::vector<char> raw = read_file(argv[9]);
    const std::string prefix = argv[10];
    const long max_blocks = argc > 11 ? atol(argv[11]) : 0;
    const bool compact = argc > 12 && !strcmp(argv[12], "compact");
    const size_t nsamp = raw.size() / 2;
    long nblocks = nsamp >= Num6x864 ? (long)((nsamp - Num6x864) / (Num6x864 / 2)) + 1 : 0;
    if (max_blocks > 0 && nblocks > max_blocks) nblocks = max_blocks;

    MSK144SearchContext ctx(center, width, step, depth, thr);
    Analytic analytic(8192);
    thrust::host_vector<Complex> a_in_host(Num6x864), a_out_host(Num6x864), host_complex_8192(8192);
    thrust::device_vector<Complex> a_in_device(Num6x864), a_out_device(Num6x864);
    SNRTracker snr_tracker;
    FILE* fb = compact ? nullptr : fopen((prefix + ".blocks.c64").c_str(), "wb");
    FILE* fi = fopen((prefix + (compact ? ".present.bin" : ".items.bin")).c_str(), "wb");
    FILE* fs = fopen((prefix + ".snr.i32").c_str(), "wb");
    FILE* fg = compact ? fopen((prefix + ".ngated.i32").c_str(), "wb") : nullptr;
    FILE* fl = compact ? fopen((prefix + ".lines.txt").c_str(), "w") : nullptr;
    ResultFilter result_filter;
    static_assert(sizeof(ResultKeeper::ResultItem) == 632, "ResultItem layout");
    for (long b = 0; b < nblocks; b++) {
        const Complex* host_data = nullptr;
        const Complex* dev_data = nullptr;
        if (read_mode == 1) {
            const short* s = reinterpret_cast<const short*>(raw.data()) + b * (Num6x864 / 2);
            thrust::host_vector<short> real16_buf(s, s + Num6x864);
            // main.cu:301-307
            float sum_rms2 = thrust::reduce(real16_buf.begin(), real16_buf.end(), 0.0f, [](float acc, short const& x) -> double {
                float v = static_cast<float>(x);
                return acc + v * v;
            });
            float rms = sqrt(sum_rms2 / Num6x864);
            float fac = 1.0f / rms;
            if (method == 1) {
                thrust::transform(real16_buf.begin(), real16_buf.end(), host_complex_8192.begin(), [fac](short v) -> Complex { return Complex(fac * v, 0.0f); });
                analytic.execute(host_complex_8192, Num6x864);
                host_data = analytic.getResultHost().data();
                dev_data = thrust::raw_pointer_cast(analytic.getResultDevice());
            } else {
                thrust::transform(real16_buf.begin(), real16_buf.end(), a_in_host.begin(), [fac](short v) -> Complex { return Complex(fac * v, 0.0f); });
                thrust::copy(a_in_host.begin(), a_in_host.end(), a_in_device.begin());
                apply_shift_filter_shift<Num6x864, 32><<<1, 32>>>(thrust::raw_pointer_cast(a_in_device.data()), thrust::raw_pointer_cast(a_out_device.data()));
                thrust::copy(a_out_device.begin(), a_out_device.end(), a_out_host.begin());
                host_data = a_out_host.data();
                dev_data = thrust::raw_pointer_cast(a_out_device.data());
            }
        } else {
            const int8_t* q = reinterpret_cast<const int8_t*>(raw.data()) + b * Num6x864;
            for (int idx = 0; idx < (int)Num6x864; idx++)  // main.cu:365-371
                a_in_host[idx] = Complex(static_cast<float>(q[idx * 2]) / 128.0f, static_cast<float>(q[idx * 2 + 1]) / 128.0f);
            thrust::copy(a_in_host.begin(), a_in_host.end(), a_in_device.begin());
            apply_filter<Num6x864, 32><<<1, 32>>>(thrust::raw_pointer_cast(a_in_device.data()), thrust::raw_pointer_cast(a_out_device.data()));
            thrust::copy(a_out_device.begin(), a_out_device.end(), a_out_host.begin());
            host_data = a_out_host.data();
            dev_data = thrust::raw_pointer_cast(a_out_device.data());
        }
        snr_tracker.process_data(host_data, Num6x864);
        if (compact) {
            // the reference's own per-block host loop (main.cu:390-419): do_decode + ResultFilter, lines tagged with
            // the block they were printed for
            result_filter.blockBegin();
            do_decode(ctx, dev_data, result_filter, snr_tracker);
            result_filter.blockEnd();
            for (auto const& elm : result_filter.getBlockResult()) {
                std::ostringstream os;
                os << "***  "
                   << "snr=" << std::setw(2) << elm.snr << "; "
                   << "f0=" << std::setw(6) << elm.f0 << "; "
                   << "num_avg=" << elm.num_avg << "; "
                   << "nbadsync=" << elm.nbadsync << "; "
                   << "pattern_idx=" << elm.pattern_idx << "; "
                   << "date=xxxxxxxxxxxxxx; "
                   << "msg='" << elm.message << "'"
                   << "; ";
                fprintf(fl, "%ld\t%s\n", b, os.str().c_str());
            }
        } else {
            // do_decode's GPU part, main.cu:461-471
            ctx.resultKeeper().clear_result();
            scan_kernel<<<ctx.getBlocks(), ctx.getThreads()>>>(ctx, dev_data);
            softbits_kernel<<<ctx.getSoftBitsBlocks(), ctx.getSoftBitsThreads()>>>(ctx, dev_data);
            index_kernel<<<1, NumIndexThreads>>>(ctx);
            const auto ldpcBlocks = ctx.resultKeeper().getIndexedCandidatesBlocks();
            if (ldpcBlocks.x > 0) ldpc_kernel<<<ldpcBlocks, 128>>>(ctx);
            if (cudaDeviceSynchronize() != cudaSuccess) {
                fprintf(stderr, "cuda error at block %ld: %s\n", b, cudaGetErrorString(cudaGetLastError()));
                return 3;
            }
        }
        thrust::host_vector<ResultKeeper::ResultItem> all = ctx.resultKeeper().get_all_results();
        if (compact) {
            int ngated = 0;
            for (size_t i = 0; i < all.size(); i++) {
                const auto& it = all[i];
                if (it.nbadsync <= thr) ngated++;
                if (!it.is_message_present) continue;
                struct {
                    int32_t block, item;
                    uint32_t pos;
                    float f0;
                    int32_t nbadsync;
                    float xb;
                    int32_t num_avg, iters, nhard, pad;
                    char message[80];
                } rec;
                static_assert(sizeof(rec) == 120, "compact record layout");
                memset(&rec, 0, sizeof(rec));
                rec.block = (int32_t)b, rec.item = (int32_t)i, rec.pos = it.pos, rec.f0 = it.f0, rec.nbadsync = it.nbadsync;
                rec.xb = it.xb, rec.num_avg = it.num_avg, rec.iters = it.ldpc_num_iterations, rec.nhard = it.ldpc_num_hard_errors;
                memcpy(rec.message, it.message, 77);
                fwrite(&rec, sizeof(rec), 1, fi);
            }
            fwrite(&ngated, 4, 1, fg);
        } else {
            fwrite(host_data, sizeof(Complex), Num6x864, fb);
            fwrite(all.data(), sizeof(ResultKeeper::ResultItem), all.size(), fi);
        }
        const int snri = snr_tracker.getSNRI();
        fwrite(&snri, 4, 1, fs);
    }
    if (fb) fclose(fb);
    if (fg) fclose(fg);
    if (fl) fclose(fl);
    fclose(fi), fclose(fs);
    fprintf(stderr, "ref_dump: %ld blocks, %zu items per block\n", nblocks, (size_t)ctx.getBlocks().x * ctx.scanDepth() * NumCandidatesPerPattern);
    return 0;
}

int main(int argc, char** argv) {
    if (argc >= 4 && !strcmp(argv[1], "ldpc")) return run_ldpc(argv[2], argv[3]);
    if (argc >= 11 && !strcmp(argv[1], "stream")) return run_stream(argc, argv);
    fprintf(stderr, "usage: see header of oracle/ref_dump.cu\n");
    return 2;
}
