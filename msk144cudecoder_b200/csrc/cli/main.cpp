// msk144cudecoder -- command-line front of libmsk144b200.so.
//
// Same surface as the reference executable (main.cu:56-70,98-425): long options only, raw samples on
// stdin (s16le mono 12 kHz, or interleaved s8 I,Q at 12 ksps), one decode line per distinct message
// per block on stdout, a final "Done", banner and warnings on stderr, exit 0 at end of input.
// Code defaults follow main.cu:124-133 (step 2, width 200, depth 4, method 2, threshold 1), not the
// help text.  This program is a thin client of the C ABI: all GPU work happens in the library.
#include <getopt.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "../../../include/msk144_b200.h"

static void show_help(const char* prog) {
    // clang-format off
    std::cout << "Calling conversion: " << prog << " {[--help] | <options> }" << std::endl;
    std::cout << " Where options are: " << std::endl;
    std::cout << "                   --help                      Show this help and exit." << std::endl;
    std::cout << "                   --center-frequency=1500.0   Center frequency in Hz." << std::endl;
    std::cout << "                   --search-step=2.0           Search step in Hz. " << std::endl;
    std::cout << "                   --search-width=100.0        Window in Hz around center frequency to find msk144 signal in." << std::endl;
    std::cout << "                   --scan-depth=[1..8]         The more depth the more averagable patterns will be tried." << std::endl;
    std::cout << "                   --read-mode=[1|2]           1=Audio,16 bit, mono, 12000sps, 1500Hz-recommended center; 2=IQ,8 bit, 12000sps, 0Hz-center. Default mode = 1." << std::endl;
    std::cout << "                   --analytic-method=[1|2]     How to convert real signal to analytic quadrature signal. 1 = FFT; 2 = Shift-left + LPF + Shift-right. Default=2." << std::endl;
    std::cout << "                   --nbadsync-threshold=[1..4] How many errors in sync pattern are acceptable to be passed to LDPC decoder." << std::endl;
    std::cout << " Extensions of this build:" << std::endl;
    std::cout << "                   --device=0                  CUDA device ordinal." << std::endl;
    std::cout << "                   --devices=0,1,...           With --channels=N: split the channels over these GPUs (one host thread and stream set each)." << std::endl;
    std::cout << "                   --batch-blocks=128          Blocks per GPU launch group when input arrives faster than real time." << std::endl;
    std::cout << "                   --per-payload-text          Decode every payload's own text (the reference reuses the first decode of a block)." << std::endl;
    std::cout << "                   --wsjtx-bp                  LDPC with WSJT-X bpdecode128_90 control flow (11 checks, early stop) instead of the reference kernel's." << std::endl;
    std::cout << "                   --channels=1                stdin carries N channels, one sample of each per time step; lines end with channel=<c>." << std::endl;
    std::cout << "                   --wav                       Skip a RIFF/WAVE header at the start of stdin (the reference decodes it as samples)." << std::endl;
    std::cout << "                   --json                      One JSON object per line instead of the '***' line format." << std::endl;
    std::cout << "                   --dedupe[=2]                One line per ping: merge the per-block repeats of a message at most N blocks apart." << std::endl;
    std::cout << "                   --unpack77-lib=PATH         Text decoding through a library exporting __packjt77_MOD_unpack77 (gfortran-built WSJT-X)." << std::endl;
    // clang-format on
}

static const char* mode_string(int m) {
    if (m == 1) return "Audio. 16 bits signed.";
    if (m == 2) return "IQ. 8+8 bits.";
    return "unknown";
}

int main(int argc, char* const argv[]) {
    msk144_config cfg;
    msk144_config_defaults(&cfg);
    cfg.flags = MSK144_FLAG_COMPAT_DECODE_CACHE | MSK144_FLAG_BLOCKING_SYNC;  // main.cu:437-445, main.cu:115
    bool center_set = false, json = false;
    int dedupe_gap = -1;  // < 0: print every per-block line, as the reference does
    int nchan = 1;
    std::vector<int32_t> devices;  // --devices: more than one entry selects the multi-GPU pool
    static struct option long_options[] = {{"help", no_argument, 0, 0},                     // 0
                                           {"center-frequency", required_argument, 0, 0},   // 1
                                           {"search-step", required_argument, 0, 0},        // 2
                                           {"search-width", required_argument, 0, 0},       // 3
                                           {"scan-depth", required_argument, 0, 0},         // 4
                                           {"read-mode", required_argument, 0, 0},          // 5
                                           {"analytic-method", required_argument, 0, 0},    // 6
                                           {"nbadsync-threshold", required_argument, 0, 0}, // 7
                                           {"device", required_argument, 0, 0},             // 8
                                           {"batch-blocks", required_argument, 0, 0},       // 9
                                           {"per-payload-text", no_argument, 0, 0},         // 10
                                           {"wav", no_argument, 0, 0},                      // 11
                                           {"json", no_argument, 0, 0},                     // 12
                                           {"dedupe", optional_argument, 0, 0},             // 13
                                           {"unpack77-lib", required_argument, 0, 0},       // 14
                                           {"channels", required_argument, 0, 0},           // 15
                                           {"devices", required_argument, 0, 0},            // 16
                                           {"wsjtx-bp", no_argument, 0, 0},                 // 17
                                           {0, 0, 0, 0}};
    while (true) {
        int option_index = 0;
        const int c = getopt_long(argc, argv, "", long_options, &option_index);
        if (c == -1) break;
        if (c != 0) continue;  // unknown option: getopt already complained on stderr (main.cu:145-189)
        switch (option_index) {
            case 0: show_help(argv[0]); return 0;
            case 1: cfg.center_frequency_hz = static_cast<float>(atof(optarg)); center_set = true; break;
            case 2: cfg.search_step_hz = static_cast<float>(atof(optarg)); break;
            case 3: cfg.search_width_hz = static_cast<float>(atof(optarg)); break;
            case 4: cfg.scan_depth = atoi(optarg); break;
            case 5: cfg.read_mode = atoi(optarg); break;
            case 6: cfg.analytic_method = atoi(optarg); break;
            case 7: cfg.nbadsync_threshold = atoi(optarg); break;
            case 8: cfg.device = atoi(optarg); break;
            case 9: cfg.max_batch_blocks = atoi(optarg); break;
            case 10: cfg.flags &= ~MSK144_FLAG_COMPAT_DECODE_CACHE; break;
            case 11: cfg.flags |= MSK144_FLAG_SKIP_WAV_HEADER; break;
            case 12: json = true; break;
            case 13: dedupe_gap = optarg ? atoi(optarg) : 2; break;
            case 14:
                if (msk144_set_unpack77_library(optarg) != 0) {
                    std::cerr << "Cannot load __packjt77_MOD_unpack77 from " << optarg << std::endl;
                    return 2;
                }
                break;
            case 15: nchan = atoi(optarg) > 0 ? atoi(optarg) : 1; break;
            case 16:
                devices.clear();
                for (const char* q = optarg; *q;) {
                    devices.push_back(static_cast<int32_t>(strtol(q, const_cast<char**>(&q), 10)));
                    while (*q == ',' || *q == ' ') q++;
                }
                break;
            case 17: cfg.flags |= MSK144_FLAG_WSJTX_BP; break;
            default: show_help(argv[0]); return 0;
        }
    }
    if (!center_set) {  // main.cu:193-208
        if (cfg.read_mode == 1) cfg.center_frequency_hz = 1500.0f;
        else if (cfg.read_mode == 2) cfg.center_frequency_hz = 0.0f;
        else {
            std::cerr << "Wrong read mode " << cfg.read_mode << std::endl;
            return 2;
        }
    }
    msk144_decoder* dec = nullptr;
    msk144_pool* pool = nullptr;
    if (devices.size() == 1) cfg.device = devices[0];
    if (devices.size() > 1) {
        if (msk144_pool_create(&cfg, devices.data(), static_cast<int32_t>(devices.size()), &pool) != MSK144_OK) {
            std::cerr << msk144_pool_last_error() << std::endl;
            return 2;
        }
        dec = msk144_pool_member(pool, 0);
    } else if (msk144_create(&cfg, &dec) != MSK144_OK) {
        std::cerr << msk144_last_error() << std::endl;
        return 2;
    }
    msk144_info info;
    msk144_get_info(dec, &info);
    // main.cu:233-252
    std::cerr << "Actual parameters:" << std::endl
              << "Center Frequency: " << cfg.center_frequency_hz << "Hz" << std::endl
              << "Search Step: " << cfg.search_step_hz << "Hz" << std::endl
              << "Search Width: " << cfg.search_width_hz << "Hz" << std::endl
              << "Scan Depth: " << info.scan_depth << std::endl
              << "Left Boundary: " << info.left_hz << "Hz" << std::endl
              << "Right Boundary: " << info.right_hz << "Hz" << std::endl
              << "Read Mode: (" << mode_string(cfg.read_mode) << ")" << std::endl;
    if (cfg.read_mode == 1) std::cerr << "Analytic Method: " << cfg.analytic_method << std::endl;
    std::cerr << "Badsync Threshold: " << cfg.nbadsync_threshold << std::endl
              << "Frequency hypotheses: " << info.num_freqs << std::endl
              << "Candidates per block: " << info.items_per_block << std::endl
              << "GPU SMs: " << info.sm_count << std::endl;
    if (pool) std::cerr << "GPUs: " << msk144_pool_size(pool) << " (" << nchan << " channels split in contiguous ranges)" << std::endl;
    std::cerr << std::endl;

    const size_t hop_bytes = MSK144_HOP_SAMPLES * 2 * static_cast<size_t>(nchan);  // both modes: 2 bytes per sample
    const size_t block_bytes = MSK144_BLOCK_SAMPLES * 2 * static_cast<size_t>(nchan);
    // one read = up to one launch group of blocks (counted over all channels)
    const size_t group_blocks = static_cast<size_t>(cfg.max_batch_blocks > 0 ? cfg.max_batch_blocks : 128);
    std::vector<unsigned char> buf(hop_bytes * std::max<size_t>(1, group_blocks / static_cast<size_t>(nchan)));
    std::vector<msk144_line> lines(4096);
    std::vector<unsigned char> carry;
    size_t total_bytes = 0;
    int rc_exit = 0;
    msk144_aggregator* agg = nullptr;
    if (dedupe_gap >= 0) msk144_aggregator_create(dedupe_gap, &agg);
    std::vector<msk144_event> events(1024);
    int64_t blocks_done = 0;
    auto print_line = [&](const msk144_line& l) {
        char out[512];
        if (json) msk144_format_line_json(&l, nullptr, out, sizeof(out));
        else msk144_format_line(&l, nullptr, out, sizeof(out));
        if (!json && nchan > 1) std::cout << out << "channel=" << l.channel << "; " << std::endl;
        else std::cout << out << std::endl;
    };
    auto drain_events = [&](int flush) {
        size_t ne = 0;
        do {
            msk144_aggregator_poll(agg, blocks_done, flush, events.data(), events.size(), &ne);
            for (size_t i = 0; i < ne; i++) print_line(events[i].best);
        } while (ne == events.size());
    };
    // stdin is drained by a reader thread while the GPU works on what has arrived so far: real-time input is processed
    // hop by hop as it comes (one read = one push), a file or a fast pipe accumulates into large launch groups instead
    // of being chopped into pipe-buffer sized pieces (a 64 KB pipe read is 12 blocks).
    struct Inbox {
        std::mutex m;
        std::condition_variable cv;
        std::vector<unsigned char> data;
        bool eof = false;
        int err = 0;
    } inbox;
    const size_t inbox_cap = std::max<size_t>(64u << 20, 4 * buf.size());  // back-pressure on a producer that outruns the GPU
    std::thread reader([&] {
        std::vector<unsigned char> tmp(1u << 20);
        while (true) {
            const ssize_t n = read(0, tmp.data(), tmp.size());
            std::unique_lock<std::mutex> lk(inbox.m);
            if (n <= 0) {
                inbox.eof = true, inbox.err = n < 0 ? errno : 0;
                inbox.cv.notify_all();
                return;
            }
            inbox.cv.wait(lk, [&] { return inbox.data.size() < inbox_cap; });
            inbox.data.insert(inbox.data.end(), tmp.data(), tmp.data() + n);
            inbox.cv.notify_all();
        }
    });
    while (true) {
        buf.clear();
        {
            std::unique_lock<std::mutex> lk(inbox.m);
            inbox.cv.wait(lk, [&] { return !inbox.data.empty() || inbox.eof; });
            buf.swap(inbox.data);
            if (buf.empty() && inbox.err) {
                errno = inbox.err;
                perror("read");
                rc_exit = 2;
            }
            inbox.cv.notify_all();
        }
        if (buf.empty()) break;
        const ssize_t n = static_cast<ssize_t>(buf.size());
        total_bytes += static_cast<size_t>(n);
        const auto t0 = std::chrono::steady_clock::now();
        const int64_t before = 0;
        (void)before;
        int prc;
        if (pool) {
            // the pool takes whole time steps (one sample of every channel): keep a partial one for the next read
            carry.insert(carry.end(), buf.data(), buf.data() + n);
            const size_t frame = 2 * static_cast<size_t>(nchan), whole = carry.size() / frame * frame;
            prc = whole ? msk144_pool_push_interleaved(pool, carry.data(), whole, nchan) : MSK144_OK;
            carry.erase(carry.begin(), carry.begin() + whole);
        } else {
            prc = nchan > 1 ? msk144_push_interleaved(dec, buf.data(), static_cast<size_t>(n), nchan)
                            : msk144_push(dec, buf.data(), static_cast<size_t>(n));
        }
        if (prc != MSK144_OK) {
            std::cerr << (pool ? msk144_pool_last_error() : msk144_last_error()) << std::endl;
            rc_exit = 2;
            break;
        }
        size_t got = 0;
        do {
            if (pool) msk144_pool_poll(pool, lines.data(), lines.size(), &got);
            else msk144_poll(dec, lines.data(), lines.size(), &got);
            if (agg) msk144_aggregator_add(agg, lines.data(), got);
            else for (size_t i = 0; i < got; i++) print_line(lines[i]);
        } while (got == lines.size());
        if (agg) {
            // blocks completed so far (WAV header bytes, if any, make this an upper bound by < 1 block: harmless,
            // an event is only held back a little longer)
            blocks_done = total_bytes >= block_bytes ? static_cast<int64_t>((total_bytes - block_bytes) / hop_bytes) + 1 : 0;
            drain_events(0);
        }
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        const double blocks_in = static_cast<double>(n) / hop_bytes;
        if (blocks_in >= 1.0 && ms > 210.0 * blocks_in)  // main.cu:397-403, scaled to the group size
            std::cerr << "Warning: Working loop takes too much time: " << ms << " ms"
                      << " of " << 210.0 * blocks_in << " ms max." << std::endl;
    }
    bool reader_finished;
    {
        std::lock_guard<std::mutex> lk(inbox.m);
        reader_finished = inbox.eof;
    }
    if (rc_exit != 0 && !reader_finished) {
        // a push failed while input is still arriving: the reader may sit in read() for ever; leave without joining it
        std::cout.flush();
        std::cerr.flush();
        _exit(rc_exit);
    }
    // end of input, or a read() error (the reader has returned in both cases): the reference ends its loop on the short
    // read and prints its epilogue, and so does this; the exit code still reports the error
    reader.join();
    // main.cu:274-278,289-293: the final short fread ends the loop
    size_t leftover;
    if (total_bytes < block_bytes) leftover = total_bytes;
    else leftover = (total_bytes - block_bytes) % hop_bytes;
    const size_t rc_items = (cfg.read_mode == 1) ? leftover / 2 : leftover;
    if (agg) {
        drain_events(1);
        msk144_aggregator_destroy(agg);
    }
    std::cerr << "Incomplete read error. rc=" << rc_items << std::endl;
    std::cout << "Done" << std::endl;
    if (pool) msk144_pool_destroy(pool);
    else msk144_destroy(dec);
    return rc_exit;
}
