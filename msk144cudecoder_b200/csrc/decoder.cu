// decoder.cu -- host orchestration and C ABI of libmsk144b200.so (include/msk144_b200.h).
//
// One decoder = one GPU + one stream.  A launch group is up to max_batch_blocks blocks (of one stream
// or of many channels):  frontend_kernel -> scan_kernel -> ldpc_kernel, then two small D2H copies
// (counters + SNR powers, then the compact list of LDPC successes).  The host then replays what
// main.cu:480-525 and result_filter.cpp do per block.  There is no CPU fallback for any GPU stage.
#include "../../include/msk144_b200.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <memory>
#include <mutex>
#include <new>
#include <stdexcept>
#include <string>
#include <vector>

#include "host/postproc.h"
#include "host/unpack77.h"
#include "kernels.cuh"
#include "ldpc_code.h"

using namespace msk;

namespace {

thread_local std::string g_err;

int fail(int code, const std::string& what) {
    g_err = what;
    return code;
}

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(MSK144_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_));         \
    } while (0)

struct Grid {  // msk_context.cuh:95-113
    int nf;
    float if1;
};
Grid make_grid(float width, float step) {
    const float half_len_in_hz = width / 2;
    const float half_len_cnt = half_len_in_hz / step;
    const int half_len = static_cast<int>(half_len_cnt);
    Grid g;
    g.nf = half_len * 2 + 1;
    g.if1 = -1 * half_len * step;
    return g;
}

void build_tables(DevTables* t) {
    std::memset(t, 0, sizeof(*t));
    const float pi = 3.141592654f;  // CUDART_PI_F
    for (size_t i = 0; i < 12; i++) {
        const float angle = i * pi / 12.0f;  // msk_context.cuh:137-145
        t->pp[i] = sinf(angle);
    }
    const int s8_org[8] = {0, 1, 1, 1, 0, 0, 1, 0};  // msk_context.cuh:149
    for (int i = 0; i < 8; i++) t->s8[i] = static_cast<float>(2 * s8_org[i] - 1);
    const int masks[8][6] = {{1, 0, 0, 0, 0, 0}, {1, 1, 0, 0, 0, 0}, {1, 1, 1, 0, 0, 0}, {1, 1, 1, 1, 0, 0},
                             {1, 1, 1, 1, 1, 0}, {1, 1, 1, 1, 1, 1}, {1, 0, 0, 1, 0, 0}, {1, 0, 0, 1, 1, 0}};
    for (int p = 0; p < 8; p++) {
        int m = 0, w = 0;
        for (int i = 0; i < 6; i++)
            if (masks[p][i]) m |= 1 << i, w++;
        t->mask[p] = static_cast<unsigned char>(m);
        t->num_avg[p] = static_cast<unsigned char>(w);
    }
    int cnt[128] = {0};
    for (int c = 0; c < 38; c++) {
        int nrw = 0;
        for (int s = 0; s < 11; s++)
            if (MSK144_NM[c][s] >= 0) nrw++;
        for (int s = 0; s < nrw; s++) {
            const int b = MSK144_NM[c][s];
            t->edge[b][cnt[b]++] = static_cast<unsigned short>(c | (s << 6) | (nrw << 10));
            t->chkmask[c][b >> 5] |= 1u << (b & 31);
        }
    }
    // x^(95-k) mod (x^13 + 0x15D7): contribution of message bit k to the CRC (ldpc_kernel.cuh:32-63)
    for (int k = 0; k < 77; k++) {
        unsigned rem = 0;
        for (int i = 0; i < 96; i++) {
            rem = (rem << 1) | (i == k ? 1u : 0u);
            if (rem & 0x2000u) rem ^= 0x2000u | 0x15D7u;
        }
        t->crc_r[k] = static_cast<unsigned short>(rem & 0x1fffu);
    }
}

// scan_kernel is instantiated per scan depth (msk_context.cuh:28-33 clamps it to 1..8)
using ScanEntry = void (*)(ScanParams);
ScanEntry scan_entry(int depth) {
    static const ScanEntry tab[8] = {scan_kernel<1>, scan_kernel<2>, scan_kernel<3>, scan_kernel<4>,
                                     scan_kernel<5>, scan_kernel<6>, scan_kernel<7>, scan_kernel<8>};
    return tab[std::max(1, std::min(8, depth)) - 1];
}
ScanEntry scan2_entry(int depth) {
    static const ScanEntry tab[8] = {scan2_kernel<1>, scan2_kernel<2>, scan2_kernel<3>, scan2_kernel<4>,
                                     scan2_kernel<5>, scan2_kernel<6>, scan2_kernel<7>, scan2_kernel<8>};
    return tab[std::max(1, std::min(8, depth)) - 1];
}
// MSK144_SCAN_FORM=2 selects the second form of the kernel (scan2.cuh: two 448-thread CTAs per SM, 110 KB each); the
// default is the first form (one 864-thread CTA per SM), 1 % faster on the deep scan (profiles/r02_scan_notes.md).
// Both produce the same bits; the GPU tests run under either.
bool scan_form2() {
    static const bool v = [] {
        const char* e = std::getenv("MSK144_SCAN_FORM");
        return e && std::atoi(e) == 2;
    }();
    return v;
}
void launch_scan(int depth, dim3 grid, cudaStream_t st, const ScanParams& sp) {
    if (scan_form2())
        scan2_entry(depth)<<<grid, SCAN2_THREADS, SCAN2_SMEM, st>>>(sp);
    else
        scan_entry(depth)<<<grid, SCAN_THREADS, SCAN_SMEM, st>>>(sp);
}

// Page-locked storage for the streaming input buffers: msk144_push uploads straight from them, so the copy is a DMA
// from the caller's bytes' first resting place instead of the driver's staging copy of a pageable vector.
template <class T>
struct PinnedAlloc {
    using value_type = T;
    PinnedAlloc() = default;
    template <class U>
    PinnedAlloc(const PinnedAlloc<U>&) {}
    T* allocate(size_t n) {
        void* p = nullptr;
        if (cudaMallocHost(&p, n * sizeof(T)) != cudaSuccess) throw std::bad_alloc();
        return static_cast<T*>(p);
    }
    void deallocate(T* p, size_t) { cudaFreeHost(p); }
    template <class U>
    bool operator==(const PinnedAlloc<U>&) const { return true; }
    template <class U>
    bool operator!=(const PinnedAlloc<U>&) const { return false; }
};
using PinnedBytes = std::vector<uint8_t, PinnedAlloc<uint8_t>>;

// Append to a streaming input buffer.  The buffer is drained to less than one block after every push, so growing it
// to exactly what this push needs (rounded up to 64 KB) instead of letting the vector double keeps the page-locked
// allocation at the size of the largest push seen, and reallocation (a cudaMallocHost plus a synchronising
// cudaFreeHost) stops after the first few pushes.  PinnedAlloc reports failure by throwing; nothing may throw across
// the C ABI, so it ends here.
int append_pinned(PinnedBytes* v, const uint8_t* b, size_t nbytes) {
    try {
        const size_t need = v->size() + nbytes;
        if (need > v->capacity()) v->reserve((need + 65535) / 65536 * 65536);
        v->insert(v->end(), b, b + nbytes);
    } catch (const std::bad_alloc&) {
        return fail(MSK144_ENOMEM, "page-locked input buffer: cudaMallocHost failed");
    } catch (const std::length_error&) {
        return fail(MSK144_ENOMEM, "input buffer larger than the address space");
    }
    return MSK144_OK;
}

struct RawResult {
    int64_t gblock;
    ResultRec r;
};

// What one channel carries from block to block: the SNR recurrence (snr_tracker.cu:39-50) and the call-sign hash
// tables of its text decoder (module state of packjt77 in the reference process).
struct ChanState {
    mskhost::SnrTracker snr;
    void* hash = nullptr;
    ChanState() : hash(msk144_hash_state_new()) {}
    ChanState(const ChanState&) = delete;
    ChanState& operator=(const ChanState&) = delete;
    ~ChanState() { msk144_hash_state_free(hash); }
    void reset() {
        snr.reset();
        msk144_hash_state_clear(hash);
    }
};

// ResultFilter output of one block (postproc.cpp) as C ABI lines
void append_block_lines(const std::vector<mskhost::Line>& bl, int64_t block_index, int channel, std::vector<msk144_line>* lines) {
    for (const auto& l : bl) {
        msk144_line o;
        std::memset(&o, 0, sizeof(o));
        o.block_index = block_index;
        o.channel = channel;
        o.snr = l.snr;
        o.f0 = l.f0;
        o.num_avg = l.num_avg;
        o.nbadsync = l.nbadsync;
        o.pattern_idx = l.pattern_idx;
        std::memcpy(o.payload, l.payload, 80);
        std::snprintf(o.text, sizeof(o.text), "%s", l.message.c_str());
        lines->push_back(o);
    }
}

}  // namespace

struct msk144_decoder {
    msk144_config cfg{};
    int nf = 0, depth = 0, items = 0, sm_count = 0;
    float if1 = 0.f;
    cudaStream_t stream = nullptr;
    struct EvQuad {
        cudaEvent_t e[4];
        int nb;
    };
    std::vector<EvQuad> ev_pool;  // one quad per launch group since the last collect
    size_t ev_used = 0;
    int64_t h2d_bytes = 0, d2h_bytes = 0;

    // Device buffers of one launch group.  Two slots: while the host fetches and unpacks group g, the
    // kernels of group g+1 already run (copies go through copy_stream, ordered by the slot's event).
    struct Slot {
        float2* d_blocks = nullptr;
        float* d_powers = nullptr;
        float* d_fac = nullptr;  // 1 / rms per block (rms_kernel)
        CandRec* d_cands = nullptr;
        QueueRec* d_queue = nullptr;
        ResultRec* d_results = nullptr;
        int* d_counters = nullptr;  // [0] queue count, [1] result count
        int* h_counters = nullptr;  // pinned
        float* h_powers = nullptr;  // pinned [cap][8]
        cudaEvent_t done = nullptr; // recorded on the launching stream after the slot's last kernel
    };
    Slot slot[2];
    int cap_blocks = 0;  // device capacity (blocks per launch group)
    size_t queue_cap = 0, result_cap = 0;
    cudaStream_t copy_stream = nullptr;
    void* d_raw = nullptr;
    size_t d_raw_bytes = 0;
    // msk144_decode_host: the input goes up in pieces on its own stream while the first launch groups already run; a
    // group waits for the piece that holds its last sample (pieces and groups both advance through d_raw in order)
    static constexpr size_t UP_PIECE = 8u << 20;
    cudaStream_t up_stream = nullptr;
    std::vector<cudaEvent_t> up_ev;
    size_t up_pieces = 0;  // pieces in flight for the current call (0: input already resident)
    long long up_waited = -1;
    // MSK144_FLAG_BLOCKING_SYNC: host waits of a live hop sleep on this event instead of spinning (main.cu:115)
    cudaEvent_t wait_ev = nullptr;
    bool sleep_wait = false;  // set per decode: the flag is on and the call is a live hop (a few blocks)
    int host_wait(cudaStream_t s);
    int upload_pieces(const void* h_src, size_t bytes);
    std::vector<float>* powers_sink = nullptr;  // msk144_decode_chunk_host: decode_common also hands the sub-block powers out
    int wait_for_input(size_t end_byte, cudaStream_t st);
    float2* d_twiddle = nullptr;
    float* d_hfilt = nullptr;
    unsigned short* d_edge_tab = nullptr;
    float2* d_mix = nullptr;  // [nf][N] mixer table e^{j phi(f, i)}

    ResultRec* h_results = nullptr;  // pinned
    size_t h_results_cap = 0;

    // accumulated by the last decode call
    msk144_stage_times times{};
    int64_t launches = 0;

    // streaming state (msk144_push / msk144_poll)
    PinnedBytes pending;  // raw bytes from the start of the next unprocessed block (page-locked)
    int share_twins = std::getenv("MSK144_NO_TWINS") ? 0 : 1;  // A/B knob; results are identical either way
    unsigned apmask[4] = {0u, 0u, 0u, 0u};  // MSK144_FLAG_WSJTX_BP: a-priori bits (msk144_set_ap_mask)
    // multi-channel streaming state (msk144_push_interleaved): raw interleaved bytes, one SNR recurrence per channel
    int il_nchan = 0;
    PinnedBytes il_pending;
    std::vector<std::unique_ptr<ChanState>> il_state;
    int64_t il_next_block = 0;
    bool wav_checked = false;      // MSK144_FLAG_SKIP_WAV_HEADER: the start of the stream has been examined
    int64_t next_block = 0;
    ChanState stream_state;
    std::deque<msk144_line> out_lines;
    std::vector<msk144_line> last_lines;      // output of the most recent batch decode (msk144_last_output)
    std::vector<msk144_result> last_results;

    ~msk144_decoder() { release(); }
    void release();
    int ensure_capacity(int nblocks);
    int ensure_raw(size_t bytes);
    int ensure_host_results(size_t n);
    int launch_frontend(Slot& sl, const void* d_in, long long chan_stride, int bpc, long long block0, int nb, cudaStream_t st);
    int launch_group(Slot& sl, const void* d_in, long long chan_stride, int bpc, long long block0, int nb, cudaStream_t st);
    int fetch_group(Slot& sl, int nb, std::vector<RawResult>* out, int64_t gblock0, std::vector<float>* powers);
    int next_events(EvQuad** q);
    void collect_times();
    // frequencies walked by one scan CTA: enough to amortise the pipeline prologue, few enough that the grid
    // still holds several waves of CTAs per SM
    int scan_fpc(int nb) const {
        int cap = 8;
        if (const char* e = std::getenv("MSK144_SCAN_FPC")) cap = std::max(1, std::atoi(e));  // tuning knob
        return std::max(1, std::min(cap, static_cast<int>((static_cast<long long>(nf) * nb) / (sm_count * 8))));
    }
    void fill_ldpc_tables(LdpcParams* lp) const {
        lp->edge_tab = d_edge_tab;
        // platanh segment constants (ldpc_kernel.cuh:75-90) and their correctly rounded reciprocals
        const float c[4] = {0.0f, 0.4064f, 0.8378f, 0.9914f}, dv[4] = {0.83f, 0.322f, 0.0524f, 0.0012f};
        for (int k = 0; k < 4; k++) {
            volatile float r = 1.0f / dv[k];  // IEEE single division, as the exhaustive division proof assumes
            lp->pc[k] = c[k], lp->pd[k] = dv[k], lp->pr[k] = r;
            lp->pnd[k] = -0.5f * dv[k], lp->pr2[k] = 2.0f * r;
        }
    }
    int decode_common(const void* d_in, size_t spc, int nchan, cudaStream_t st, int64_t block_index_base,
                      ChanState* const* carry, std::vector<msk144_line>* lines, std::vector<msk144_result>* results);
};

void msk144_decoder::release() {
    cudaSetDevice(cfg.device);
    for (Slot& sl : slot) {
        cudaFree(sl.d_blocks), cudaFree(sl.d_powers), cudaFree(sl.d_fac), cudaFree(sl.d_cands), cudaFree(sl.d_queue), cudaFree(sl.d_results);
        cudaFree(sl.d_counters);
        if (sl.h_counters) cudaFreeHost(sl.h_counters);
        if (sl.h_powers) cudaFreeHost(sl.h_powers);
        if (sl.done) cudaEventDestroy(sl.done);
        sl = Slot{};
    }
    cudaFree(d_raw), cudaFree(d_twiddle), cudaFree(d_hfilt), cudaFree(d_edge_tab), cudaFree(d_mix);
    d_edge_tab = nullptr, d_mix = nullptr, d_raw = nullptr, d_twiddle = nullptr, d_hfilt = nullptr;
    if (h_results) cudaFreeHost(h_results);
    h_results = nullptr;
    for (auto& q : ev_pool)
        for (auto& e : q.e)
            if (e) cudaEventDestroy(e), e = nullptr;
    ev_pool.clear();
    ev_used = 0;
    if (stream) cudaStreamDestroy(stream), stream = nullptr;
    if (copy_stream) cudaStreamDestroy(copy_stream), copy_stream = nullptr;
    if (wait_ev) cudaEventDestroy(wait_ev), wait_ev = nullptr;
    for (auto& e : up_ev) cudaEventDestroy(e);
    up_ev.clear();
    if (up_stream) cudaStreamDestroy(up_stream), up_stream = nullptr;
    cap_blocks = 0;
}

int msk144_decoder::ensure_capacity(int nblocks) {
    if (nblocks <= cap_blocks) return 0;
    CK(cudaDeviceSynchronize());
    cap_blocks = 0;
    const size_t nb = static_cast<size_t>(nblocks);
    queue_cap = nb * items;  // worst case: every candidate passes the sync gate
    result_cap = nb * items;
    if (queue_cap > 0x7fffffffull) return fail(MSK144_EINVAL, "batch too large");
    for (Slot& sl : slot) {
        cudaFree(sl.d_blocks), cudaFree(sl.d_powers), cudaFree(sl.d_fac), cudaFree(sl.d_cands), cudaFree(sl.d_queue), cudaFree(sl.d_results);
        sl.d_blocks = nullptr, sl.d_powers = nullptr, sl.d_fac = nullptr, sl.d_cands = nullptr, sl.d_queue = nullptr, sl.d_results = nullptr;
        if (sl.h_powers) cudaFreeHost(sl.h_powers), sl.h_powers = nullptr;
        CK(cudaMalloc(&sl.d_blocks, nb * N * sizeof(float2)));
        CK(cudaMalloc(&sl.d_powers, nb * 8 * sizeof(float)));
        CK(cudaMalloc(&sl.d_fac, nb * sizeof(float)));
        CK(cudaMalloc(&sl.d_cands, nb * items * sizeof(CandRec)));
        CK(cudaMalloc(&sl.d_queue, queue_cap * sizeof(QueueRec)));
        CK(cudaMalloc(&sl.d_results, result_cap * sizeof(ResultRec)));
        CK(cudaMallocHost(&sl.h_powers, nb * 8 * sizeof(float)));
    }
    cap_blocks = nblocks;
    return 0;
}

int msk144_decoder::ensure_raw(size_t bytes) {
    if (bytes <= d_raw_bytes) return 0;
    cudaFree(d_raw);
    d_raw = nullptr, d_raw_bytes = 0;
    CK(cudaMalloc(&d_raw, bytes));
    d_raw_bytes = bytes;
    return 0;
}

int msk144_decoder::host_wait(cudaStream_t s) {
    if (sleep_wait && wait_ev) {
        CK(cudaEventRecord(wait_ev, s));
        CK(cudaEventSynchronize(wait_ev));
    } else {
        CK(cudaStreamSynchronize(s));
    }
    return 0;
}

int msk144_decoder::upload_pieces(const void* h_src, size_t bytes) {
    if (!up_stream) CK(cudaStreamCreateWithFlags(&up_stream, cudaStreamNonBlocking));
    up_pieces = (bytes + UP_PIECE - 1) / UP_PIECE;
    up_waited = -1;
    while (up_ev.size() < up_pieces) {
        cudaEvent_t e;
        CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        up_ev.push_back(e);
    }
    for (size_t i = 0; i < up_pieces; i++) {
        const size_t off = i * UP_PIECE, len = std::min(UP_PIECE, bytes - off);
        CK(cudaMemcpyAsync(static_cast<char*>(d_raw) + off, static_cast<const char*>(h_src) + off, len, cudaMemcpyHostToDevice, up_stream));
        CK(cudaEventRecord(up_ev[i], up_stream));
    }
    h2d_bytes += bytes;
    return 0;
}

int msk144_decoder::wait_for_input(size_t end_byte, cudaStream_t st) {
    if (up_pieces == 0 || end_byte == 0) return 0;
    const long long piece = static_cast<long long>(std::min((end_byte - 1) / UP_PIECE, up_pieces - 1));
    if (piece > up_waited) {
        CK(cudaStreamWaitEvent(st, up_ev[static_cast<size_t>(piece)], 0));
        up_waited = piece;
    }
    return 0;
}

int msk144_decoder::ensure_host_results(size_t n) {
    if (n <= h_results_cap) return 0;
    if (h_results) cudaFreeHost(h_results), h_results = nullptr;
    h_results_cap = 0;
    size_t want = std::max<size_t>(n, 4096);
    want += want / 2;
    CK(cudaMallocHost(&h_results, want * sizeof(ResultRec)));
    h_results_cap = want;
    return 0;
}

int msk144_decoder::next_events(EvQuad** q) {
    if (ev_used == ev_pool.size()) {
        EvQuad n{};
        for (auto& e : n.e) CK(cudaEventCreate(&e));
        ev_pool.push_back(n);
    }
    *q = &ev_pool[ev_used++];
    return 0;
}

// Sum the event pairs of every launch group recorded since the last collect.  Caller has synchronised.
void msk144_decoder::collect_times() {
    for (size_t i = 0; i < ev_used; i++) {
        EvQuad& q = ev_pool[i];
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, q.e[0], q.e[1]) == cudaSuccess) times.frontend_ms += ms;
        if (cudaEventElapsedTime(&ms, q.e[1], q.e[2]) == cudaSuccess) times.scan_ms += ms;
        if (cudaEventElapsedTime(&ms, q.e[2], q.e[3]) == cudaSuccess) times.ldpc_ms += ms;
        if (cudaEventElapsedTime(&ms, q.e[0], q.e[3]) == cudaSuccess) times.total_ms += ms;
        times.blocks += q.nb;
        times.groups += 1;
    }
    ev_used = 0;
}

// The front end of one launch group.  With more blocks than SMs the per-block 1/rms (a sequential FP32 sum in the
// reference, main.cu:301-307) is taken out of the front-end CTAs into rms_kernel, which runs 32 of those chains per warp.
int msk144_decoder::launch_frontend(Slot& sl, const void* d_in, long long chan_stride, int bpc, long long block0, int nb,
                                    cudaStream_t st) {
    const float* d_fac = nullptr;
    if (cfg.read_mode == 1 && nb > sm_count) {
        RmsParams rp;
        rp.raw = static_cast<const short*>(d_in);
        rp.chan_stride = chan_stride;
        rp.bpc = bpc;
        rp.block0 = block0;
        rp.nblocks = nb;
        rp.fac = sl.d_fac;
        rms_kernel<<<(nb + 31) / 32, 32, 0, st>>>(rp);
        d_fac = sl.d_fac;
        launches += 1;
    }
    if (cfg.read_mode == 1 && cfg.analytic_method == 1) {
        FftParams fp;
        fp.raw = static_cast<const short*>(d_in);
        fp.chan_stride = chan_stride;
        fp.bpc = bpc;
        fp.block0 = block0;
        fp.twiddle = d_twiddle;
        fp.hfilt = d_hfilt;
        fp.out = sl.d_blocks;
        fp.powers = sl.d_powers;
        fp.nblocks = nb;
        fp.fac = d_fac;
        fft_frontend_kernel<<<std::min(nb, sm_count), FFT_THREADS, FFT_SMEM, st>>>(fp);
    } else {
        FrontParams fp;
        fp.raw = d_in;
        fp.chan_stride = chan_stride;
        fp.bpc = bpc;
        fp.read_mode = cfg.read_mode;
        fp.block0 = block0;
        fp.out = sl.d_blocks;
        fp.powers = sl.d_powers;
        fp.fac = d_fac;
        frontend_kernel<<<nb, FRONT_THREADS, FRONT_SMEM, st>>>(fp);
    }
    CK(cudaGetLastError());
    return MSK144_OK;
}

int msk144_decoder::launch_group(Slot& sl, const void* d_in, long long chan_stride, int bpc, long long block0, int nb,
                                 cudaStream_t st) {
    EvQuad* evq = nullptr;
    if (int rc = next_events(&evq)) return rc;
    cudaEvent_t* ev = evq->e;
    evq->nb = nb;
    CK(cudaMemsetAsync(sl.d_counters, 0, 4 * sizeof(int), st));
    CK(cudaEventRecord(ev[0], st));
    if (int rc = launch_frontend(sl, d_in, chan_stride, bpc, block0, nb, st)) return rc;
    CK(cudaGetLastError());
    CK(cudaEventRecord(ev[1], st));
    ScanParams sp;
    sp.blocks = sl.d_blocks;
    sp.mix = d_mix;
    sp.nf = nf;
    sp.depth = depth;
    sp.thr = cfg.nbadsync_threshold;
    sp.fpc = scan_fpc(nb);
    sp.center = cfg.center_frequency_hz;
    sp.if1 = if1;
    sp.step = cfg.search_step_hz;
    sp.block0 = 0;
    sp.cands = sl.d_cands;
    sp.queue = sl.d_queue;
    sp.queue_count = sl.d_counters;
    sp.queue_cap = static_cast<int>(std::min<size_t>(queue_cap, static_cast<size_t>(nb) * items));
    sp.share_twins = share_twins;
    dim3 sgrid((nf + sp.fpc - 1) / sp.fpc, nb);
    launch_scan(depth, sgrid, st, sp);
    CK(cudaGetLastError());
    CK(cudaEventRecord(ev[2], st));
    LdpcParams lp{};
    lp.llr = reinterpret_cast<const float*>(sl.d_queue) + 8;
    lp.stride = sizeof(QueueRec) / sizeof(float);
    lp.count_ptr = sl.d_counters;
    lp.count = 0;
    lp.cap = sp.queue_cap;
    lp.queue = sl.d_queue;
    lp.results = sl.d_results;
    lp.result_count = sl.d_counters + 1;
    lp.result_cap = static_cast<int>(std::min<size_t>(result_cap, 0x7fffffffull));
    lp.flat_out = nullptr;
    std::memcpy(lp.apmask, apmask, sizeof(lp.apmask));
    fill_ldpc_tables(&lp);
    // MSK144_LDPC_CTAS=1: one persistent LDPC CTA per SM instead of two (experiments with two decoders sharing a GPU)
    static const int ldpc_ctas = [] {
        const char* e = std::getenv("MSK144_LDPC_CTAS");
        return e ? std::max(1, std::min(LDPC_CTAS_PER_SM, std::atoi(e))) : LDPC_CTAS_PER_SM;
    }();
    if (cfg.flags & MSK144_FLAG_WSJTX_BP)
        ldpc_kernel<true><<<sm_count * ldpc_ctas, LDPC_WARPS * 32, 0, st>>>(lp);
    else
        ldpc_kernel<false><<<sm_count * ldpc_ctas, LDPC_WARPS * 32, 0, st>>>(lp);
    CK(cudaGetLastError());
    CK(cudaEventRecord(ev[3], st));
    CK(cudaEventRecord(sl.done, st));
    launches += 3;
    return 0;
}

int msk144_decoder::fetch_group(Slot& sl, int nb, std::vector<RawResult>* out, int64_t gblock0,
                                std::vector<float>* powers) {
    cudaStream_t cs = copy_stream;
    CK(cudaStreamWaitEvent(cs, sl.done, 0));
    CK(cudaMemcpyAsync(sl.h_counters, sl.d_counters, 4 * sizeof(int), cudaMemcpyDeviceToHost, cs));
    CK(cudaMemcpyAsync(sl.h_powers, sl.d_powers, static_cast<size_t>(nb) * 8 * sizeof(float), cudaMemcpyDeviceToHost, cs));
    if (int rc = host_wait(cs)) return rc;
    const int qn = sl.h_counters[0], rn = sl.h_counters[1];
    if (static_cast<size_t>(qn) > queue_cap || static_cast<size_t>(rn) > result_cap)
        return fail(MSK144_EOVERFLOW, "device queue overflow");
    d2h_bytes += 4 * sizeof(int) + static_cast<size_t>(nb) * 8 * sizeof(float);
    times.llr_codewords += qn;
    times.ldpc_successes += rn;
    if (powers) powers->insert(powers->end(), sl.h_powers, sl.h_powers + static_cast<size_t>(nb) * 8);
    if (rn > 0) {
        if (int rc = ensure_host_results(rn)) return rc;
        CK(cudaMemcpyAsync(h_results, sl.d_results, static_cast<size_t>(rn) * sizeof(ResultRec), cudaMemcpyDeviceToHost, cs));
        if (int rc = host_wait(cs)) return rc;
        d2h_bytes += static_cast<size_t>(rn) * sizeof(ResultRec);
        if (out)
            for (int i = 0; i < rn; i++) out->push_back(RawResult{gblock0 + h_results[i].block, h_results[i]});
    }
    return 0;
}

int msk144_decoder::decode_common(const void* d_in, size_t spc, int nchan, cudaStream_t st, int64_t block_index_base,
                                  ChanState* const* carry, std::vector<msk144_line>* lines,
                                  std::vector<msk144_result>* results) {
    times = msk144_stage_times{};
    ev_used = 0;
    const int64_t bpc = spc >= static_cast<size_t>(N) ? static_cast<int64_t>((spc - N) / (N / 2)) + 1 : 0;
    const int64_t total = bpc * nchan;
    if (total == 0) return 0;
    if (bpc > 0x7fffffff) return fail(MSK144_EINVAL, "channel too long for one call");
    const int group = std::max(1, cfg.max_batch_blocks);
    if (int rc = ensure_capacity(static_cast<int>(std::min<int64_t>(group, total)))) return rc;
    // a live receiver completes one block per 216 ms hop: its waits sleep; a file or a fast pipe completes a launch
    // group per call, where a sleeping wait would cost throughput (measured: 6 % of the CLI's steady-state rate)
    sleep_wait = (cfg.flags & MSK144_FLAG_BLOCKING_SYNC) != 0 && total <= 8;
    std::vector<RawResult> raw;
    std::vector<float> powers;
    const bool compat = (cfg.flags & MSK144_FLAG_COMPAT_DECODE_CACHE) != 0;
    ChanState local;
    // host post-processing of one fetched group, block by block in stream order (main.cu:388-419); it runs
    // while the GPU works on the next group
    auto process = [&](int64_t g0, int nb) {
        std::sort(raw.begin(), raw.end(), [](const RawResult& a, const RawResult& b) {
            if (a.gblock != b.gblock) return a.gblock < b.gblock;
            return a.r.item < b.r.item;
        });
        size_t ri = 0;
        for (int64_t gb = g0; gb < g0 + nb; gb++) {
            const int ch = static_cast<int>(gb / bpc);
            const int64_t k = gb % bpc;
            if (k == 0) local.reset();  // every channel starts a fresh SNR recurrence and fresh call-sign tables ...
            ChanState* cs = carry ? carry[ch] : &local;  // ... unless the caller carries them per channel
            mskhost::SnrTracker* snr = &cs->snr;
            snr->process_powers(&powers[static_cast<size_t>(gb - g0) * 8]);
            std::vector<mskhost::Present> present;
            while (ri < raw.size() && raw[ri].gblock == gb) {
                const ResultRec& r = raw[ri].r;
                mskhost::Present p;
                p.item = r.item;
                p.f0 = r.f0;
                p.pattern_idx = static_cast<int>((r.item / NCAND) % depth);
                static const int navg[8] = {1, 2, 3, 4, 5, 6, 2, 3};
                p.num_avg = navg[p.pattern_idx];
                p.nbadsync = r.nbadsync;
                std::memset(p.message, 0, sizeof(p.message));
                for (int b = 0; b < 77; b++) p.message[b] = (r.bits[b >> 5] >> (b & 31)) & 1u;
                present.push_back(p);
                if (results) {
                    msk144_result o;
                    std::memset(&o, 0, sizeof(o));
                    o.block_index = block_index_base + k;
                    o.channel = ch;
                    o.item_index = r.item;
                    o.freq_idx = r.item / (NCAND * depth);
                    o.pattern_idx = p.pattern_idx;
                    o.slot = r.item % NCAND;
                    o.pos = r.pos;
                    o.f0 = r.f0;
                    o.xb = r.xb;
                    o.num_avg = p.num_avg;
                    o.nbadsync = r.nbadsync;
                    o.ldpc_iters = r.iters;
                    o.ldpc_nhard = r.nhard;
                    std::memcpy(o.message, p.message, 80);
                    results->push_back(o);
                }
                ri++;
            }
            if (lines && !present.empty()) {
                append_block_lines(mskhost::block_lines(present, compat, snr->snr_i(), cs->hash), block_index_base + k, ch, lines);
            }
        }
        if (powers_sink) powers_sink->insert(powers_sink->end(), powers.begin(), powers.begin() + static_cast<size_t>(nb) * 8);
        raw.clear();
        powers.clear();
    };
    // two-slot software pipeline: the kernels of group g+1 are enqueued before the host fetches group g
    int64_t prev_g0 = -1;
    int prev_nb = 0, cur = 0;
    for (int64_t g0 = 0; g0 < total; g0 += group) {
        const int nb = static_cast<int>(std::min<int64_t>(group, total - g0));
        {
            const int64_t gl = g0 + nb - 1;  // last block of the group: channel gl / bpc, block gl % bpc of it
            const size_t end_sample = static_cast<size_t>(gl / bpc) * spc + static_cast<size_t>(gl % bpc) * (N / 2) + N;
            if (int rc = wait_for_input(end_sample * 2, st)) return rc;
        }
        if (int rc = launch_group(slot[cur], d_in, static_cast<long long>(spc), static_cast<int>(bpc), g0, nb, st)) return rc;
        if (prev_g0 >= 0) {
            if (int rc = fetch_group(slot[cur ^ 1], prev_nb, &raw, prev_g0, &powers)) return rc;
            process(prev_g0, prev_nb);
        }
        prev_g0 = g0, prev_nb = nb, cur ^= 1;
    }
    if (int rc = fetch_group(slot[cur ^ 1], prev_nb, &raw, prev_g0, &powers)) return rc;
    process(prev_g0, prev_nb);
    if (int rc = host_wait(st)) return rc;
    collect_times();
    times.kernel_launches = launches;
    return 0;
}

// ========================================================================================= C ABI
extern "C" {

const char* msk144_last_error(void) { return g_err.c_str(); }
int msk144_fail_(int code, const char* what) { return fail(code, what ? what : ""); }  // for the other translation units

void msk144_config_defaults(msk144_config* cfg) {
    if (!cfg) return;
    std::memset(cfg, 0, sizeof(*cfg));
    cfg->center_frequency_hz = 1500.0f;  // main.cu:125
    cfg->search_step_hz = 2.0f;          // main.cu:129
    cfg->search_width_hz = 200.0f;       // main.cu:130
    cfg->scan_depth = 4;                 // main.cu:131
    cfg->read_mode = 1;
    cfg->analytic_method = 2;            // main.cu:132
    cfg->nbadsync_threshold = 1;         // main.cu:133
    cfg->device = 0;
    cfg->max_batch_blocks = 128;
    cfg->flags = 0;
}

int msk144_create(const msk144_config* cfg, msk144_decoder** out) {
    if (!cfg || !out) return fail(MSK144_EINVAL, "null argument");
    *out = nullptr;
    if (!(cfg->search_step_hz > 0.0f)) return fail(MSK144_EINVAL, "search step must be > 0");  // msk_context.cuh:97
    if (cfg->read_mode != 1 && cfg->read_mode != 2) return fail(MSK144_EINVAL, "read mode must be 1 or 2");
    if (cfg->read_mode == 1 && cfg->analytic_method != 1 && cfg->analytic_method != 2)
        return fail(MSK144_EINVAL, "analytic method must be 1 or 2");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(MSK144_ENODEV, "no CUDA device: this library has no CPU path");
    if (cfg->device < 0 || cfg->device >= ndev) return fail(MSK144_EINVAL, "bad device ordinal");
    CK(cudaSetDevice(cfg->device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, cfg->device));
    if (prop.major < 10)
        return fail(MSK144_ENODEV, std::string("device is sm_") + std::to_string(prop.major * 10 + prop.minor) +
                                       ": kernels are built for sm_100a only");
    msk144_decoder* d = new msk144_decoder();
    d->cfg = *cfg;
    if (d->cfg.max_batch_blocks <= 0) d->cfg.max_batch_blocks = 128;
    d->depth = std::max(1, std::min(8, cfg->scan_depth));  // msk_context.cuh:28-33
    const Grid g = make_grid(cfg->search_width_hz, cfg->search_step_hz);
    d->nf = g.nf;
    d->if1 = g.if1;
    d->items = d->nf * d->depth * NCAND;
    d->sm_count = prop.multiProcessorCount;
    auto bail = [&](int rc) {
        delete d;
        return rc;
    };
#define CKD(call)                                                                                          \
    do {                                                                                                   \
        cudaError_t e_ = (call);                                                                           \
        if (e_ != cudaSuccess) return bail(fail(MSK144_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_))); \
    } while (0)
    CKD(cudaStreamCreateWithFlags(&d->stream, cudaStreamNonBlocking));
    DevTables tab;
    build_tables(&tab);
    CKD(cudaMemcpyToSymbol(c_tab, &tab, sizeof(tab)));
    CKD(cudaMalloc(&d->d_edge_tab, sizeof(tab.edge)));
    CKD(cudaMemcpy(d->d_edge_tab, tab.edge, sizeof(tab.edge), cudaMemcpyHostToDevice));
    CKD(cudaFuncSetAttribute(frontend_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(FRONT_SMEM)));
    CKD(cudaFuncSetAttribute(fft_frontend_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(FFT_SMEM)));
    CKD(cudaFuncSetAttribute(scan_entry(d->depth), cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(SCAN_SMEM)));
    CKD(cudaFuncSetAttribute(scan2_entry(d->depth), cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(SCAN2_SMEM)));
    CKD(cudaFuncSetAttribute(scan2_entry(d->depth), cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    {
        // mixer table, once per decoder (scan_kernel.cuh:45-69 evaluates it per block and frequency)
        const size_t bytes = static_cast<size_t>(d->nf) * N * sizeof(float2);
        if (bytes > (64ull << 30)) return bail(fail(MSK144_EINVAL, "frequency grid too large for the mixer table"));
        CKD(cudaMalloc(&d->d_mix, bytes));
        MixParams mp;
        mp.table = d->d_mix;
        mp.nf = d->nf;
        mp.center = cfg->center_frequency_hz;
        mp.if1 = d->if1;
        mp.step = cfg->search_step_hz;
        if (d->nf > 65535) return bail(fail(MSK144_EINVAL, "more than 65535 search frequencies"));
        mix_table_kernel<<<dim3((N + 255) / 256, d->nf), 256, 0, d->stream>>>(mp);
        CKD(cudaGetLastError());
        CKD(cudaStreamSynchronize(d->stream));
        d->launches += 1;
    }
    CKD(cudaStreamCreateWithFlags(&d->copy_stream, cudaStreamNonBlocking));
    if (cfg->flags & MSK144_FLAG_BLOCKING_SYNC) CKD(cudaEventCreateWithFlags(&d->wait_ev, cudaEventBlockingSync | cudaEventDisableTiming));
    for (auto& sl : d->slot) {
        CKD(cudaMalloc(&sl.d_counters, 4 * sizeof(int)));
        CKD(cudaMallocHost(&sl.h_counters, 4 * sizeof(int)));
        CKD(cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
    }
    if (cfg->read_mode == 1 && cfg->analytic_method == 1) {
        // analytic_fft.cu:32-58: raised-cosine band-pass, 1500 +- 1000 Hz, beta 0.1; DC halved, upper half zero
        std::vector<float2> tw(NFFT / 2);
        for (int k = 0; k < NFFT / 2; k++) {
            const double a = -2.0 * M_PI * k / NFFT;
            tw[k] = make_float2(static_cast<float>(std::cos(a)), static_cast<float>(std::sin(a)));
        }
        std::vector<float> h(NFFT, 0.0f);
        const float df = 12000.0f / NFFT;
        const float pi = 3.141592654f;
        const float t = 1.0f / 2000.0f, beta = 0.1f;
        for (int i = 0; i < NFFT / 2; i++) {
            const float ff = i * df;
            const float f = ff - 1500.0f;
            h[i] = 1.0f;
            if (fabsf(f) > (1 - beta) / (2 * t) && fabsf(f) <= (1 + beta) / (2 * t))
                h[i] = h[i] * 0.5f * (1.0f + cosf((pi * t / beta) * (fabsf(f) - (1 - beta) / (2 * t))));
            else if (fabsf(f) > (1 + beta) / (2 * t))
                h[i] = 0;
        }
        h[0] = 0.5f * h[0];
        CKD(cudaMalloc(&d->d_twiddle, tw.size() * sizeof(float2)));
        CKD(cudaMalloc(&d->d_hfilt, h.size() * sizeof(float)));
        CKD(cudaMemcpy(d->d_twiddle, tw.data(), tw.size() * sizeof(float2), cudaMemcpyHostToDevice));
        CKD(cudaMemcpy(d->d_hfilt, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
#undef CKD
    *out = d;
    return MSK144_OK;
}

void msk144_destroy(msk144_decoder* dec) { delete dec; }

int msk144_get_info(const msk144_decoder* dec, msk144_info* info) {
    if (!dec || !info) return fail(MSK144_EINVAL, "null argument");
    std::memset(info, 0, sizeof(*info));
    info->num_freqs = dec->nf;
    info->scan_depth = dec->depth;
    info->items_per_block = dec->items;
    info->sm_count = dec->sm_count;
    info->left_hz = dec->cfg.center_frequency_hz + dec->if1;  // msk_context.cuh:127-129
    info->right_hz = dec->cfg.center_frequency_hz + dec->if1 + (dec->nf - 1) * dec->cfg.search_step_hz;
    info->max_batch_blocks = dec->cfg.max_batch_blocks;
    return MSK144_OK;
}

int msk144_get_stage_times(msk144_decoder* dec, msk144_stage_times* t) {
    if (!dec || !t) return fail(MSK144_EINVAL, "null argument");
    if (dec->ev_used > 0) {  // groups launched by msk144_run_device and not yet timed
        CK(cudaSetDevice(dec->cfg.device));
        CK(cudaDeviceSynchronize());
        dec->collect_times();
    }
    *t = dec->times;
    t->kernel_launches = dec->launches;
    t->h2d_bytes = dec->h2d_bytes;
    t->d2h_bytes = dec->d2h_bytes;
    dec->times = msk144_stage_times{};
    return MSK144_OK;
}

static void copy_out(const std::vector<msk144_line>& lv, const std::vector<msk144_result>& rv, msk144_line* lines,
                     size_t line_cap, size_t* nlines, msk144_result* results, size_t result_cap, size_t* nresults,
                     int* rc) {
    if (nlines) *nlines = lv.size();
    if (nresults) *nresults = rv.size();
    if (lines) std::memcpy(lines, lv.data(), std::min(line_cap, lv.size()) * sizeof(msk144_line));
    if (results) std::memcpy(results, rv.data(), std::min(result_cap, rv.size()) * sizeof(msk144_result));
    if ((lines && lv.size() > line_cap) || (results && rv.size() > result_cap))
        *rc = fail(MSK144_EOVERFLOW, "output array too small");
}

int msk144_decode_device(msk144_decoder* dec, const void* d_samples, size_t spc, int nchan, void* stream,
                         msk144_line* lines, size_t line_cap, size_t* nlines, msk144_result* results,
                         size_t result_cap, size_t* nresults) {
    if (!dec || !d_samples || nchan < 1) return fail(MSK144_EINVAL, "bad argument");
    CK(cudaSetDevice(dec->cfg.device));
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : dec->stream;
    std::vector<msk144_line> lv;
    std::vector<msk144_result> rv;
    if (int rc = dec->decode_common(d_samples, spc, nchan, st, 0, nullptr, &lv, results || nresults ? &rv : nullptr))
        return rc;
    int rc = MSK144_OK;
    copy_out(lv, rv, lines, line_cap, nlines, results, result_cap, nresults, &rc);
    dec->last_lines = std::move(lv);
    dec->last_results = std::move(rv);
    return rc;
}

int msk144_last_output(msk144_decoder* dec, msk144_line* lines, size_t line_cap, size_t* nlines, msk144_result* results,
                       size_t result_cap, size_t* nresults) {
    if (!dec) return fail(MSK144_EINVAL, "null argument");
    int rc = MSK144_OK;
    copy_out(dec->last_lines, dec->last_results, lines, line_cap, nlines, results, result_cap, nresults, &rc);
    return rc;
}

int msk144_decode_host(msk144_decoder* dec, const void* samples, size_t spc, int nchan, msk144_line* lines,
                       size_t line_cap, size_t* nlines, msk144_result* results, size_t result_cap, size_t* nresults) {
    if (!dec || !samples || nchan < 1) return fail(MSK144_EINVAL, "bad argument");
    CK(cudaSetDevice(dec->cfg.device));
    const size_t bytes = spc * static_cast<size_t>(nchan) * 2;  // s16 mono or s8 I + s8 Q: 2 bytes per sample
    if (int rc = dec->ensure_raw(bytes)) return rc;
    // one exit path: a call without blocks, or an error (in the upload itself included), may leave pieces in flight;
    // they must not outlive the caller's buffer, and a stale piece count must not make the next call wait on events
    // of this one
    int rc = dec->upload_pieces(samples, bytes);
    std::string upload_err;
    if (rc != MSK144_OK)
        upload_err = g_err;
    else
        rc = msk144_decode_device(dec, dec->d_raw, spc, nchan, dec->stream, lines, line_cap, nlines, results, result_cap,
                                  nresults);
    dec->up_pieces = 0;
    if (dec->up_stream) cudaStreamSynchronize(dec->up_stream);
    if (!upload_err.empty()) return fail(rc, upload_err);
    return rc;
}

// ---- one channel split in time (SURVEY.md section 8e "Partitioning"): GPU stages per piece, one ordered host pass ----
int msk144_decode_chunk_host(msk144_decoder* dec, const void* samples, size_t nsamples, int64_t block_base,
                             msk144_result* results, size_t result_cap, size_t* nresults, float* powers8,
                             size_t power_cap_blocks, int64_t* nblocks_out) {
    if (!dec || !samples) return fail(MSK144_EINVAL, "bad argument");
    const int64_t nb = nsamples >= static_cast<size_t>(N) ? static_cast<int64_t>((nsamples - N) / (N / 2)) + 1 : 0;
    if (nblocks_out) *nblocks_out = nb;
    if (nresults) *nresults = 0;
    if (powers8 && power_cap_blocks < static_cast<size_t>(nb)) return fail(MSK144_EINVAL, "powers8 holds fewer blocks than the piece has");
    CK(cudaSetDevice(dec->cfg.device));
    const size_t bytes = nsamples * 2;
    if (int rc = dec->ensure_raw(bytes)) return rc;
    std::vector<msk144_result> rv;
    std::vector<float> pw;
    int rc = dec->upload_pieces(samples, bytes);
    std::string err;
    if (rc == MSK144_OK) {
        dec->powers_sink = &pw;
        // lines == nullptr: no SNR index is used and no text is decoded here; the caller replays both in stream order
        rc = dec->decode_common(dec->d_raw, nsamples, 1, dec->stream, block_base, nullptr, nullptr, &rv);
        dec->powers_sink = nullptr;
    }
    if (rc != MSK144_OK) err = g_err;
    dec->up_pieces = 0;
    if (dec->up_stream) cudaStreamSynchronize(dec->up_stream);
    if (rc != MSK144_OK) return fail(rc, err);
    if (powers8) std::memcpy(powers8, pw.data(), pw.size() * sizeof(float));
    copy_out({}, rv, nullptr, 0, nullptr, results, result_cap, nresults, &rc);
    dec->last_lines.clear();
    dec->last_results = std::move(rv);  // after MSK144_EOVERFLOW: msk144_last_output with an array of *nresults
    return rc;
}

int msk144_replay_channel(int32_t flags, int32_t channel, int64_t first_block, int64_t nblocks, const float* powers8,
                          const msk144_result* results, size_t nresults, msk144_line* lines, size_t line_cap,
                          size_t* nlines) {
    if (nblocks < 0 || (nblocks && !powers8) || (nresults && !results)) return fail(MSK144_EINVAL, "bad argument");
    const bool compat = (flags & MSK144_FLAG_COMPAT_DECODE_CACHE) != 0;
    ChanState cs;  // a fresh SNR recurrence and fresh call-sign tables, as at the start of a channel
    std::vector<msk144_line> lv;
    size_t ri = 0;
    for (int64_t k = 0; k < nblocks; k++) {
        const int64_t gb = first_block + k;
        cs.snr.process_powers(powers8 + static_cast<size_t>(k) * 8);
        if (ri < nresults && results[ri].block_index < gb) return fail(MSK144_EINVAL, "results are not in block order");
        std::vector<mskhost::Present> present;
        while (ri < nresults && results[ri].block_index == gb) {
            const msk144_result& r = results[ri++];
            if (!present.empty() && r.item_index <= present.back().item) return fail(MSK144_EINVAL, "results of a block are not in item order");
            mskhost::Present p;
            p.item = r.item_index;
            p.f0 = r.f0;
            p.pattern_idx = static_cast<int>(r.pattern_idx);
            p.num_avg = r.num_avg;
            p.nbadsync = r.nbadsync;
            std::memcpy(p.message, r.message, sizeof(p.message));
            present.push_back(p);
        }
        if (!present.empty()) append_block_lines(mskhost::block_lines(present, compat, cs.snr.snr_i(), cs.hash), gb, channel, &lv);
    }
    if (ri != nresults) return fail(MSK144_EINVAL, "results beyond the last block");
    int rc = MSK144_OK;
    copy_out(lv, {}, lines, line_cap, nlines, nullptr, 0, nullptr, &rc);
    return rc;
}

int msk144_deinterleave_device(msk144_decoder* dec, const void* d_in, void* d_out, size_t spc, int nchan, void* stream) {
    if (!dec || !d_in || !d_out || nchan < 1 || d_in == d_out) return fail(MSK144_EINVAL, "bad argument");
    if (spc == 0) return MSK144_OK;
    CK(cudaSetDevice(dec->cfg.device));
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : dec->stream;
    const dim3 grid(static_cast<unsigned>((spc + DEINT_TILE - 1) / DEINT_TILE), static_cast<unsigned>((nchan + DEINT_TILE - 1) / DEINT_TILE));
    if (grid.y > 65535u) return fail(MSK144_EINVAL, "too many channels");
    const bool vec = nchan % 4 == 0 && spc % 4 == 0 && reinterpret_cast<uintptr_t>(d_in) % 8 == 0 &&
                     reinterpret_cast<uintptr_t>(d_out) % 8 == 0;
    const auto* in = static_cast<const unsigned short*>(d_in);
    auto* out = static_cast<unsigned short*>(d_out);
    if (vec)
        deinterleave_kernel<true><<<grid, DEINT_THREADS, 0, st>>>(in, out, static_cast<long long>(spc), nchan);
    else
        deinterleave_kernel<false><<<grid, DEINT_THREADS, 0, st>>>(in, out, static_cast<long long>(spc), nchan);
    CK(cudaGetLastError());
    dec->launches += 1;
    return MSK144_OK;
}

int msk144_decode_host_interleaved(msk144_decoder* dec, const void* samples, size_t spc, int nchan, msk144_line* lines,
                                   size_t line_cap, size_t* nlines, msk144_result* results, size_t result_cap,
                                   size_t* nresults) {
    if (!dec || !samples || nchan < 1) return fail(MSK144_EINVAL, "bad argument");
    CK(cudaSetDevice(dec->cfg.device));
    const size_t bytes = spc * static_cast<size_t>(nchan) * 2;
    // one allocation: [channel-major copy | interleaved input]; the decode reads the first half
    if (int rc = dec->ensure_raw(2 * bytes + 16)) return rc;
    char* d_cm = static_cast<char*>(dec->d_raw);
    char* d_il = d_cm + ((bytes + 15) / 16) * 16;
    CK(cudaMemcpyAsync(d_il, samples, bytes, cudaMemcpyHostToDevice, dec->stream));
    dec->h2d_bytes += bytes;
    if (int rc = msk144_deinterleave_device(dec, d_il, d_cm, spc, nchan, dec->stream)) return rc;
    return msk144_decode_device(dec, d_cm, spc, nchan, dec->stream, lines, line_cap, nlines, results, result_cap, nresults);
}

int msk144_run_device(msk144_decoder* dec, const void* d_samples, size_t spc, int nchan, void* stream,
                      int64_t* nblocks_out) {
    if (!dec || !d_samples || nchan < 1) return fail(MSK144_EINVAL, "bad argument");
    CK(cudaSetDevice(dec->cfg.device));
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : dec->stream;
    const int64_t bpc = spc >= static_cast<size_t>(N) ? static_cast<int64_t>((spc - N) / (N / 2)) + 1 : 0;
    const int64_t total = bpc * nchan;
    if (nblocks_out) *nblocks_out = total;
    if (total == 0) return MSK144_OK;
    const int group = std::max(1, dec->cfg.max_batch_blocks);
    if (int rc = dec->ensure_capacity(static_cast<int>(std::min<int64_t>(group, total)))) return rc;
    for (int64_t g0 = 0; g0 < total; g0 += group) {
        const int nb = static_cast<int>(std::min<int64_t>(group, total - g0));
        // alternate the two buffer slots like the decode path (same stream: groups still run in order)
        if (int rc = dec->launch_group(dec->slot[(g0 / group) & 1], d_samples, static_cast<long long>(spc), static_cast<int>(bpc), g0, nb, st)) return rc;
    }
    return MSK144_OK;
}

// ---- streaming -----------------------------------------------------------------------------------
int msk144_reset_stream(msk144_decoder* dec) {
    if (!dec) return fail(MSK144_EINVAL, "null argument");
    dec->pending.clear();
    dec->il_pending.clear(), dec->il_state.clear(), dec->il_nchan = 0, dec->il_next_block = 0;
    dec->wav_checked = false;
    dec->next_block = 0;
    dec->stream_state.reset();
    dec->out_lines.clear();
    return MSK144_OK;
}

int msk144_push(msk144_decoder* dec, const void* bytes, size_t nbytes) {
    if (!dec || (!bytes && nbytes)) return fail(MSK144_EINVAL, "bad argument");
    CK(cudaSetDevice(dec->cfg.device));
    const uint8_t* b = static_cast<const uint8_t*>(bytes);
    if (int rc = append_pinned(&dec->pending, b, nbytes)) return rc;
    if ((dec->cfg.flags & MSK144_FLAG_SKIP_WAV_HEADER) && !dec->wav_checked) {
        // start of the stream: recognise a RIFF/WAVE header and drop it (adapters.cpp)
        const int64_t off = msk144_wav_data_offset(dec->pending.data(), dec->pending.size(), nullptr, nullptr, nullptr);
        if (off == MSK144_EAGAIN) return MSK144_OK;  // header not complete yet
        if (off == MSK144_EINVAL) return fail(MSK144_EINVAL, "RIFF input is not PCM");
        dec->wav_checked = true;
        if (off > 0) dec->pending.erase(dec->pending.begin(), dec->pending.begin() + off);
    }
    const size_t nsamp = dec->pending.size() / 2;
    if (nsamp < static_cast<size_t>(N)) return MSK144_OK;
    const size_t nblocks = (nsamp - N) / (N / 2) + 1;
    const size_t used_samples = (nblocks - 1) * (N / 2) + N;
    if (int rc = dec->ensure_raw(used_samples * 2)) return rc;
    CK(cudaMemcpyAsync(dec->d_raw, dec->pending.data(), used_samples * 2, cudaMemcpyHostToDevice, dec->stream));
    dec->h2d_bytes += used_samples * 2;
    std::vector<msk144_line> lv;
    ChanState* const carry1[1] = {&dec->stream_state};
    if (int rc = dec->decode_common(dec->d_raw, used_samples, 1, dec->stream, dec->next_block, carry1, &lv, nullptr))
        return rc;
    for (const auto& l : lv) dec->out_lines.push_back(l);
    dec->next_block += static_cast<int64_t>(nblocks);
    // keep the second half of the last block: it is the first half of the next one (main.cu:284-288)
    dec->pending.erase(dec->pending.begin(), dec->pending.begin() + nblocks * (N / 2) * 2);
    return MSK144_OK;
}

int msk144_push_interleaved(msk144_decoder* dec, const void* bytes, size_t nbytes, int nchan) {
    if (!dec || (!bytes && nbytes) || nchan < 1) return fail(MSK144_EINVAL, "bad argument");
    if (dec->il_nchan == 0) {
        dec->il_nchan = nchan;
        dec->il_state.clear();
        for (int c = 0; c < nchan; c++) dec->il_state.emplace_back(new ChanState());
    } else if (dec->il_nchan != nchan) {
        return fail(MSK144_EINVAL, "channel count changed in mid-stream (msk144_reset_stream first)");
    }
    CK(cudaSetDevice(dec->cfg.device));
    const uint8_t* b = static_cast<const uint8_t*>(bytes);
    if (int rc = append_pinned(&dec->il_pending, b, nbytes)) return rc;
    const size_t frame = 2 * static_cast<size_t>(nchan);  // one time step of every channel
    const size_t nsteps = dec->il_pending.size() / frame;
    if (nsteps < static_cast<size_t>(N)) return MSK144_OK;
    const size_t nblocks = (nsteps - N) / (N / 2) + 1;
    const size_t used = (nblocks - 1) * (N / 2) + N;
    const size_t ubytes = used * frame;
    if (int rc = dec->ensure_raw(2 * ubytes + 16)) return rc;
    char* d_cm = static_cast<char*>(dec->d_raw);
    char* d_il = d_cm + ((ubytes + 15) / 16) * 16;
    CK(cudaMemcpyAsync(d_il, dec->il_pending.data(), ubytes, cudaMemcpyHostToDevice, dec->stream));
    dec->h2d_bytes += ubytes;
    if (int rc = msk144_deinterleave_device(dec, d_il, d_cm, used, nchan, dec->stream)) return rc;
    std::vector<msk144_line> lv;
    std::vector<ChanState*> carry(static_cast<size_t>(nchan));
    for (int c = 0; c < nchan; c++) carry[c] = dec->il_state[c].get();
    if (int rc = dec->decode_common(d_cm, used, nchan, dec->stream, dec->il_next_block, carry.data(), &lv, nullptr))
        return rc;
    // time order (block, then channel), whatever number of blocks this push happened to complete: the lines of a stream
    // do not depend on how the caller chunks its pushes
    std::stable_sort(lv.begin(), lv.end(), [](const msk144_line& a, const msk144_line& b) {
        return a.block_index != b.block_index ? a.block_index < b.block_index : a.channel < b.channel;
    });
    for (const auto& l : lv) dec->out_lines.push_back(l);
    dec->il_next_block += static_cast<int64_t>(nblocks);
    dec->il_pending.erase(dec->il_pending.begin(), dec->il_pending.begin() + nblocks * (N / 2) * frame);
    return MSK144_OK;
}

int msk144_poll(msk144_decoder* dec, msk144_line* lines, size_t cap, size_t* nlines) {
    if (!dec || !nlines) return fail(MSK144_EINVAL, "null argument");
    size_t n = 0;
    while (n < cap && !dec->out_lines.empty()) {
        lines[n++] = dec->out_lines.front();
        dec->out_lines.pop_front();
    }
    *nlines = n;
    return MSK144_OK;
}

// ---- stage level -----------------------------------------------------------------------------------
int msk144_frontend_host(msk144_decoder* dec, const void* samples, size_t nsamples, float* blocks_re_im,
                         float* powers8, int64_t* nblocks_out) {
    if (!dec || !samples) return fail(MSK144_EINVAL, "bad argument");
    CK(cudaSetDevice(dec->cfg.device));
    const int64_t total = nsamples >= static_cast<size_t>(N) ? static_cast<int64_t>((nsamples - N) / (N / 2)) + 1 : 0;
    if (nblocks_out) *nblocks_out = total;
    if (total == 0) return MSK144_OK;
    if (int rc = dec->ensure_raw(nsamples * 2)) return rc;
    CK(cudaMemcpyAsync(dec->d_raw, samples, nsamples * 2, cudaMemcpyHostToDevice, dec->stream));
    const int group = std::max(1, dec->cfg.max_batch_blocks);
    if (int rc = dec->ensure_capacity(static_cast<int>(std::min<int64_t>(group, total)))) return rc;
    // run only the front-end kernel of each group
    for (int64_t g0 = 0; g0 < total; g0 += group) {
        const int nb = static_cast<int>(std::min<int64_t>(group, total - g0));
        if (int rc = dec->launch_frontend(dec->slot[0], dec->d_raw, static_cast<long long>(nsamples), static_cast<int>(total), g0, nb,
                                          dec->stream))
            return rc;
        CK(cudaGetLastError());
        dec->launches += 1;
        if (blocks_re_im)
            CK(cudaMemcpyAsync(blocks_re_im + static_cast<size_t>(g0) * N * 2, dec->slot[0].d_blocks,
                               static_cast<size_t>(nb) * N * sizeof(float2), cudaMemcpyDeviceToHost, dec->stream));
        if (powers8)
            CK(cudaMemcpyAsync(powers8 + static_cast<size_t>(g0) * 8, dec->slot[0].d_powers, static_cast<size_t>(nb) * 8 * sizeof(float),
                               cudaMemcpyDeviceToHost, dec->stream));
        CK(cudaStreamSynchronize(dec->stream));
    }
    return MSK144_OK;
}

int msk144_scan_host(msk144_decoder* dec, const float* blocks_re_im, int64_t nblocks, msk144_candidate* cands,
                     float* llr, uint32_t* llr_item, size_t llr_cap, size_t* nllr) {
    if (!dec || !blocks_re_im || nblocks < 1) return fail(MSK144_EINVAL, "bad argument");
    CK(cudaSetDevice(dec->cfg.device));
    const int group = std::max(1, dec->cfg.max_batch_blocks);
    if (int rc = dec->ensure_capacity(static_cast<int>(std::min<int64_t>(group, nblocks)))) return rc;
    size_t total_llr = 0;
    int rc_out = MSK144_OK;
    std::vector<QueueRec> hq;
    for (int64_t g0 = 0; g0 < nblocks; g0 += group) {
        const int nb = static_cast<int>(std::min<int64_t>(group, nblocks - g0));
        CK(cudaMemcpyAsync(dec->slot[0].d_blocks, blocks_re_im + static_cast<size_t>(g0) * N * 2,
                           static_cast<size_t>(nb) * N * sizeof(float2), cudaMemcpyHostToDevice, dec->stream));
        CK(cudaMemsetAsync(dec->slot[0].d_counters, 0, 4 * sizeof(int), dec->stream));
        ScanParams sp;
        sp.blocks = dec->slot[0].d_blocks;
        sp.mix = dec->d_mix;
        sp.nf = dec->nf;
        sp.depth = dec->depth;
        sp.thr = dec->cfg.nbadsync_threshold;
        sp.fpc = dec->scan_fpc(nb);
        sp.center = dec->cfg.center_frequency_hz;
        sp.if1 = dec->if1;
        sp.step = dec->cfg.search_step_hz;
        sp.block0 = 0;
        sp.cands = dec->slot[0].d_cands;
        sp.queue = dec->slot[0].d_queue;
        sp.queue_count = dec->slot[0].d_counters;
        sp.queue_cap = static_cast<int>(static_cast<size_t>(nb) * dec->items);
        sp.share_twins = 0;  // the stage API returns every gated candidate's own LLR row
        dim3 sgrid((dec->nf + sp.fpc - 1) / sp.fpc, nb);
        launch_scan(dec->depth, sgrid, dec->stream, sp);
        CK(cudaGetLastError());
        dec->launches += 1;
        CK(cudaMemcpyAsync(dec->slot[0].h_counters, dec->slot[0].d_counters, 4 * sizeof(int), cudaMemcpyDeviceToHost, dec->stream));
        if (cands)
            CK(cudaMemcpyAsync(cands + static_cast<size_t>(g0) * dec->items, dec->slot[0].d_cands,
                               static_cast<size_t>(nb) * dec->items * sizeof(CandRec), cudaMemcpyDeviceToHost, dec->stream));
        CK(cudaStreamSynchronize(dec->stream));
        const int qn = dec->slot[0].h_counters[0];
        if (llr || llr_item) {
            hq.resize(qn);
            CK(cudaMemcpy(hq.data(), dec->slot[0].d_queue, static_cast<size_t>(qn) * sizeof(QueueRec), cudaMemcpyDeviceToHost));
            for (int i = 0; i < qn; i++) {
                const size_t row = total_llr + i;
                if (row >= llr_cap) {
                    rc_out = MSK144_EOVERFLOW;
                    continue;
                }
                if (llr) std::memcpy(llr + row * 128, hq[i].llr, 128 * sizeof(float));
                if (llr_item) llr_item[row] = static_cast<uint32_t>((g0 + hq[i].block) * dec->items + hq[i].item);
            }
        }
        if (cands) {  // make llr_slot global across groups
            msk144_candidate* c = cands + static_cast<size_t>(g0) * dec->items;
            for (size_t i = 0; i < static_cast<size_t>(nb) * dec->items; i++)
                if (c[i].llr_slot >= 0) c[i].llr_slot += static_cast<int32_t>(total_llr);
        }
        total_llr += qn;
    }
    if (nllr) *nllr = total_llr;
    if (rc_out != MSK144_OK) return fail(rc_out, "llr array too small");
    return MSK144_OK;
}

int msk144_ldpc_decode_device(msk144_decoder* dec, const float* d_llr, size_t ncw, uint32_t* d_out, void* stream) {
    if (!dec || !d_llr || !d_out) return fail(MSK144_EINVAL, "bad argument");
    if (ncw > 0x7fffffffull) return fail(MSK144_EINVAL, "too many codewords for one call");
    CK(cudaSetDevice(dec->cfg.device));
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : dec->stream;
    LdpcParams lp{};
    lp.llr = d_llr;
    lp.stride = 128;
    lp.count_ptr = nullptr;
    lp.count = static_cast<int>(ncw);
    lp.cap = static_cast<int>(ncw);
    lp.flat_out = reinterpret_cast<uint4*>(d_out);
    dec->fill_ldpc_tables(&lp);
    const size_t npairs = (ncw + 1) / 2;
    const int grid = static_cast<int>(std::min<size_t>(static_cast<size_t>(dec->sm_count) * LDPC_CTAS_PER_SM, (npairs + LDPC_WARPS - 1) / LDPC_WARPS));
    if (grid > 0) {
        // caller-supplied LLRs: iteration counts differ from codeword to codeword, so the streaming form (each half
        // of a warp refills on its own) is used; MSK144_LDPC_NO_REFILL=1 selects the pair form for A/B runs
        static const bool no_refill = std::getenv("MSK144_LDPC_NO_REFILL") != nullptr;
        if (dec->cfg.flags & MSK144_FLAG_WSJTX_BP) {
            std::memcpy(lp.apmask, dec->apmask, sizeof(lp.apmask));
            ldpc_kernel<true><<<grid, LDPC_WARPS * 32, 0, st>>>(lp);  // the pair form carries the WSJT-X control flow
        } else if (no_refill)
            ldpc_kernel<false><<<grid, LDPC_WARPS * 32, 0, st>>>(lp);
        else
            ldpc_stream_kernel<<<grid, LDPC_WARPS * 32, 0, st>>>(lp);
        CK(cudaGetLastError());
        dec->launches += 1;
    }
    return MSK144_OK;
}

int msk144_ldpc_decode_host(msk144_decoder* dec, const float* llr, size_t ncw, uint8_t* found, uint8_t* msg77,
                            int32_t* iters, int32_t* nhard) {
    if (!dec || !llr) return fail(MSK144_EINVAL, "bad argument");
    if (ncw == 0) return MSK144_OK;
    CK(cudaSetDevice(dec->cfg.device));
    float* d_llr = nullptr;
    uint32_t* d_out = nullptr;
    CK(cudaMalloc(&d_llr, ncw * 128 * sizeof(float)));
    if (cudaMalloc(&d_out, ncw * 16) != cudaSuccess) {
        cudaFree(d_llr);
        return fail(MSK144_ENOMEM, "cudaMalloc");
    }
    int rc = MSK144_OK;
    std::vector<uint32_t> out(ncw * 4);
    do {
        if (cudaMemcpyAsync(d_llr, llr, ncw * 128 * sizeof(float), cudaMemcpyHostToDevice, dec->stream) != cudaSuccess) {
            rc = fail(MSK144_ECUDA, "H2D");
            break;
        }
        rc = msk144_ldpc_decode_device(dec, d_llr, ncw, d_out, dec->stream);
        if (rc) break;
        if (cudaMemcpyAsync(out.data(), d_out, ncw * 16, cudaMemcpyDeviceToHost, dec->stream) != cudaSuccess ||
            cudaStreamSynchronize(dec->stream) != cudaSuccess) {
            rc = fail(MSK144_ECUDA, std::string("ldpc: ") + cudaGetErrorString(cudaGetLastError()));
            break;
        }
    } while (0);
    cudaFree(d_llr);
    cudaFree(d_out);
    if (rc) return rc;
    for (size_t i = 0; i < ncw; i++) {
        const uint32_t* w = &out[i * 4];
        const bool f = (w[2] >> 31) != 0;
        if (found) found[i] = f ? 1 : 0;
        if (msg77)
            for (int b = 0; b < 77; b++) msg77[i * 77 + b] = (w[b >> 5] >> (b & 31)) & 1u;
        if (iters) iters[i] = static_cast<int32_t>(w[3] & 0xff);
        if (nhard) {
            const int nh = static_cast<int>((w[3] >> 8) & 0xff);
            nhard[i] = (nh == 0xff) ? -1 : nh;
        }
    }
    return MSK144_OK;
}

int msk144_set_ap_mask(msk144_decoder* dec, const uint8_t* apmask128) {
    if (!dec) return fail(MSK144_EINVAL, "null argument");
    if (!(dec->cfg.flags & MSK144_FLAG_WSJTX_BP)) return fail(MSK144_EINVAL, "a-priori masks belong to MSK144_FLAG_WSJTX_BP");
    for (int i = 0; i < 4; i++) dec->apmask[i] = 0u;
    if (apmask128)
        for (int b = 0; b < 128; b++)
            if (apmask128[b]) dec->apmask[b >> 5] |= 1u << (b & 31);
    return MSK144_OK;
}

int msk144_selftest_ldpc_math(msk144_decoder* dec, uint64_t* tanh_mismatches, uint64_t* platanh_mismatches) {
    if (!dec) return fail(MSK144_EINVAL, "null argument");
    CK(cudaSetDevice(dec->cfg.device));
    unsigned long long* d_m = nullptr;
    CK(cudaMalloc(&d_m, 2 * sizeof(unsigned long long)));
    cudaMemsetAsync(d_m, 0, 2 * sizeof(unsigned long long), dec->stream);
    LdpcParams lp{};
    dec->fill_ldpc_tables(&lp);
    ldpc_math_selftest_kernel<<<dec->sm_count * 8, 256, 0, dec->stream>>>(lp, d_m);
    unsigned long long h[2] = {~0ull, ~0ull};
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(h, d_m, sizeof(h), cudaMemcpyDeviceToHost, dec->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(dec->stream);
    cudaFree(d_m);
    if (e != cudaSuccess) return fail(MSK144_ECUDA, std::string("selftest: ") + cudaGetErrorString(e));
    dec->launches += 1;
    if (tanh_mismatches) *tanh_mismatches = h[0];
    if (platanh_mismatches) *platanh_mismatches = h[1];
    return MSK144_OK;
}

int msk144_measure_fp32_peak(msk144_decoder* dec, double* tflops) {
    if (!dec || !tflops) return fail(MSK144_EINVAL, "null argument");
    CK(cudaSetDevice(dec->cfg.device));
    float* d_sink = nullptr;
    CK(cudaMalloc(&d_sink, sizeof(float)));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const int grid = dec->sm_count * 4, threads = 512, iters = 8192;  // 2048 threads = 16 warps per scheduler
    double best = 0.0;
    for (int rep = 0; rep < 4; rep++) {  // first pass warms up; best of the rest
        cudaEventRecord(e0, dec->stream);
        fp32_peak_kernel<<<grid, threads, 0, dec->stream>>>(d_sink, iters, 0.999f, 0.001f);
        cudaEventRecord(e1, dec->stream);
        if (cudaStreamSynchronize(dec->stream) != cudaSuccess) break;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double fl = 2.0 * 32.0 * iters * static_cast<double>(grid) * threads;
        if (rep > 0 && ms > 0.f) best = std::max(best, fl / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0), cudaEventDestroy(e1);
    cudaFree(d_sink);
    dec->launches += 4;
    if (cudaGetLastError() != cudaSuccess || best <= 0.0) return fail(MSK144_ECUDA, "fp32 peak probe failed");
    *tflops = best;
    return MSK144_OK;
}

const char* msk144_build_info(void) {
#define MSK_STR2(x) #x
#define MSK_STR(x) MSK_STR2(x)
    return "nvcc " MSK_STR(__CUDACC_VER_MAJOR__) "." MSK_STR(__CUDACC_VER_MINOR__) "." MSK_STR(__CUDACC_VER_BUILD__)
           "; sm_100a; " __DATE__;
}

#ifdef MSK_SCAN_PROF
// debug build only: reads and clears the per-phase cycle counters of scan_kernel
int msk144_debug_scan_prof(unsigned long long* out8) {
    unsigned long long z[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out8, g_scan_prof, sizeof(z));
    cudaMemcpyToSymbol(g_scan_prof, z, sizeof(z));
    return 0;
}
#endif

#ifdef MSK_LDPC_PROF
int msk144_debug_ldpc_prof(unsigned long long* out8) {
    unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out8, g_ldpc_prof, sizeof(z));
    cudaMemcpyToSymbol(g_ldpc_prof, z, sizeof(z));
    return 0;
}
#endif

int msk144_decode_message(const uint8_t* msg77, char* text) {
    std::string s;
    const bool ok = mskhost::decode_message(msg77, &s);
    if (text) std::snprintf(text, 38, "%s", s.c_str());
    return ok ? 1 : 0;
}

int msk144_format_line(const msk144_line* line, const char* date14, char* out, size_t cap) {
    if (!line || !out) return fail(MSK144_EINVAL, "null argument");
    const std::string stamp = date14 ? std::string(date14) : mskhost::now_stamp();
    const std::string s = mskhost::format_line(line->snr, line->f0, line->num_avg, line->nbadsync, line->pattern_idx,
                                               stamp.c_str(), line->text);
    std::snprintf(out, cap, "%s", s.c_str());
    return static_cast<int>(s.size());
}

}  // extern "C"

// =====================================================================================================
// msk144_frontend_device_noalloc: the method-2 / IQ front end for callers that must not allocate.
//
// The pipeline's front end keeps 1 / rms of every block in a cudaMalloc'ed array between rms_kernel and
// frontend_kernel (Slot::d_fac).  Here that array is the file-scope device array below, which the CUDA module of
// this library carries on every GPU that loads it: 32 KB per GPU, from the first use of the library on that GPU
// until the process exits.  Nothing in this section calls cudaMalloc / cudaFree / cudaMallocHost or any other
// allocating routine, and the success path does not touch the host heap either.
// =====================================================================================================
static __device__ __align__(256) float g_noalloc_fac[MSK144_NOALLOC_MAX_BLOCKS];

namespace {

// In host code `g_noalloc_fac` is only the host-side handle of the symbol, not a device address: a kernel handed
// that name would read host memory or fault.  The device address differs from GPU to GPU, so it is looked up
// once per device, with that device current, and kept here.
constexpr int NOALLOC_MAX_DEVICES = 64;
struct NoAllocDevice {
    std::atomic<int> ready{0};
    float* fac = nullptr;
};
NoAllocDevice g_noalloc_dev[NOALLOC_MAX_DEVICES];
std::mutex g_noalloc_mutex;

// Set-up for the CURRENT device (ordinal `dev`): device address of the scratch array, and the opt-in for the
// 82 KB of dynamic shared memory frontend_kernel needs (per device, so it cannot lean on msk144_create having
// run on some other GPU).  Neither call allocates.
int noalloc_setup(int dev, float** fac) {
    NoAllocDevice& d = g_noalloc_dev[dev];
    if (d.ready.load(std::memory_order_acquire)) {
        *fac = d.fac;
        return MSK144_OK;
    }
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!d.ready.load(std::memory_order_relaxed)) {
        void* p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_noalloc_fac));
        if (!p) return fail(MSK144_ECUDA, "cudaGetSymbolAddress returned no address for the scratch array");
        CK(cudaFuncSetAttribute(frontend_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(FRONT_SMEM)));
        d.fac = static_cast<float*>(p);
        d.ready.store(1, std::memory_order_release);
    }
    *fac = d.fac;
    return MSK144_OK;
}

}  // namespace

extern "C" int msk144_frontend_device_noalloc(const void* d_samples, size_t nsamples, int read_mode, float* d_blocks_re_im,
                                              float* d_powers8, void* stream, int64_t* nblocks_out) {
    if (nblocks_out) *nblocks_out = 0;
    if (!d_samples || !d_blocks_re_im || !d_powers8) return fail(MSK144_EINVAL, "null device pointer");
    if (read_mode != 1 && read_mode != 2) return fail(MSK144_EINVAL, "read_mode must be 1 (s16 audio) or 2 (s8 I/Q pairs)");
    if ((reinterpret_cast<uintptr_t>(d_blocks_re_im) & 7) || (reinterpret_cast<uintptr_t>(d_powers8) & 3) ||
        (read_mode == 1 && (reinterpret_cast<uintptr_t>(d_samples) & 1)))
        return fail(MSK144_EINVAL, "misaligned device pointer (blocks: 8 bytes, powers: 4, s16 samples: 2)");
    const int64_t total = nsamples >= static_cast<size_t>(N) ? static_cast<int64_t>((nsamples - N) / (N / 2)) + 1 : 0;
    if (total > MSK144_NOALLOC_MAX_BLOCKS) {
        char msg[192];
        std::snprintf(msg, sizeof(msg),
                      "%lld blocks in one call, but the static scratch of msk144_frontend_device_noalloc holds %d: pass at most %lld samples per call",
                      static_cast<long long>(total), MSK144_NOALLOC_MAX_BLOCKS,
                      static_cast<long long>(MSK144_NOALLOC_MAX_BLOCKS + 1) * (N / 2));
        return fail(MSK144_EINVAL, msg);
    }
    if (nblocks_out) *nblocks_out = total;
    if (total == 0) return MSK144_OK;
    int dev = -1;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return fail(MSK144_EINVAL, "device ordinal beyond the per-device table (64)");
    float* fac = nullptr;
    if (int rc = noalloc_setup(dev, &fac)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int nb = static_cast<int>(total);
    if (read_mode == 1) {
        // every entry frontend_kernel reads below is written here first, so what an earlier call left in the array
        // never shows; nothing in it has to start at zero
        RmsParams rp;
        rp.raw = static_cast<const short*>(d_samples);
        rp.chan_stride = static_cast<long long>(nsamples);
        rp.bpc = nb;
        rp.block0 = 0;
        rp.nblocks = nb;
        rp.fac = fac;
        rms_kernel<<<(nb + 31) / 32, 32, 0, st>>>(rp);
        CK(cudaGetLastError());
    }
    FrontParams fp;
    fp.raw = d_samples;
    fp.chan_stride = static_cast<long long>(nsamples);
    fp.bpc = nb;
    fp.read_mode = read_mode;
    fp.block0 = 0;
    fp.out = reinterpret_cast<float2*>(d_blocks_re_im);
    fp.powers = d_powers8;
    fp.fac = read_mode == 1 ? fac : nullptr;  // the IQ path has no rms step (main.cu:365-371) and uses no scratch
    frontend_kernel<<<nb, FRONT_THREADS, FRONT_SMEM, st>>>(fp);
    CK(cudaGetLastError());
    return MSK144_OK;
}
