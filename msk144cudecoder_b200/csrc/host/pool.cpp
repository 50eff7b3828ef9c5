// pool.cpp -- one decoder per GPU behind one handle (include/msk144_b200.h, msk144_pool_*).
//
// The reference is a single-GPU, single-channel process (main.cu:98-115: one cudaSetDevice-less context,
// no threads).  Channels are independent units, so a box with several GPUs is used by
// giving every GPU a contiguous range of channels, one host thread and its own CUDA streams, and merging the
// decode lines on the host: no data crosses GPUs.  Everything here goes through the single-GPU C ABI; the
// output is, line for line and in order, what ONE decoder returns for the same input.
#include <algorithm>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../../include/msk144_b200.h"

namespace {
thread_local std::string g_pool_err;
int pfail(int code, const std::string& what) {
    g_pool_err = what;
    return code;
}
// contiguous, balanced: the first nchan % n members get one extra channel (== sharding.channels_of_rank)
void split(int nchan, int n, std::vector<int>* first, std::vector<int>* count) {
    first->assign(n, 0), count->assign(n, 0);
    const int base = nchan / n, extra = nchan % n;
    int start = 0;
    for (int i = 0; i < n; i++) {
        (*first)[i] = start;
        (*count)[i] = base + (i < extra ? 1 : 0);
        start += (*count)[i];
    }
}
}  // namespace

struct msk144_pool {
    msk144_config cfg{};
    std::vector<msk144_decoder*> dec;
    std::vector<int> device;
    std::vector<msk144_line> last_lines;      // merged output of the most recent msk144_pool_decode_host
    std::vector<msk144_result> last_results;
    // streaming: channel ranges fixed by the first push
    int il_nchan = 0;
    std::vector<int> first, count;
    std::vector<msk144_line> out_lines;
    size_t out_head = 0;
    // a member failed in mid-stream: the members are no longer at the same block, so every further push is refused
    // until msk144_pool_reset_stream (lines already produced stay pollable)
    bool stream_failed = false;
    std::string stream_error;
};

extern "C" {

const char* msk144_pool_last_error(void) { return g_pool_err.c_str(); }

int msk144_pool_create(const msk144_config* cfg, const int32_t* devices, int32_t ndevices, msk144_pool** out) {
    if (!cfg || !out || !devices || ndevices < 1) return pfail(MSK144_EINVAL, "bad argument");
    *out = nullptr;
    msk144_pool* p = new msk144_pool();
    p->cfg = *cfg;
    for (int i = 0; i < ndevices; i++) {
        msk144_config c = *cfg;
        c.device = devices[i];
        msk144_decoder* d = nullptr;
        const int rc = msk144_create(&c, &d);
        if (rc != MSK144_OK) {
            const std::string why = msk144_last_error();
            for (auto* q : p->dec) msk144_destroy(q);
            delete p;
            return pfail(rc, "device " + std::to_string(devices[i]) + ": " + why);
        }
        p->dec.push_back(d);
        p->device.push_back(devices[i]);
    }
    *out = p;
    return MSK144_OK;
}

void msk144_pool_destroy(msk144_pool* p) {
    if (!p) return;
    for (auto* d : p->dec) msk144_destroy(d);
    delete p;
}

int msk144_pool_size(const msk144_pool* p) { return p ? static_cast<int>(p->dec.size()) : 0; }

msk144_decoder* msk144_pool_member(msk144_pool* p, int32_t i) {
    return (p && i >= 0 && i < static_cast<int>(p->dec.size())) ? p->dec[i] : nullptr;
}

// Fewer channels than members: channel c gets members [mfirst[c], mfirst[c] + mcount[c]) and its blocks are cut into
// that many contiguous ranges (SURVEY.md section 8e "Partitioning").  A member runs the GPU stages of its range from
// samples [2592 b0, 2592 (b1 - 1) + 5184) of the channel; what crosses blocks (SNR recurrence, hash tables) is replayed
// here once per channel, in block order, over the members' powers and LDPC successes.
static int pool_decode_time_split(msk144_pool* p, const void* samples, size_t spc, int nchan) {
    const int n = static_cast<int>(p->dec.size());
    const int64_t bpc = spc >= MSK144_BLOCK_SAMPLES ? static_cast<int64_t>((spc - MSK144_BLOCK_SAMPLES) / MSK144_HOP_SAMPLES) + 1 : 0;
    std::vector<int> mfirst, mcount;
    split(n, nchan, &mfirst, &mcount);
    struct Part {
        int channel = -1;
        int64_t b0 = 0, b1 = 0;
        std::vector<msk144_result> results;
        std::vector<float> powers;
        int rc = MSK144_OK;
        std::string err;
    };
    std::vector<Part> part(n);
    for (int c = 0; c < nchan; c++) {
        const int64_t base = bpc / mcount[c], extra = bpc % mcount[c];
        int64_t b = 0;
        for (int j = 0; j < mcount[c]; j++) {
            Part& pt = part[mfirst[c] + j];
            pt.channel = c, pt.b0 = b, pt.b1 = b + base + (j < extra ? 1 : 0);
            b = pt.b1;
        }
    }
    auto work = [&](int i) {
        Part& pt = part[i];
        const int64_t nb = pt.b1 - pt.b0;
        if (pt.channel < 0 || nb <= 0) return;
        const char* src = static_cast<const char*>(samples) +
                          (static_cast<size_t>(pt.channel) * spc + static_cast<size_t>(pt.b0) * MSK144_HOP_SAMPLES) * 2;
        const size_t ns = static_cast<size_t>(nb - 1) * MSK144_HOP_SAMPLES + MSK144_BLOCK_SAMPLES;
        size_t rcap = 1u << 16, nr = 0;
        int64_t got = 0;
        pt.results.resize(rcap);
        pt.powers.resize(static_cast<size_t>(nb) * 8);
        pt.rc = msk144_decode_chunk_host(p->dec[i], src, ns, pt.b0, pt.results.data(), rcap, &nr, pt.powers.data(),
                                         static_cast<size_t>(nb), &got);
        if (pt.rc == MSK144_EOVERFLOW) {
            rcap = nr;
            pt.results.resize(rcap);
            pt.rc = msk144_last_output(p->dec[i], nullptr, 0, nullptr, pt.results.data(), rcap, &nr);
        }
        if (pt.rc != MSK144_OK) pt.err = msk144_last_error();
        pt.results.resize(std::min(nr, rcap));
    };
    std::vector<std::thread> th;
    for (int i = 1; i < n; i++) th.emplace_back(work, i);
    work(0);
    for (auto& t : th) t.join();
    for (int i = 0; i < n; i++)
        if (part[i].rc != MSK144_OK) return pfail(part[i].rc, "device " + std::to_string(p->device[i]) + ": " + part[i].err);
    p->last_lines.clear(), p->last_results.clear();
    std::vector<float> powers;
    std::vector<msk144_line> lv;
    for (int c = 0; c < nchan; c++) {
        const size_t r0 = p->last_results.size();
        powers.clear();
        for (int j = 0; j < mcount[c]; j++) {
            Part& pt = part[mfirst[c] + j];
            powers.insert(powers.end(), pt.powers.begin(), pt.powers.end());
            p->last_results.insert(p->last_results.end(), pt.results.begin(), pt.results.end());
        }
        size_t nl = 0;
        lv.resize(std::max<size_t>(lv.size(), 4096));
        int rc = msk144_replay_channel(p->cfg.flags, c, 0, bpc, powers.data(), p->last_results.data() + r0,
                                       p->last_results.size() - r0, lv.data(), lv.size(), &nl);
        if (rc == MSK144_EOVERFLOW) {
            lv.resize(nl);
            rc = msk144_replay_channel(p->cfg.flags, c, 0, bpc, powers.data(), p->last_results.data() + r0,
                                       p->last_results.size() - r0, lv.data(), lv.size(), &nl);
        }
        if (rc != MSK144_OK) return pfail(rc, std::string("replay: ") + msk144_last_error());
        p->last_lines.insert(p->last_lines.end(), lv.begin(), lv.begin() + nl);
        for (size_t k = r0; k < p->last_results.size(); k++) p->last_results[k].channel = c;
    }
    return MSK144_OK;
}

int msk144_pool_decode_host(msk144_pool* p, const void* samples, size_t spc, int nchan, msk144_line* lines,
                            size_t line_cap, size_t* nlines, msk144_result* results, size_t result_cap,
                            size_t* nresults) {
    if (!p || !samples || nchan < 1) return pfail(MSK144_EINVAL, "bad argument");
    const int n = static_cast<int>(p->dec.size());
    if (nchan < n) {
        if (int rc = pool_decode_time_split(p, samples, spc, nchan)) return rc;
        return msk144_pool_last_output(p, lines, line_cap, nlines, results, result_cap, nresults);
    }
    std::vector<int> first, count;
    split(nchan, n, &first, &count);
    struct Part {
        std::vector<msk144_line> lines;
        std::vector<msk144_result> results;
        int rc = MSK144_OK;
        std::string err;
    };
    std::vector<Part> part(n);
    const bool want_results = results != nullptr || nresults != nullptr;
    auto work = [&](int i) {
        Part& pt = part[i];
        if (count[i] == 0) return;
        const char* src = static_cast<const char*>(samples) + static_cast<size_t>(first[i]) * spc * 2;
        size_t lcap = 4096, rcap = want_results ? (1u << 16) : 0;
        pt.lines.resize(lcap);
        pt.results.resize(rcap);
        size_t nl = 0, nr = 0;
        pt.rc = msk144_decode_host(p->dec[i], src, spc, count[i], pt.lines.data(), lcap, &nl,
                                   rcap ? pt.results.data() : nullptr, rcap, rcap ? &nr : nullptr);
        if (pt.rc == MSK144_EOVERFLOW) {  // the member keeps its output: fetch it into arrays of the reported sizes
            lcap = std::max(lcap, nl), rcap = rcap ? std::max(rcap, nr) : 0;
            pt.lines.resize(lcap);
            pt.results.resize(rcap);
            pt.rc = msk144_last_output(p->dec[i], pt.lines.data(), lcap, &nl, rcap ? pt.results.data() : nullptr, rcap,
                                       rcap ? &nr : nullptr);
        }
        if (pt.rc != MSK144_OK) pt.err = msk144_last_error();
        pt.lines.resize(std::min(nl, lcap));
        pt.results.resize(std::min(nr, rcap));
        for (auto& l : pt.lines) l.channel += first[i];
        for (auto& r : pt.results) r.channel += first[i];
    };
    std::vector<std::thread> th;
    for (int i = 1; i < n; i++) th.emplace_back(work, i);
    work(0);
    for (auto& t : th) t.join();
    for (int i = 0; i < n; i++)
        if (part[i].rc != MSK144_OK) return pfail(part[i].rc, "device " + std::to_string(p->device[i]) + ": " + part[i].err);
    // members hold contiguous channel ranges in order, and each returns its lines channel-major: concatenation
    // is the order a single decoder produces
    p->last_lines.clear(), p->last_results.clear();
    for (int i = 0; i < n; i++) {
        p->last_lines.insert(p->last_lines.end(), part[i].lines.begin(), part[i].lines.end());
        p->last_results.insert(p->last_results.end(), part[i].results.begin(), part[i].results.end());
    }
    return msk144_pool_last_output(p, lines, line_cap, nlines, results, result_cap, nresults);
}

int msk144_pool_last_output(msk144_pool* p, msk144_line* lines, size_t line_cap, size_t* nlines, msk144_result* results,
                            size_t result_cap, size_t* nresults) {
    if (!p) return pfail(MSK144_EINVAL, "null argument");
    const size_t nl = p->last_lines.size(), nr = p->last_results.size();
    if (lines) std::memcpy(lines, p->last_lines.data(), std::min(nl, line_cap) * sizeof(msk144_line));
    if (results) std::memcpy(results, p->last_results.data(), std::min(nr, result_cap) * sizeof(msk144_result));
    if (nlines) *nlines = nl;
    if (nresults) *nresults = nr;
    if ((lines && nl > line_cap) || (results && nr > result_cap)) return pfail(MSK144_EOVERFLOW, "output array too small");
    return MSK144_OK;
}

int msk144_pool_push_interleaved(msk144_pool* p, const void* bytes, size_t nbytes, int nchan) {
    if (!p || (!bytes && nbytes) || nchan < 1) return pfail(MSK144_EINVAL, "bad argument");
    if (p->stream_failed)
        return pfail(MSK144_EINVAL, "an earlier push failed (" + p->stream_error + "): msk144_pool_reset_stream before pushing again");
    const int n = static_cast<int>(p->dec.size());
    if (p->il_nchan == 0) {
        p->il_nchan = nchan;
        split(nchan, n, &p->first, &p->count);
    } else if (p->il_nchan != nchan) {
        return pfail(MSK144_EINVAL, "channel count changed in mid-stream (msk144_pool_reset_stream first)");
    }
    const size_t frame = 2 * static_cast<size_t>(nchan);
    if (nbytes % frame) return pfail(MSK144_EINVAL, "push a whole number of time steps (2 * nchan bytes each)");
    const size_t steps = nbytes / frame;
    struct Part {
        std::vector<unsigned char> sub;
        std::vector<msk144_line> lines;
        int rc = MSK144_OK;
        std::string err;
    };
    std::vector<Part> part(n);
    auto work = [&](int i) {
        Part& pt = part[i];
        const int c0 = p->first[i], cn = p->count[i];
        if (cn == 0) return;
        // this member's columns of every time step
        const size_t sub_frame = 2 * static_cast<size_t>(cn);
        pt.sub.resize(steps * sub_frame);
        const unsigned char* src = static_cast<const unsigned char*>(bytes) + 2 * static_cast<size_t>(c0);
        for (size_t t = 0; t < steps; t++) std::memcpy(pt.sub.data() + t * sub_frame, src + t * frame, sub_frame);
        // the interleaved form for a single channel too: it is what ONE decoder would run for this column, and unlike
        // msk144_push it never looks at MSK144_FLAG_SKIP_WAV_HEADER (a multi-channel stream has no RIFF header)
        pt.rc = msk144_push_interleaved(p->dec[i], pt.sub.data(), pt.sub.size(), cn);
        if (pt.rc != MSK144_OK) {
            pt.err = msk144_last_error();
            return;
        }
        std::vector<msk144_line> buf(1024);
        size_t got;
        do {
            got = 0;
            const int prc = msk144_poll(p->dec[i], buf.data(), buf.size(), &got);
            if (prc != MSK144_OK) {
                pt.rc = prc, pt.err = msk144_last_error();
                return;
            }
            for (size_t k = 0; k < got; k++) {
                buf[k].channel += c0;
                pt.lines.push_back(buf[k]);
            }
        } while (got == buf.size());
    };
    std::vector<std::thread> th;
    for (int i = 1; i < n; i++) th.emplace_back(work, i);
    work(0);
    for (auto& t : th) t.join();
    // members hold contiguous channel ranges in order; (block, channel) order is what one decoder's
    // msk144_push_interleaved emits, independent of how the pushes were chunked.  If a member failed, the lines the
    // others produced from this push are still delivered (they have consumed their columns and will not produce them
    // again), and the stream is marked failed.
    std::vector<msk144_line> merged;
    for (int i = 0; i < n; i++) merged.insert(merged.end(), part[i].lines.begin(), part[i].lines.end());
    std::stable_sort(merged.begin(), merged.end(), [](const msk144_line& a, const msk144_line& b) {
        return a.block_index != b.block_index ? a.block_index < b.block_index : a.channel < b.channel;
    });
    p->out_lines.insert(p->out_lines.end(), merged.begin(), merged.end());
    for (int i = 0; i < n; i++)
        if (part[i].rc != MSK144_OK) {
            p->stream_failed = true;
            p->stream_error = "device " + std::to_string(p->device[i]) + ": " + part[i].err;
            return pfail(part[i].rc, p->stream_error);
        }
    return MSK144_OK;
}

int msk144_pool_poll(msk144_pool* p, msk144_line* lines, size_t cap, size_t* nlines) {
    if (!p || !nlines) return pfail(MSK144_EINVAL, "null argument");
    size_t k = 0;
    while (k < cap && p->out_head < p->out_lines.size()) lines[k++] = p->out_lines[p->out_head++];
    if (p->out_head == p->out_lines.size()) p->out_lines.clear(), p->out_head = 0;
    *nlines = k;
    return MSK144_OK;
}

int msk144_pool_reset_stream(msk144_pool* p) {
    if (!p) return pfail(MSK144_EINVAL, "null argument");
    for (auto* d : p->dec) msk144_reset_stream(d);
    p->il_nchan = 0, p->first.clear(), p->count.clear(), p->out_lines.clear(), p->out_head = 0;
    p->stream_failed = false, p->stream_error.clear();
    return MSK144_OK;
}

}  // extern "C"
