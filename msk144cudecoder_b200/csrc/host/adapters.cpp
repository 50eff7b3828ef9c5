// adapters.cpp -- the steps either side of the hot path (none of this exists in the
// reference, whose stdin is headerless raw samples and whose stdout repeats a message once per 216 ms block):
//   * RIFF/WAVE header recognition, so that a .wav file can be piped in without its header being decoded as
//     44 bytes of samples (what the reference does with demo/0001.wav: README.md:62, main.cu:273)
//   * JSON-lines rendering of a decode line
//   * aggregation of the per-block lines of one ping (the same text in overlapping / adjacent blocks) into one
//     event carrying the best line
// Host only; no GPU is needed for any of it (covered by the CPU test suite).
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../../../include/msk144_b200.h"
#include "postproc.h"

namespace {
uint32_t rd32(const uint8_t* p) { return p[0] | (p[1] << 8) | (p[2] << 16) | (static_cast<uint32_t>(p[3]) << 24); }
uint32_t rd16(const uint8_t* p) { return p[0] | (p[1] << 8); }

std::string json_escape(const char* s) {
    std::string o;
    for (; *s; ++s) {
        const unsigned char c = static_cast<unsigned char>(*s);
        if (c == '"' || c == '\\') {
            o += '\\';
            o += static_cast<char>(c);
        } else if (c < 0x20 || c >= 0x7f) {
            char b[8];
            std::snprintf(b, sizeof(b), "\\u%04x", c);
            o += b;
        } else {
            o += static_cast<char>(c);
        }
    }
    return o;
}

bool better(const msk144_line& a, const msk144_line& b) {  // the per-block rule of result_filter.cpp:61-69, then snr
    if (a.num_avg != b.num_avg) return a.num_avg < b.num_avg;
    if (a.nbadsync != b.nbadsync) return a.nbadsync < b.nbadsync;
    return a.snr > b.snr;
}
}  // namespace

struct msk144_aggregator {
    int64_t max_gap;
    // open events by (channel, text); closed events in closing order
    std::map<std::pair<int32_t, std::string>, msk144_event> open;
    std::vector<msk144_event> closed;
    void close_upto(int64_t block, int32_t channel, bool all_channels) {
        std::vector<msk144_event> out;
        for (auto it = open.begin(); it != open.end();) {
            const bool chan_ok = all_channels || it->first.first == channel;
            if (chan_ok && it->second.last_block + max_gap < block) {
                out.push_back(it->second);
                it = open.erase(it);
            } else {
                ++it;
            }
        }
        // deterministic order: by first block, then channel, then text (map order is already channel, text)
        std::stable_sort(out.begin(), out.end(), [](const msk144_event& a, const msk144_event& b) {
            if (a.first_block != b.first_block) return a.first_block < b.first_block;
            return a.best.channel < b.best.channel;
        });
        closed.insert(closed.end(), out.begin(), out.end());
    }
};

extern "C" {

int64_t msk144_wav_data_offset(const void* bytes, size_t n, int32_t* channels, int32_t* sample_rate, int32_t* bits) {
    const uint8_t* p = static_cast<const uint8_t*>(bytes);
    if (!p) return MSK144_EINVAL;
    if (std::memcmp(p, "RIFF", n < 4 ? n : 4) != 0) return 0;  // raw samples
    if (n < 12) return MSK144_EAGAIN;
    if (std::memcmp(p, "RIFF", 4) != 0 || std::memcmp(p + 8, "WAVE", 4) != 0) return 0;  // raw samples
    size_t off = 12;
    bool have_fmt = false;
    while (true) {
        if (off + 8 > n) return MSK144_EAGAIN;
        const uint32_t len = rd32(p + off + 4);
        if (std::memcmp(p + off, "fmt ", 4) == 0) {
            if (len < 16) return MSK144_EINVAL;
            if (off + 8 + 16 > n) return MSK144_EAGAIN;
            const uint32_t tag = rd16(p + off + 8);
            if (tag != 1 && tag != 0xFFFE) return MSK144_EINVAL;  // PCM (or extensible PCM) only
            if (channels) *channels = static_cast<int32_t>(rd16(p + off + 10));
            if (sample_rate) *sample_rate = static_cast<int32_t>(rd32(p + off + 12));
            if (bits) *bits = static_cast<int32_t>(rd16(p + off + 22));
            have_fmt = true;
        } else if (std::memcmp(p + off, "data", 4) == 0) {
            if (!have_fmt) return MSK144_EINVAL;
            return static_cast<int64_t>(off + 8);
        }
        off += 8 + static_cast<size_t>(len) + (len & 1u);  // chunks are word aligned
        if (off > (1u << 20)) return MSK144_EINVAL;          // no data chunk within the first MiB
    }
}

int msk144_format_line_json(const msk144_line* line, const char* date14, char* out, size_t cap) {
    if (!line || !out) return MSK144_EINVAL;
    const std::string stamp = date14 ? std::string(date14) : mskhost::now_stamp();
    char payload[78];
    for (int i = 0; i < 77; i++) payload[i] = static_cast<char>('0' + (line->payload[i] & 1));
    payload[77] = 0;
    char buf[512];
    const int n = std::snprintf(buf, sizeof(buf),
                                "{\"channel\":%d,\"block\":%lld,\"t\":%.3f,\"snr\":%d,\"f0\":%.9g,\"num_avg\":%d,\"nbadsync\":%d,"
                                "\"pattern_idx\":%d,\"date\":\"%s\",\"msg\":\"%s\",\"payload\":\"%s\"}",
                                line->channel, static_cast<long long>(line->block_index),
                                static_cast<double>(line->block_index) * MSK144_HOP_SAMPLES / 12000.0, line->snr,
                                static_cast<double>(line->f0), line->num_avg, line->nbadsync, line->pattern_idx,
                                json_escape(stamp.c_str()).c_str(), json_escape(line->text).c_str(), payload);
    std::snprintf(out, cap, "%s", buf);
    return n;
}

int msk144_aggregator_create(int32_t max_gap_blocks, msk144_aggregator** out) {
    if (!out || max_gap_blocks < 0) return MSK144_EINVAL;
    *out = new msk144_aggregator{max_gap_blocks, {}, {}};
    return MSK144_OK;
}

void msk144_aggregator_destroy(msk144_aggregator* agg) { delete agg; }

int msk144_aggregator_add(msk144_aggregator* agg, const msk144_line* lines, size_t n) {
    if (!agg || (!lines && n)) return MSK144_EINVAL;
    for (size_t i = 0; i < n; i++) {
        const msk144_line& l = lines[i];
        // lines of a channel arrive in block order: anything this line's block has left behind is finished
        agg->close_upto(l.block_index, l.channel, false);
        const auto key = std::make_pair(l.channel, std::string(l.text));
        auto it = agg->open.find(key);
        if (it == agg->open.end()) {
            msk144_event e;
            std::memset(&e, 0, sizeof(e));
            e.first_block = e.last_block = l.block_index;
            e.nlines = 1;
            e.best = l;
            agg->open.emplace(key, e);
        } else {
            msk144_event& e = it->second;
            if (l.block_index < e.first_block) e.first_block = l.block_index;
            if (l.block_index > e.last_block) e.last_block = l.block_index;
            e.nlines++;
            if (better(l, e.best)) e.best = l;
        }
    }
    return MSK144_OK;
}

int msk144_aggregator_poll(msk144_aggregator* agg, int64_t next_block, int flush, msk144_event* out, size_t cap, size_t* nout) {
    if (!agg || !nout || (!out && cap)) return MSK144_EINVAL;
    if (flush)
        agg->close_upto(INT64_MAX, 0, true);
    else
        agg->close_upto(next_block, 0, true);
    size_t n = 0;
    while (n < cap && n < agg->closed.size()) {
        out[n] = agg->closed[n];
        n++;
    }
    agg->closed.erase(agg->closed.begin(), agg->closed.begin() + static_cast<long>(n));
    *nout = n;
    return MSK144_OK;
}

}  // extern "C"
