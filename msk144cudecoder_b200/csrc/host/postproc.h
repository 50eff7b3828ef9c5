// postproc.h -- host side of the per-block loop: SNR tracker, i3/n3 gate + unpack77, per-block
// result filter, output line format.  Mirrors snr_tracker.{h,cu}, decode_softbits.{h,cpp},
// result_filter.{h,cpp} and main.cu:406-419,480-525 of the reference.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace mskhost {

// snr_tracker.cu:21-69.  Takes the 8 sub-block powers the front-end kernel already computed.
class SnrTracker {
public:
    void reset() { noise_ = 0.0f, snr_ = 0.0f; }
    void process_powers(const float* arr8);
    float snr_f() const { return snr_; }
    int snr_i() const { return static_cast<int>(snr_); }

private:
    float noise_ = 0.0f, snr_ = 0.0f;
};

// decode_softbits.cpp:23-55.  hash_state: the stream's call-sign tables (unpack77.h), NULL = process-wide.
bool decode_message(const uint8_t* msg77, std::string* text, void* hash_state = nullptr);

struct Present {  // one LDPC success of a block, in item-index order
    uint32_t item;
    float f0;
    int num_avg, nbadsync, pattern_idx;
    uint8_t message[80];
};

struct Line {  // result_filter.h:17-29
    int snr;
    float f0;
    int num_avg, nbadsync, pattern_idx;
    std::string message;
    uint8_t payload[80];
};

// main.cu:480-525 (host loop incl. decode_cache) followed by ResultFilter::blockEnd
// (result_filter.cpp:43-74).  `items` must be sorted by item index.
std::vector<Line> block_lines(const std::vector<Present>& items, bool compat_decode_cache, int snr,
                              void* hash_state = nullptr);

// main.cu:406-419.  date14: "%Y%m%d%H%M%S".
std::string format_line(int snr, float f0, int num_avg, int nbadsync, int pattern_idx, const char* date14,
                        const char* text);
std::string now_stamp();  // result_filter.cpp:13-21

}  // namespace mskhost
