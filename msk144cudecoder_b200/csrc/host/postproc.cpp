// postproc.cpp -- see postproc.h
#include "postproc.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <ctime>
#include <iomanip>
#include <map>
#include <set>
#include <sstream>

#include "unpack77.h"

namespace mskhost {

void SnrTracker::process_powers(const float* arr) {
    // snr_tracker.cu:34-69
    float summ = 0.0f;
    for (int i = 0; i < 8; i++) summ += arr[i];
    const float avg = summ / 8;
    const float peak = *std::max_element(arr, arr + 8);
    if (noise_ <= 0.0f)
        noise_ = avg;  // initial
    else if (avg > noise_)
        noise_ = 0.9f * noise_ + 0.1f * avg;  // slow to rise
    else
        noise_ = avg;  // quick to fall
    if (noise_ > 0.0f)
        snr_ = 10.0f * std::log10(peak / noise_ - 1.0f);
    else
        snr_ = 0.0f;
    if (snr_ > 24.0f) snr_ = 24.0f;
    if (snr_ < -8.0f) snr_ = -8.0f;
}

bool decode_message(const uint8_t* m, std::string* text, void* hash_state) {
    const int n3 = (m[71] << 2) | (m[72] << 1) | m[73];
    const int i3 = (m[74] << 2) | (m[75] << 1) | m[76];
    if ((i3 == 0 && (n3 == 1 || n3 == 3 || n3 == 4 || n3 > 5)) || i3 == 3 || i3 > 5) return false;
    char c77[77];
    for (int i = 0; i < 77; i++) c77[i] = static_cast<char>('0' + m[i]);
    const int nrx = 1;  // receive semantics
    char msg[37];
    int ok = 0;
    msk144_unpack77_in(hash_state, c77, &nrx, msg, &ok);
    if (!ok) return false;
    int len = 37;
    while (len > 0 && msg[len - 1] == ' ') len--;
    text->assign(msg, msg + len);
    return true;
}

std::vector<Line> block_lines(const std::vector<Present>& items, bool compat, int snr, void* hash_state) {
    std::vector<Line> scope;
    std::map<std::string, std::pair<bool, std::string>> cache;
    bool have_first = false;
    std::pair<bool, std::string> first;
    for (const Present& it : items) {
        std::pair<bool, std::string> res;
        if (compat) {
            // The reference keys its per-block cache with a comparator that never compares
            // (main.cu:437-445), so the first candidate's outcome is reused for the whole block.
            if (!have_first) {
                first.first = decode_message(it.message, &first.second, hash_state);
                have_first = true;
            }
            res = first;
        } else {
            const std::string key(reinterpret_cast<const char*>(it.message), 77);
            auto f = cache.find(key);
            if (f == cache.end()) {
                std::pair<bool, std::string> r;
                r.first = decode_message(it.message, &r.second, hash_state);
                f = cache.emplace(key, r).first;
            }
            res = f->second;
        }
        if (!res.first) continue;
        Line l;
        l.snr = snr;
        l.f0 = it.f0;
        l.num_avg = it.num_avg;
        l.nbadsync = it.nbadsync;
        l.pattern_idx = it.pattern_idx;
        l.message = res.second;
        std::memcpy(l.payload, it.message, 80);
        scope.push_back(l);
    }
    // ResultFilter::blockEnd: one line per distinct text, lowest num_avg then lowest nbadsync
    std::vector<Line> result;
    std::set<std::string> only;
    for (const Line& e : scope) only.insert(e.message);
    for (const std::string& msg : only) {
        std::vector<Line> filtered;
        std::copy_if(scope.begin(), scope.end(), std::back_inserter(filtered),
                     [&msg](const Line& x) { return x.message == msg; });
        std::sort(filtered.begin(), filtered.end(), [](const Line& a, const Line& b) {
            if (a.num_avg == b.num_avg) return a.nbadsync < b.nbadsync;
            return a.num_avg < b.num_avg;
        });
        result.push_back(filtered.front());
    }
    return result;
}

std::string format_line(int snr, float f0, int num_avg, int nbadsync, int pattern_idx, const char* date14,
                        const char* text) {
    std::ostringstream os;  // same stream manipulators as main.cu:409-417
    os << "***  "
       << "snr=" << std::setw(2) << snr << "; "
       << "f0=" << std::setw(6) << f0 << "; "
       << "num_avg=" << num_avg << "; "
       << "nbadsync=" << nbadsync << "; "
       << "pattern_idx=" << pattern_idx << "; "
       << "date=" << date14 << "; "
       << "msg='" << text << "'"
       << "; ";
    return os.str();
}

std::string now_stamp() {
    const std::time_t now_c = std::time(nullptr);
    std::tm tmv;
    localtime_r(&now_c, &tmv);
    char buff[70];
    std::strftime(buff, sizeof buff, "%Y%m%d%H%M%S", &tmv);
    return std::string(buff);
}

}  // namespace mskhost
