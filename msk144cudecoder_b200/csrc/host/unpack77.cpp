// unpack77.cpp -- 77-bit WSJT-X message payload -> text.
//
// The reference calls the gfortran symbol __packjt77_MOD_unpack77 of WSJT-X
// (lib/77bit/packjt77.f90, pinned d28164e9) through f_interop.cpp:61-71 from
// decode_softbits.cpp:41-44.  That Fortran is not in the reference tree (empty submodule) and there
// is no Fortran compiler here, so this is a C++ statement of the published 77-bit message protocol
// (Franke/Somerville/Taylor, "The FT4 and FT8 Communication Protocols", QEX 2020, tables 1-3) with
// the same calling convention.  Text parity with WSJT-X is therefore UNPINNED; payload
// bits are pinned.  Covered: every i3.n3 the reference lets through (decode_softbits.cpp:30):
// 0.0 free text, 0.5 telemetry, 1/2 standard (+/R,/P), 4 non-standard call, 5 EU VHF (hashed);
// 0.2 is retired in the protocol and reports failure.  Call-sign hash tables persist across calls
// as in WSJT-X (nrx=1 receive semantics; no "my call"/"dx call" is ever set by the reference).
#include "unpack77.h"

#include <dlfcn.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>

namespace {

constexpr int NTOKENS = 2063592;
constexpr int MAX22 = 4194304;
constexpr int MAXGRID4 = 32400;
constexpr int MAXHASH = 20;

// Call-sign hash tables (WSJT-X keeps them in module variables: one set per process).  Here one set per stream:
// every decoder channel owns a HashState (msk144_hash_state_new) and names it for the duration of a text decode, so
// what channel A heard never changes how channel B's <...> fields print, and concurrent decoders do not share state.
// The process-wide set remains for callers of the bare gfortran-ABI symbol.
struct HashState {
    std::unordered_map<int, std::string> calls10, calls12;   // sparse: a channel hears a handful of calls
    int ihash22[MAXHASH];
    std::string calls22[MAXHASH];
    HashState() {
        for (int& h : ihash22) h = -1;
    }
};
HashState* g_hash = nullptr;
thread_local HashState* t_hash = nullptr;  // the state named by the caller for this thread, if any
std::mutex g_mu;

HashState& hs() {
    if (t_hash) return *t_hash;
    if (!g_hash) g_hash = new HashState();
    return *g_hash;
}

std::string rtrim(std::string s) {
    size_t n = s.size();
    while (n > 0 && s[n - 1] == ' ') n--;
    s.resize(n);
    return s;
}
std::string ltrim(std::string s) {
    size_t i = 0;
    while (i < s.size() && s[i] == ' ') i++;
    return s.substr(i);
}

uint64_t bits(const char* c77, int start, int len) {
    uint64_t v = 0;
    for (int i = 0; i < len; i++) v = (v << 1) | (uint64_t)(c77[start + i] == '1');
    return v;
}

// m-bit hash of a call sign: base-38 value of the 11-character left-justified call, times a fixed
// 64-bit multiplier, top m bits.
int ihashcall(const std::string& call, int m) {
    static const char* A = " 0123456789ABCDEFGHIJKLMNOPQRSTUVWXYZ/";
    std::string c = call;
    if (!c.empty() && c[0] == '<') c = c.substr(1);
    size_t gt = c.find('>');
    if (gt != std::string::npos) c = c.substr(0, gt);
    c.resize(11, ' ');
    uint64_t n8 = 0;
    for (int i = 0; i < 11; i++) {
        const char* p = std::strchr(A, c[i]);
        const int j = (p && c[i]) ? (int)(p - A) : -1;  // unknown characters count as -1 (index()-1)
        n8 = 38 * n8 + (uint64_t)(int64_t)j;
    }
    return (int)((47055833459ULL * n8) >> (64 - m));
}

void save_hash_call(const std::string& c13) {
    std::string cw = c13;
    if (cw.empty() || cw[0] == ' ' || cw.compare(0, 5, "<...>") == 0) return;
    if (cw[0] == '<') cw = cw.substr(1);
    size_t gt = cw.find('>');
    if (gt != std::string::npos) cw = cw.substr(0, gt);
    cw = rtrim(cw);
    if (cw.size() < 3) return;
    HashState& h = hs();
    const int n10 = ihashcall(cw, 10);
    if (n10 >= 0 && n10 <= 1023) h.calls10[n10] = cw;
    const int n12 = ihashcall(cw, 12);
    if (n12 >= 0 && n12 <= 4095) h.calls12[n12] = cw;
    const int n22 = ihashcall(cw, 22);
    bool exists = false;
    for (int i = 0; i < MAXHASH; i++)
        if (h.ihash22[i] == n22) {
            h.calls22[i] = cw;
            exists = true;
        }
    if (exists) return;
    for (int i = MAXHASH - 1; i >= 1; i--) {
        h.ihash22[i] = h.ihash22[i - 1];
        h.calls22[i] = h.calls22[i - 1];
    }
    h.ihash22[0] = n22;
    h.calls22[0] = cw;
}

std::string hash12(int n12) {
    if (n12 < 0 || n12 > 4095) return "<...>";
    const auto& tab = hs().calls12;
    const auto f = tab.find(n12);
    if (f == tab.end() || f->second.empty()) return "<...>";
    return "<" + f->second + ">";
}
std::string hash22(int n22) {
    HashState& h = hs();
    for (int i = 0; i < MAXHASH; i++)
        if (h.ihash22[i] == n22 && n22 >= 0) return "<" + h.calls22[i] + ">";
    return "<...>";
}

// 28-bit field -> token, hashed call or standard call.
bool unpack28(int n28, std::string* out) {
    static const char* C1 = " 0123456789ABCDEFGHIJKLMNOPQRSTUVWXYZ";
    static const char* C2 = "0123456789ABCDEFGHIJKLMNOPQRSTUVWXYZ";
    static const char* C3 = "0123456789";
    static const char* C4 = " ABCDEFGHIJKLMNOPQRSTUVWXYZ";
    std::string c13;
    bool done = false;
    if (n28 < NTOKENS) {
        if (n28 == 0) c13 = "DE", done = true;
        else if (n28 == 1) c13 = "QRZ", done = true;
        else if (n28 == 2) c13 = "CQ", done = true;
        else if (n28 <= 1002) {
            char b[16];
            std::snprintf(b, sizeof(b), "CQ_%03d", n28 - 3);
            c13 = b;
            done = true;
        } else if (n28 <= 532443) {
            int n = n28 - 1003;
            const int i1 = n / (27 * 27 * 27);
            n -= 27 * 27 * 27 * i1;
            const int i2 = n / (27 * 27);
            n -= 27 * 27 * i2;
            const int i3 = n / 27;
            const int i4 = n - 27 * i3;
            std::string s;
            s += C4[i1];
            s += C4[i2];
            s += C4[i3];
            s += C4[i4];
            c13 = "CQ_" + ltrim(s);
            done = true;
        }
    }
    if (!done) {
        n28 -= NTOKENS;
        if (n28 < MAX22) {
            c13 = hash22(n28);
        } else {
            int n = n28 - MAX22;
            const int i1 = n / (36 * 10 * 27 * 27 * 27);
            n -= 36 * 10 * 27 * 27 * 27 * i1;
            const int i2 = n / (10 * 27 * 27 * 27);
            n -= 10 * 27 * 27 * 27 * i2;
            const int i3 = n / (27 * 27 * 27);
            n -= 27 * 27 * 27 * i3;
            const int i4 = n / (27 * 27);
            n -= 27 * 27 * i4;
            const int i5 = n / 27;
            const int i6 = n - 27 * i5;
            if (i1 > 36) {  // beyond the defined range of the field
                *out = "QU1RK";
                return false;
            }
            std::string s;
            s += C1[i1];
            s += C2[i2];
            s += C3[i3];
            s += C4[i4];
            s += C4[i5];
            s += C4[i6];
            c13 = ltrim(s);
            // the protocol's work-arounds for two prefixes that do not fit the 6-character pattern (the packer maps
            // 3DA0xxx -> 3D0xxx and 3Xa... -> Qa...; undone here)
            if (c13.compare(0, 3, "3D0") == 0 && c13.size() > 3 && c13[3] != ' ') c13 = "3DA0" + c13.substr(3);
            else if (c13.size() > 1 && c13[0] == 'Q' && c13[1] >= 'A' && c13[1] <= 'Z') c13 = "3X" + c13.substr(1);
        }
    }
    c13 = rtrim(c13);
    // an embedded blank means the field did not hold a valid call
    if (c13.find(' ') != std::string::npos) {
        *out = "QU1RK";
        return false;
    }
    *out = c13;
    return true;
}

bool to_grid4(int n, std::string* g) {
    const int j1 = n / (18 * 10 * 10);
    if (j1 < 0 || j1 > 17) return false;
    n -= j1 * 18 * 10 * 10;
    const int j2 = n / (10 * 10);
    if (j2 < 0 || j2 > 17) return false;
    n -= j2 * 10 * 10;
    const int j3 = n / 10;
    if (j3 < 0 || j3 > 9) return false;
    const int j4 = n - j3 * 10;
    if (j4 < 0 || j4 > 9) return false;
    g->clear();
    *g += (char)('A' + j1);
    *g += (char)('A' + j2);
    *g += (char)('0' + j3);
    *g += (char)('0' + j4);
    return true;
}

bool to_grid6(int n, std::string* g) {
    const int j1 = n / (18 * 10 * 10 * 24 * 24);
    if (j1 < 0 || j1 > 17) return false;
    n -= j1 * 18 * 10 * 10 * 24 * 24;
    const int j2 = n / (10 * 10 * 24 * 24);
    if (j2 < 0 || j2 > 17) return false;
    n -= j2 * 10 * 10 * 24 * 24;
    const int j3 = n / (10 * 24 * 24);
    n -= j3 * 10 * 24 * 24;
    const int j4 = n / (24 * 24);
    n -= j4 * 24 * 24;
    const int j5 = n / 24;
    const int j6 = n - j5 * 24;
    if (j3 > 9 || j4 > 9 || j5 > 23 || j6 > 23) return false;
    g->clear();
    *g += (char)('A' + j1);
    *g += (char)('A' + j2);
    *g += (char)('0' + j3);
    *g += (char)('0' + j4);
    *g += (char)('A' + j5);
    *g += (char)('A' + j6);
    return true;
}

void finish_call(std::string* call, int ip, int i3) {
    if (call->compare(0, 3, "CQ_") == 0) (*call)[2] = ' ';
    if (call->find('<') == std::string::npos) {
        // position of the first blank in the 13-character field: len+1 (1-based) for a trimmed call
        const int i = (int)call->size() + 1;
        if (call->find(' ') != std::string::npos) return;  // "CQ xxx": never gets a suffix or a hash
        if (i >= 4 && ip == 1 && i3 == 1) *call += "/R";
        if (i >= 4 && ip == 1 && i3 == 2) *call += "/P";
        if (i >= 4) save_hash_call(*call);
    }
}

bool unpack77_impl(const char* c77, int /*nrx*/, std::string* msg) {
    for (int i = 0; i < 77; i++)
        if (c77[i] != '0' && c77[i] != '1') {
            *msg = "failed unpack";
            return false;
        }
    const int n3 = (int)bits(c77, 71, 3);
    const int i3 = (int)bits(c77, 74, 3);
    bool ok = true;
    msg->clear();

    if (i3 == 0 && n3 == 0) {
        // 0.0 free text: 71 bits, base 42, 13 characters
        static const char* A = " 0123456789ABCDEFGHIJKLMNOPQRSTUVWXYZ+-./?";
        unsigned __int128 v = 0;
        for (int i = 0; i < 71; i++) v = (v << 1) | (unsigned)(c77[i] == '1');
        char t[14];
        for (int i = 12; i >= 0; i--) {
            t[i] = A[(int)(v % 42)];
            v /= 42;
        }
        t[13] = 0;
        *msg = ltrim(std::string(t));
        if (msg->empty()) ok = false;
    } else if (i3 == 0 && n3 == 5) {
        // 0.5 telemetry: 23 + 24 + 24 bits as 18 hex digits, leading zeros dropped
        char t[24];
        std::snprintf(t, sizeof(t), "%06X%06X%06X", (unsigned)bits(c77, 0, 23), (unsigned)bits(c77, 23, 24),
                      (unsigned)bits(c77, 47, 24));
        std::string s(t);
        size_t i = 0;
        while (i < 18 && s[i] == '0') i++;
        *msg = s.substr(i);
    } else if (i3 == 1 || i3 == 2) {
        const int n28a = (int)bits(c77, 0, 28), ipa = (int)bits(c77, 28, 1);
        const int n28b = (int)bits(c77, 29, 28), ipb = (int)bits(c77, 57, 1);
        const int ir = (int)bits(c77, 58, 1), igrid4 = (int)bits(c77, 59, 15);
        std::string c1, c2;
        if (!unpack28(n28a, &c1)) ok = false;
        if (!unpack28(n28b, &c2)) ok = false;
        finish_call(&c1, ipa, i3);
        finish_call(&c2, ipb, i3);
        if (igrid4 <= MAXGRID4) {
            std::string g4;
            if (!to_grid4(igrid4, &g4)) ok = false;
            *msg = c1 + " " + c2 + (ir ? " R " : " ") + g4;
            if (msg->compare(0, 3, "CQ ") == 0 && ir == 1) ok = false;
        } else {
            const int irpt = igrid4 - MAXGRID4;
            if (irpt == 1) *msg = c1 + " " + c2;
            if (irpt == 2) *msg = c1 + " " + c2 + " RRR";
            if (irpt == 3) *msg = c1 + " " + c2 + " RR73";
            if (irpt == 4) *msg = c1 + " " + c2 + " 73";
            if (irpt >= 5) {
                int isnr = irpt - 35;
                if (isnr > 50) isnr -= 101;
                char r[16];
                std::snprintf(r, sizeof(r), "%c%02d", isnr < 0 ? '-' : '+', isnr < 0 ? -isnr : isnr);
                *msg = c1 + " " + c2 + (ir ? " R" : " ") + r;
            }
            if (msg->compare(0, 3, "CQ ") == 0 && irpt >= 2) ok = false;
        }
    } else if (i3 == 4) {
        static const char* A = " 0123456789ABCDEFGHIJKLMNOPQRSTUVWXYZ/";
        const int n12 = (int)bits(c77, 0, 12);
        uint64_t n58 = bits(c77, 12, 58);
        const int iflip = (int)bits(c77, 70, 1), nrpt = (int)bits(c77, 71, 2), icq = (int)bits(c77, 73, 1);
        char c11[12];
        for (int i = 10; i >= 0; i--) {
            c11[i] = A[n58 % 38];
            n58 /= 38;
        }
        c11[11] = 0;
        const std::string full = rtrim(ltrim(std::string(c11)));
        const std::string c3 = hash12(n12);
        std::string c1, c2;
        if (iflip == 0) {
            c1 = c3;
            c2 = full;
        } else {
            c1 = full;
            c2 = c3;
        }
        save_hash_call(full);
        if (icq == 0) {
            if (nrpt == 0) *msg = c1 + " " + c2;
            if (nrpt == 1) *msg = c1 + " " + c2 + " RRR";
            if (nrpt == 2) *msg = c1 + " " + c2 + " RR73";
            if (nrpt == 3) *msg = c1 + " " + c2 + " 73";
        } else {
            *msg = "CQ " + c2;
        }
    } else if (i3 == 5) {
        const int n12 = (int)bits(c77, 0, 12), n22 = (int)bits(c77, 12, 22);
        const int ir = (int)bits(c77, 34, 1), irpt = (int)bits(c77, 35, 3);
        const int iserial = (int)bits(c77, 38, 11), igrid6 = (int)bits(c77, 49, 25);
        std::string g6;
        if (igrid6 > 18662399 || !to_grid6(igrid6, &g6)) {
            *msg = "";
            return false;
        }
        char ex[16];
        std::snprintf(ex, sizeof(ex), "%2d%04d", 52 + irpt, iserial);
        *msg = hash12(n12) + " " + hash22(n22) + (ir ? " R " : " ") + ex + " " + g6;
    } else {
        // 0.1, 0.3, 0.4, 0.6, 3.x are filtered by the caller (decode_softbits.cpp:30); 0.2 is retired
        ok = false;
    }
    if (msg->compare(0, 4, "CQ <") == 0) ok = false;
    if (msg->size() > 37) msg->resize(37);
    return ok;
}



// ---- pluggable seam (f_interop.cpp:10-29,61-71): a gfortran-built WSJT-X library exporting the module procedure
typedef void (*unpack77_fn)(char*, const int*, char*, int*, size_t, size_t);
unpack77_fn g_ext = nullptr;
void* g_ext_handle = nullptr;
bool g_env_checked = false;

int load_external(const char* path) {  // g_mu held
    if (g_ext_handle) dlclose(g_ext_handle);
    g_ext_handle = nullptr, g_ext = nullptr;
    if (!path || !*path) return 0;
    void* h = dlopen(path, RTLD_NOW | RTLD_LOCAL);
    if (!h) return -1;
    // dlsym on the handle searches that library (and its dependencies), not this one, so the user's symbol wins
    // over the alias exported below
    void* f = dlsym(h, "__packjt77_MOD_unpack77");
    if (!f) {
        dlclose(h);
        return -1;
    }
    g_ext_handle = h;
    g_ext = reinterpret_cast<unpack77_fn>(f);
    return 0;
}

}  // namespace

extern "C" {

int msk144_set_unpack77_library(const char* path) {
    std::lock_guard<std::mutex> lk(g_mu);
    g_env_checked = true;  // an explicit choice overrides the environment
    return load_external(path) == 0 ? 0 : -1;
}

int msk144_unpack77_is_external(void) {
    std::lock_guard<std::mutex> lk(g_mu);
    return g_ext ? 1 : 0;
}

void msk144_unpack77(const char c77[77], const int* nrx, char msg[37], int* ok) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_env_checked) {
        g_env_checked = true;
        if (const char* e = std::getenv("MSK144_UNPACK77_LIB")) {
            if (load_external(e) != 0) std::fprintf(stderr, "msk144: cannot use MSK144_UNPACK77_LIB=%s, built-in unpack77 stays\n", e);
        }
    }
    if (g_ext) {
        char in[77];
        std::memcpy(in, c77, 77);
        const int one = 1;
        int okv = 0;  // Fortran LOGICAL, 4 bytes
        std::memset(msg, ' ', 37);
        g_ext(in, nrx ? nrx : &one, msg, &okv, 77, 37);
        *ok = okv ? 1 : 0;
        return;
    }
    std::string s;
    const bool r = unpack77_impl(c77, nrx ? *nrx : 1, &s);
    std::memset(msg, ' ', 37);
    std::memcpy(msg, s.data(), s.size() < 37 ? s.size() : 37);
    *ok = r ? 1 : 0;
}

void msk144_unpack77_reset(void) {
    std::lock_guard<std::mutex> lk(g_mu);
    delete g_hash;
    g_hash = nullptr;
}

void* msk144_hash_state_new(void) { return new HashState(); }
void msk144_hash_state_free(void* st) { delete static_cast<HashState*>(st); }
void msk144_hash_state_clear(void* st) {
    if (st) *static_cast<HashState*>(st) = HashState();
}
void msk144_unpack77_in(void* st, const char c77[77], const int* nrx, char msg[37], int* ok) {
    HashState* prev = t_hash;
    t_hash = static_cast<HashState*>(st);
    msk144_unpack77(c77, nrx, msg, ok);
    t_hash = prev;
}

// gfortran module-procedure ABI: character arguments carry hidden trailing lengths (f_interop.cpp:10-29)
void __packjt77_MOD_unpack77(char c77[], const int* nrx, char msg[], int* ok, size_t, size_t) {
    msk144_unpack77(c77, nrx, msg, ok);
}

}  // extern "C"
