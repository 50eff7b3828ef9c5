/*
 * msk144_oracle.h -- C API of the CPU oracle.
 *
 * TEST INFRASTRUCTURE ONLY.  This is a plain C++ restatement, on the CPU, of the
 * algorithm of alexander-sholohov/msk144cudecoder (the "reference").  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load it.  The product (msk144cudecoder_b200/) never links, imports or calls it.
 *
 * Parity status: the reference ships no golden vectors and no tests.  The oracle is
 * pinned against outputs of the reference itself, recompiled for sm_100a from
 * /root/reference/src and run on a B200 (oracle/build_ref.sh, oracle/ref_dump.cu,
 * tests/golden/).  Text (unpack77) has no source in the reference tree at all
 * (WSJT-X submodule absent): text parity is unpinned, payload-bit parity is pinned.
 *
 * Every function cites the reference file:line it follows.
 */
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_NSAMP = 5184, ORC_NFRAME = 864, ORC_NLLR = 128, ORC_NMSG = 77 };

/* One candidate record == ResultKeeper::ResultItem (result_keeper.cuh:16-30). */
typedef struct orc_item {
    uint32_t freq_idx;    /* reference: block_idx (CUDA block == frequency hypothesis) */
    uint32_t pattern_idx;
    uint32_t pos;
    float    f0;
    int32_t  nbadsync;
    float    xb;
    int32_t  num_avg;
    float    llr[ORC_NLLR]; /* softbits_wo_sync */
    int32_t  is_message_present;
    int32_t  ldpc_iters;
    int32_t  ldpc_nhard;
    uint8_t  message[80];   /* 77 used, values 0/1 */
} orc_item;

/* One printed decode line (ResultFilter::ResultItem, result_filter.h:17-29). */
typedef struct orc_line {
    int64_t block_index;
    int32_t snr;
    float   f0;
    int32_t num_avg;
    int32_t nbadsync;
    int32_t pattern_idx;
    uint8_t payload[80];   /* 77 bits of the candidate the line was printed for */
    char    text[40];
} orc_line;

typedef struct orc_config {
    float center_hz;
    float width_hz;
    float step_hz;
    int   scan_depth;
    int   nbadsync_threshold;
    int   read_mode;        /* 1 = s16 audio, 2 = s8 IQ */
    int   analytic_method;  /* 1 = FFT, 2 = shift/LPF/shift */
    int   scan_mode;        /* 0 = reference summation order (brute force), 1 = linear reformulation */
    int   compat_cache;     /* 1 = reproduce the decode_cache comparator quirk (main.cu:430-447) */
    int   nthreads;         /* OpenMP threads, 0 = default */
    int   wsjtx_bp;         /* 1 = LDPC with WSJT-X bpdecode128_90 control flow (11 checks, early stop) instead of the
                               reference kernel's (10 checks, none) */
} orc_config;

/* ---- grid / tables (msk_context.cuh:95-113,137-248) ---- */
int   orc_num_freqs(float width_hz, float step_hz);
float orc_freq(float center_hz, float width_hz, float step_hz, int idx);
int   orc_effective_depth(int scan_depth);
void  orc_sync_waveform(float* cb42_re_im /*84*/, float* pp12 /*12*/);
void  orc_pattern(int pattern_idx, int* num_avg, int* mask6);

/* ---- front end ---- */
/* main.cu:300-332 + analytic2.cuh:235-258 (method 2) or analytic_fft.cu:18-157 (method 1) */
void orc_frontend_audio(const int16_t* s16 /*5184*/, int method, float* out_re_im /*2*5184*/);
/* main.cu:365-374 + analytic2.cuh:260-281 */
void orc_frontend_iq(const int8_t* iq /*2*5184*/, float* out_re_im /*2*5184*/);

/* ---- SNR tracker (snr_tracker.cu:21-69) ---- */
void  orc_snr_powers(const float* blk_re_im, float* powers8);
/* state[0] = noise power (0 at start); returns the float snr, *snri = truncated int */
float orc_snr_update(float* state, const float* powers8, int* snri);

/* ---- scan (scan_kernel.cuh:27-364) ---- */
/* items: nf*depth*8 records; fills freq_idx, pattern_idx, pos, xb, f0, num_avg */
void orc_scan(const float* blk_re_im, const orc_config* cfg, orc_item* items);

/* ---- softbits for one candidate (softbits_kernel.cuh:9-249) ---- */
void orc_softbits(const float* blk_re_im, float freq_hz, int pattern_idx, uint32_t pos,
                  float* llr128, int* nbadsync, float* softbits144 /* may be NULL */);

/* ---- LDPC(128,90) (ldpc_kernel.cuh:100-249, ldpc_context.cuh) ---- */
int      orc_ldpc_decode(const float* llr128, uint8_t* msg77, int* iters, int* nhard);
void     orc_ldpc_decode_batch(const float* llr, size_t ncw, uint8_t* found, uint8_t* msg77,
                               int32_t* iters, int32_t* nhard, int nthreads);
uint32_t orc_crc13(const uint8_t* bits77);
/* the same decoder with WSJT-X bpdecode128_90 control flow (11 checks, early stop, a-priori mask; see the .cpp);
 * apmask128: 128 bytes of 0/1 common to all codewords, or NULL */
void     orc_ldpc_decode_wsjtx_batch(const float* llr, size_t ncw, const uint8_t* apmask128, uint8_t* found,
                                     uint8_t* msg77, int32_t* iters, int32_t* nhard, int nthreads);
void     orc_ldpc_encode(const uint8_t* msg77, uint8_t* cw128);
void     orc_ldpc_parity_matrix(uint8_t* h /*38*128*/);

/* ---- whole block / whole stream (main.cu:261-425,428-532) ---- */
/* Runs scan + softbits + gate + LDPC on one filtered block.  items: nf*depth*8. */
void orc_decode_block(const float* blk_re_im, const orc_config* cfg, orc_item* items);

/* Host post-processing of one block's items: decode_message filter, unpack77, ResultFilter.
 * Returns number of lines written (<= cap). */
int orc_block_lines(const orc_item* items, int nitems, const orc_config* cfg, int snr,
                    int64_t block_index, orc_line* lines, int cap);

/* Whole stream: bytes exactly as they arrive on stdin.  Returns number of lines (may exceed cap;
 * only cap are written).  nblocks_out = number of blocks processed.  max_blocks <= 0: all. */
int64_t orc_decode_stream(const void* bytes, size_t nbytes, const orc_config* cfg,
                          int64_t first_block, int64_t max_blocks,
                          orc_line* lines, int64_t cap, int64_t* nblocks_out);

/* decode_softbits.cpp:23-55: 1 = found, text right-trimmed in text37 (NUL terminated, >=38 bytes) */
int orc_decode_message(const uint8_t* msg77, char* text37);

/* printed line, main.cu:406-419, date masked as 14 'x' when date==NULL */
int orc_format_line(const orc_line* line, const char* date14, char* out, int cap);

#ifdef __cplusplus
}
#endif
