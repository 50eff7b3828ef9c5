/*
 * msk144_b200.h -- C ABI of libmsk144b200.so, the B200-native MSK144 stream decoder.
 *
 * Drop-in boundary for the hot path of alexander-sholohov/msk144cudecoder ("the reference"):
 *   sample block -> analytic signal -> sync search -> soft bits -> LDPC(128,90) -> 77-bit payload -> text.
 * The reference has no library boundary of its own: its de-facto operator API is
 *   MSK144SearchContext ctx(center, width, step, depth, thr)          (msk_context.cuh:23)
 *   do_decode(ctx, device_block5184, ResultFilter&, SNRTracker&)      (main.cu:53,428-532)
 *   ResultKeeper::get_all_results() -> ResultItem[]                   (result_keeper.cuh:123)
 * inside the process loop of main.cu:261-425.  Every entry point below names the piece it replaces.
 *
 * Conventions: plain C, no exceptions cross the boundary; every function returns 0 on success or a
 * negative MSK144_E* code, and msk144_last_error() gives the text for the calling thread.  One
 * decoder owns one GPU and its streams; a decoder is used from one thread at a time.  There is NO
 * CPU fallback: creation fails if no sm_100-class device is usable.
 */
#ifndef MSK144_B200_H
#define MSK144_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MSK144_BLOCK_SAMPLES 5184 /* common.h:15  Num6x864 */
#define MSK144_HOP_SAMPLES 2592   /* main.cu:284-288 */
#define MSK144_NLLR 128           /* common.h:26 */
#define MSK144_NMSG 77            /* common.h:27 */
#define MSK144_NCAND 8            /* common.h:34 */

enum {
    MSK144_OK = 0,
    MSK144_EINVAL = -1,  /* bad argument / configuration */
    MSK144_ECUDA = -2,   /* CUDA runtime error (the reference throws std::runtime_error) */
    MSK144_ENODEV = -3,  /* no usable GPU */
    MSK144_ENOMEM = -4,
    MSK144_EOVERFLOW = -5, /* an output array was too small; counts are still returned */
    MSK144_EAGAIN = -6     /* more input bytes are needed to decide (msk144_wav_data_offset) */
};

/* flags */
enum {
    /* Reproduce the reference's decode_cache comparator quirk (main.cu:430-447): within one block every
     * candidate after the first decoded one is reported with the FIRST candidate's text.  The CLI sets
     * it (output identical to the reference); clear it for per-payload text. */
    MSK144_FLAG_COMPAT_DECODE_CACHE = 1,
    /* Streaming surface only: if the pushed stream starts with a RIFF/WAVE header, skip to the first byte of
     * its "data" chunk instead of decoding the header as samples (the reference does the latter: README.md:62
     * pipes demo/0001.wav straight in).  Off by default for compatibility; the CLI sets it for --wav. */
    MSK144_FLAG_SKIP_WAV_HEADER = 2,
    /* LDPC with the control flow of WSJT-X's bpdecode128_90 instead of the reference kernel's:
     * iter = 0..10 (11 parity checks, the reference has 10: common.h:29, ldpc_kernel.cuh:161), the first hard decision
     * that satisfies all parities and the CRC ends the decode (accepted if nharderror < 18), early stop when the number
     * of unsatisfied checks has not decreased for three checks running (iter >= 5, more than 10 unsatisfied), optional
     * a-priori mask (msk144_set_ap_mask).  Same arithmetic.  Off by default: the drop-in output is the reference's.
     * bpdecode128_90.f90 is not on this box; the mode is restated from the published algorithm (parity unpinned). */
    MSK144_FLAG_WSJTX_BP = 4,
    /* A live receiver's host thread sleeps while it waits for the GPU instead of spinning, as the reference's whole
     * process does (cudaDeviceScheduleBlockingSync, main.cu:115): the waits of a call that completes at most 8 blocks
     * (one 216 ms hop completes one) go through a blocking-sync event.  Calls that complete more (a file, a fast pipe)
     * keep spinning: there a sleeping wait costs throughput (6 % of the CLI's steady-state rate, measured).  The CLI
     * sets it. */
    MSK144_FLAG_BLOCKING_SYNC = 8
};

/* The seven CLI values of main.cu:124-133 + placement.  Zero-initialise, then msk144_config_defaults(). */
typedef struct msk144_config {
    float center_frequency_hz; /* --center-frequency; 1500 audio / 0 IQ (main.cu:193-208) */
    float search_step_hz;      /* --search-step, 2.0 */
    float search_width_hz;     /* --search-width, 200.0 */
    int32_t scan_depth;        /* --scan-depth, 4; clamped to [1,8] (msk_context.cuh:28-33) */
    int32_t read_mode;         /* --read-mode: 1 = s16 mono audio, 2 = s8 I/Q pairs (main.cu:79-83) */
    int32_t analytic_method;   /* --analytic-method: 1 = FFT, 2 = shift/LPF/shift (audio only) */
    int32_t nbadsync_threshold; /* --nbadsync-threshold, 1 */
    int32_t device;            /* CUDA device ordinal */
    int32_t max_batch_blocks;  /* blocks per launch group (0 = default 128) */
    int32_t flags;             /* MSK144_FLAG_* */
    int32_t reserved[6];
} msk144_config;

/* One printed decode line == ResultFilter::ResultItem (result_filter.h:17-29) + provenance. */
typedef struct msk144_line {
    int64_t block_index; /* block number within its channel, hop = 2592 samples */
    int32_t channel;
    int32_t snr;
    float f0;
    int32_t num_avg;
    int32_t nbadsync;
    int32_t pattern_idx;
    uint8_t payload[80]; /* 77 bits (0/1) of the candidate this line was printed for */
    char text[40];       /* NUL terminated, right-trimmed (decode_softbits.cpp:50-53) */
} msk144_line;

/* One LDPC success == the fields of ResultKeeper::ResultItem (result_keeper.cuh:16-30) that leave
 * the GPU.  item_index = (freq_idx * depth + pattern_idx) * 8 + slot (result_keeper.cuh:85-91). */
typedef struct msk144_result {
    int64_t block_index;
    int32_t channel;
    uint32_t item_index;
    uint32_t freq_idx;
    uint32_t pattern_idx;
    uint32_t slot;
    uint32_t pos;
    float f0;
    float xb;
    int32_t num_avg;
    int32_t nbadsync;
    int32_t ldpc_iters;
    int32_t ldpc_nhard;
    uint8_t message[80]; /* 77 bits, 0/1 */
} msk144_result;

/* One scan candidate (every (frequency, pattern, slot), decoded or not). */
typedef struct msk144_candidate {
    uint32_t pos;     /* 0..5375, as the reference reports it (scan_kernel.cuh:358-364) */
    float xb;
    int32_t nbadsync;
    int32_t llr_slot; /* index into the LLR queue if nbadsync <= threshold, else -1 */
} msk144_candidate;

typedef struct msk144_info {
    int32_t num_freqs;       /* NF, msk_context.cuh:102-108 */
    int32_t scan_depth;      /* effective depth */
    int32_t items_per_block; /* NF * depth * 8 */
    int32_t sm_count;
    float left_hz, right_hz; /* msk_context.cuh:127-129 */
    int32_t max_batch_blocks;
    int32_t reserved[9];
} msk144_info;

/* Per-kernel device time, CUDA events recorded on the launching stream around every kernel of every
 * launch group since the last msk144_get_stage_times() / decode call (milliseconds, summed). */
typedef struct msk144_stage_times {
    float frontend_ms, scan_ms, ldpc_ms, total_ms;
    int64_t blocks;        /* blocks in the group */
    int64_t llr_codewords; /* candidates that passed the nbadsync gate */
    int64_t ldpc_successes;
    int64_t kernel_launches; /* kernels launched by this library since creation */
    int64_t h2d_bytes;       /* host->device bytes copied by the library since creation */
    int64_t d2h_bytes;       /* device->host bytes copied by the library since creation */
    int64_t groups;          /* launch groups timed */
} msk144_stage_times;

typedef struct msk144_decoder msk144_decoder;

const char* msk144_last_error(void);
void msk144_config_defaults(msk144_config* cfg); /* code defaults of main.cu:124-133 */

/* MSK144SearchContext ctor + Analytic ctor + LDPCContext::init (main.cu:224-226). */
int msk144_create(const msk144_config* cfg, msk144_decoder** out);
void msk144_destroy(msk144_decoder* dec);
int msk144_get_info(const msk144_decoder* dec, msk144_info* info);
/* Synchronises the stream(s) used since the last call, sums the event pairs, resets the accumulation. */
int msk144_get_stage_times(msk144_decoder* dec, msk144_stage_times* t);

/* ---- streaming surface: the while(true) loop of main.cu:261-425 ------------------------------------
 * push raw stdin bytes (s16le mono, or s8 I,Q pairs); the library owns the half-block overlap, runs
 * every complete block, keeps the SNR recurrence, and queues decode lines in output order. */
int msk144_push(msk144_decoder* dec, const void* bytes, size_t nbytes);
/* The same loop for `nchan` channels arriving interleaved (one 2-byte sample of every channel per time step): one
 * overlap buffer and one SNR recurrence per channel; lines carry their channel and come out in time order (block, then
 * channel), however the caller chunks its pushes.  nchan is fixed until msk144_reset_stream().  Do not mix with
 * msk144_push on one decoder. */
int msk144_push_interleaved(msk144_decoder* dec, const void* bytes, size_t nbytes, int nchan);
int msk144_poll(msk144_decoder* dec, msk144_line* lines, size_t cap, size_t* nlines);
int msk144_reset_stream(msk144_decoder* dec);

/* ---- batch surface: nchan independent channels, each samples_per_channel samples, back to back.
 * Blocks of a channel start every 2592 samples; channels with fewer than 5184 samples yield nothing.
 * Each channel starts a fresh SNR recurrence.  lines / results may be NULL (cap 0).
 * `_host`: samples in host memory (pinned or pageable) -- H2D and D2H are inside the call.
 * `_device`: samples already resident in device memory; `stream` is a cudaStream_t (NULL = the
 * decoder's own stream).  Both return after the results are on the host. */
int msk144_decode_host(msk144_decoder* dec, const void* samples, size_t samples_per_channel, int nchan,
                       msk144_line* lines, size_t line_cap, size_t* nlines,
                       msk144_result* results, size_t result_cap, size_t* nresults);
int msk144_decode_device(msk144_decoder* dec, const void* d_samples, size_t samples_per_channel, int nchan,
                         void* stream, msk144_line* lines, size_t line_cap, size_t* nlines,
                         msk144_result* results, size_t result_cap, size_t* nresults);

/* The complete output of the decoder's most recent msk144_decode_* call, which it keeps until the next one: after
 * MSK144_EOVERFLOW (counts returned, arrays too small) call this with arrays of the reported sizes instead of decoding
 * again.  Either array may be NULL. */
int msk144_last_output(msk144_decoder* dec, msk144_line* lines, size_t line_cap, size_t* nlines,
                       msk144_result* results, size_t result_cap, size_t* nresults);

/* ---- ONE channel split in time (SURVEY.md section 8e "Partitioning"; no counterpart in the reference, which is one
 * process on one GPU).  The only things that cross blocks are on the host: the SNR recurrence (snr_tracker.cu:39-50,
 * fed by 8 sub-block powers per block) and the call-sign hash tables of the text decoder; on the device a block needs
 * nothing but its own 5184 samples, and consecutive blocks overlap by half (main.cu:284-288).  So a long stream is cut
 * into pieces of consecutive blocks, each piece runs the GPU stages on any decoder (any GPU, any process), and ONE
 * ordered host pass over the pieces' output prints exactly the lines of an undivided decode.
 * `_chunk_host`: the piece that starts at block `block_base` of its channel: samples[nsamples] begins at sample
 * 2592 x block_base of the channel and holds nb = (nsamples - 5184) / 2592 + 1 blocks (the piece of blocks [b0, b1) is
 * samples [2592 b0, 2592 (b1 - 1) + 5184)).  Returns every LDPC success (block_index = block_base + k, channel 0, in
 * (block, item) order) and powers8[nb][8]; no SNR, no text.  MSK144_EOVERFLOW: fetch with msk144_last_output. */
int msk144_decode_chunk_host(msk144_decoder* dec, const void* samples, size_t nsamples, int64_t block_base,
                             msk144_result* results, size_t result_cap, size_t* nresults, float* powers8,
                             size_t power_cap_blocks, int64_t* nblocks_out);
/* `_replay_channel`: main.cu:388-419 for blocks [first_block, first_block + nblocks) of one channel, from the start of
 * the channel's state: SNR recurrence over powers8[nblocks][8], then per block the i3/n3 gate, unpack77 (hash tables in
 * stream order) and ResultFilter over that block's results.  `results` = the pieces' results concatenated in block
 * order.  Host only, no GPU.  flags: MSK144_FLAG_COMPAT_DECODE_CACHE as for msk144_create; `channel` is copied into
 * the lines.  MSK144_EOVERFLOW: *nlines is the size needed, call again (the pass is cheap). */
int msk144_replay_channel(int32_t flags, int32_t channel, int64_t first_block, int64_t nblocks, const float* powers8,
                          const msk144_result* results, size_t nresults, msk144_line* lines, size_t line_cap,
                          size_t* nlines);

/* Multi-channel input as a channeliser / multi-receiver front end delivers it: one sample of every channel per time
 * step (samples[t][c], 2 bytes each).  The reference reads one channel from stdin (main.cu:261-298); this is the ingest
 * adapter of SURVEY.md section 8f.  `_device`: transposes d_in [spc][nchan] into d_out [nchan][spc] (the layout of
 * msk144_decode_device) on `stream`; d_in and d_out must not overlap.  `_host_interleaved`: H2D, transpose, decode. */
int msk144_deinterleave_device(msk144_decoder* dec, const void* d_in, void* d_out, size_t samples_per_channel, int nchan,
                               void* stream);
int msk144_decode_host_interleaved(msk144_decoder* dec, const void* samples, size_t samples_per_channel, int nchan,
                                   msk144_line* lines, size_t line_cap, size_t* nlines,
                                   msk144_result* results, size_t result_cap, size_t* nresults);

/* Device-only pass for measurement: runs the kernels on `stream` over resident input and leaves the
 * results on the device; no host synchronisation, no D2H.  *nblocks_out = blocks processed. */
int msk144_run_device(msk144_decoder* dec, const void* d_samples, size_t samples_per_channel, int nchan,
                      void* stream, int64_t* nblocks_out);

/* ---- stage-level entry points (tests, micro-benchmarks) ---------------------------------------------- */
/* main.cu:300-332,365-377 + analytic2.cuh / analytic_fft.cu: raw samples of `nblocks` consecutive
 * blocks of ONE channel -> filtered complex blocks [nblocks][5184][2] and SNR sub-block powers
 * [nblocks][8] (snr_tracker.cu:23-32). */
int msk144_frontend_host(msk144_decoder* dec, const void* samples, size_t nsamples, float* blocks_re_im,
                         float* powers8, int64_t* nblocks_out);
/* scan_kernel + softbits_kernel + index_kernel on already filtered blocks (host, [nblocks][5184][2]).
 * cands: [nblocks][items_per_block]; llr: [llr_cap][128] compacted in arbitrary order, llr_item[i] =
 * block * items_per_block + item_index of row i. */
int msk144_scan_host(msk144_decoder* dec, const float* blocks_re_im, int64_t nblocks, msk144_candidate* cands,
                     float* llr, uint32_t* llr_item, size_t llr_cap, size_t* nllr);
/* ldpc_kernel on caller-supplied LLRs [ncw][128] (BASELINE config 4). found/iters/nhard: [ncw];
 * msg77: [ncw][77] of 0/1. */
int msk144_ldpc_decode_host(msk144_decoder* dec, const float* llr, size_t ncw, uint8_t* found, uint8_t* msg77,
                            int32_t* iters, int32_t* nhard);
/* Same on device-resident LLRs, asynchronous on `stream`; d_out: [ncw] x 4 uint32 words
 * {bits 0..31, 32..63, 64..76 | found<<31, iters | nhard<<8}. */
int msk144_ldpc_decode_device(msk144_decoder* dec, const float* d_llr, size_t ncw, uint32_t* d_out, void* stream);

/* MSK144_FLAG_WSJTX_BP only: bits with apmask128[b] != 0 keep zn = llr in every iteration (bpdecode128_90's apmask;
 * the caller puts its a-priori knowledge into their LLRs).  NULL clears the mask.  Applies to every later decode. */
int msk144_set_ap_mask(msk144_decoder* dec, const uint8_t* apmask128);

/* Device self-test of the LDPC kernel's packed math: all 2^32 float inputs through the f32x2 tanh and the
 * division-free platanh against libdevice tanhf and the reference's platanh (ldpc_kernel.cuh:65-93) with an
 * IEEE division.  Counts of mismatching inputs; both must be 0 for bit-exact LDPC parity. */
int msk144_selftest_ldpc_math(msk144_decoder* dec, uint64_t* tanh_mismatches, uint64_t* platanh_mismatches);

/* ---- host post-processing, exported for the CLI and for callers that bring their own GPU stage ---- */
/* decode_softbits.cpp:23-55 (i3/n3 gate + unpack77 + right trim). 1 = found. text: >= 38 bytes. */
int msk144_decode_message(const uint8_t* msg77, char* text);
/* main.cu:406-419 line format; date14 = "%Y%m%d%H%M%S" local time, NULL = now. Returns length. */
int msk144_format_line(const msk144_line* line, const char* date14, char* out, size_t cap);


/* ---- adapters either side of the path (no counterpart in the reference) -------- */
/* Pluggable text seam == f_interop.cpp:61-71: the reference resolves `__packjt77_MOD_unpack77` of a gfortran-built
 * WSJT-X at link time (bpdecode.dll via LoadLibrary on Win32, f_interop.cpp:40-53).  Here the built-in C++ unpack77 is
 * the default; this call (or the environment variable MSK144_UNPACK77_LIB, read at the first text decode) dlopen()s a
 * user-supplied shared library exporting that symbol and routes every text decode through it.  NULL = built-in. */
int msk144_set_unpack77_library(const char* path);
/* 1 if an external unpack77 is active. */
int msk144_unpack77_is_external(void);

/* RIFF/WAVE header recognition on the first bytes of a stream.  Returns the byte offset of the first sample of the
 * "data" chunk (> 0), 0 if the bytes are not a RIFF file (raw samples), MSK144_EAGAIN if the header is not complete
 * within `n` bytes, MSK144_EINVAL for a RIFF file that is not PCM.  channels / sample_rate / bits may be NULL. */
int64_t msk144_wav_data_offset(const void* bytes, size_t n, int32_t* channels, int32_t* sample_rate, int32_t* bits);

/* One decode line as a JSON object on one line (machine-readable twin of msk144_format_line). Returns length. */
int msk144_format_line_json(const msk144_line* line, const char* date14, char* out, size_t cap);

/* Cross-block aggregation: the reference prints a message once per 216 ms block it decodes in (blocks overlap by
 * half, main.cu:284-288), so one ping yields several lines.  An event merges the lines of one (channel, text) whose
 * blocks are at most max_gap_blocks apart and carries the best line (lowest num_avg, then nbadsync, then highest snr). */
typedef struct msk144_event {
    int64_t first_block, last_block;
    int32_t nlines;
    int32_t reserved;
    msk144_line best;
} msk144_event;
typedef struct msk144_aggregator msk144_aggregator;
int msk144_aggregator_create(int32_t max_gap_blocks, msk144_aggregator** out);
void msk144_aggregator_destroy(msk144_aggregator* agg);
/* lines of each channel in block order (as msk144_poll / msk144_decode_* deliver them) */
int msk144_aggregator_add(msk144_aggregator* agg, const msk144_line* lines, size_t n);
/* events that can no longer grow: last_block + max_gap_blocks < next_block (the index of the next block the
 * stream will produce), or all of them if flush != 0 */
int msk144_aggregator_poll(msk144_aggregator* agg, int64_t next_block, int flush, msk144_event* out, size_t cap,
                           size_t* nout);

/* ---- several GPUs of one box (channels are independent, no exchange step) -------------
 * The reference is one process, one GPU, one channel (main.cu:98-115,261-298).  A pool owns one decoder per listed
 * device; channels are split into contiguous, balanced ranges (the first nchan % ndevices members get one more),
 * every member runs on its own host thread with its own CUDA streams, and the decode lines are merged on the host.
 * Output (content and order) equals what a single decoder returns for the same input.  A device may be listed more
 * than once (two decoders sharing a GPU).  Errors: msk144_pool_last_error() for the calling thread. */
typedef struct msk144_pool msk144_pool;
const char* msk144_pool_last_error(void);
int msk144_pool_create(const msk144_config* cfg, const int32_t* devices, int32_t ndevices, msk144_pool** out);
void msk144_pool_destroy(msk144_pool* pool);
int msk144_pool_size(const msk144_pool* pool);
msk144_decoder* msk144_pool_member(msk144_pool* pool, int32_t i); /* for msk144_get_stage_times etc. */
/* == msk144_decode_host over all members.  With fewer channels than members the channels are ALSO split in time: each
 * channel gets a balanced share of the members, every one of them runs msk144_decode_chunk_host on a contiguous range
 * of the channel's blocks, and the pool runs msk144_replay_channel once per channel; same lines, same order. */
int msk144_pool_decode_host(msk144_pool* pool, const void* samples, size_t samples_per_channel, int nchan,
                            msk144_line* lines, size_t line_cap, size_t* nlines,
                            msk144_result* results, size_t result_cap, size_t* nresults);
/* == msk144_last_output: the merged output of the pool's most recent msk144_pool_decode_host */
int msk144_pool_last_output(msk144_pool* pool, msk144_line* lines, size_t line_cap, size_t* nlines,
                            msk144_result* results, size_t result_cap, size_t* nresults);
/* == msk144_push_interleaved / msk144_poll / msk144_reset_stream over all members; nbytes must be a whole number of
 * time steps (2 * nchan bytes each) */
int msk144_pool_push_interleaved(msk144_pool* pool, const void* bytes, size_t nbytes, int nchan);
int msk144_pool_poll(msk144_pool* pool, msk144_line* lines, size_t cap, size_t* nlines);
int msk144_pool_reset_stream(msk144_pool* pool);

/* ---- channeliser in front of the decoder (ingest side) ---------------------------------------
 * The reference reads ONE channel that something else has already brought to 12 ksps; its README documents the chain
 *   rtl_sdr -s 1920000 ... | csdr convert_u8_f | fir_decimate_cc 8 | 5 | 4 | gain_ff 100 | convert_f_s8   (README.md:75-78)
 * This does that step on the GPU for any number of channels of one wideband capture at once: a uniform polyphase
 * filter bank with decimation M = input rate / 12000.  Channel c is centred bins[c] x 12 kHz from the centre of the
 * capture (|bin| < M, negative = below) and leaves as s8 I,Q pairs at 12 ksps, the input format of --read-mode=2:
 *   out = saturate_s8(round_to_nearest_even(127 x gain x y)),  y = low-passed sample relative to full scale
 *   (full scale: u8 (v - 127.5) / 127.5, s8 v / 128, s16 v / 32768; saturation at +-127).
 * Prototype low-pass: 8 M taps, Kaiser beta 7.857, -6 dB at 6 kHz, unit DC gain (msk144_channelizer_prototype returns it).
 * csdr itself is not on this box: the filter is this library's own design, not csdr's.  No CPU fallback. */
enum { MSK144_IQ_U8 = 1, MSK144_IQ_S8 = 2, MSK144_IQ_S16 = 3 };
typedef struct msk144_channelizer_config {
    int32_t input_format; /* MSK144_IQ_*: interleaved I,Q */
    int32_t decimation;   /* M, 2..256: input rate = M x 12000 samples/s (160 for rtl_sdr -s 1920000) */
    float gain;           /* csdr gain_ff; 100 in the README's recipe */
    int32_t device;       /* CUDA device ordinal */
    int32_t reserved[4];
} msk144_channelizer_config;
typedef struct msk144_channelizer msk144_channelizer;
int msk144_channelizer_create(const msk144_channelizer_config* cfg, const int32_t* bins, int32_t nbins, msk144_channelizer** out);
void msk144_channelizer_destroy(msk144_channelizer* chz);
int msk144_channelizer_prototype(const msk144_channelizer* chz, float* taps, size_t cap, size_t* ntaps);
int64_t msk144_channelizer_launches(const msk144_channelizer* chz); /* kernels launched since creation */
/* Resident input -> resident output, asynchronous on `stream` (cudaStream_t, NULL = the channeliser's own).  d_in points
 * at the first sample of frame 0 (a frame = M input samples = one output sample per channel); `history_frames` (0..7)
 * whole frames BEFORE d_in are valid and are used as filter history, anything earlier counts as zero.  Output sample
 * (channel c, frame f) is the byte pair at d_out + 2 (c x out_chan_stride + f x out_frame_stride): strides (nframes
 * rounded up to 8, 1) give the channel-major layout of msk144_decode_device, (1, nchan) the time-major layout of
 * msk144_push_interleaved. */
int msk144_channelize_device(msk144_channelizer* chz, const void* d_in, size_t nframes, int32_t history_frames, void* d_out,
                             size_t out_chan_stride, size_t out_frame_stride, void* stream);
/* One capture in host memory -> out[nchan][*nframes][2], *nframes = whole frames in nbytes; zero history. */
int msk144_channelize_host(msk144_channelizer* chz, const void* in, size_t nbytes, int8_t* out, size_t out_cap_frames, size_t* nframes);
/* Streaming: push any number of bytes; whole frames come back time-major, out[*nframes][nchan][2] (feed them to
 * msk144_push_interleaved, or msk144_push for one channel); the filter history and a partial frame are kept inside.
 * MSK144_EOVERFLOW (nothing consumed, *nframes = frames needed) if out_cap_frames is too small. */
int msk144_channelizer_push(msk144_channelizer* chz, const void* bytes, size_t nbytes, int8_t* out, size_t out_cap_frames, size_t* nframes);
int msk144_channelizer_reset(msk144_channelizer* chz);

/* ---- front end without any allocation --------------------------------------------------------------------------
 * No counterpart in the reference (a stand-alone process that allocates its block buffers itself: main.cu:214-216).
 * For code that holds device buffers but is not allowed to allocate at the point where it wants the front end run:
 * a plugin running inside a host application that owns the whole memory budget (a torch caching allocator, an SDR
 * framework's arena), a real-time receive loop in which cudaMalloc / cudaFree would stall every stream of the device
 * (both synchronise it), or a process that has deliberately handed all free HBM to something else.  Every other entry
 * point of this library needs a decoder, and msk144_create allocates.  This one needs none.
 *
 * It runs the method-2 front end (--analytic-method=2 for s16 audio, the IQ low-pass for --read-mode=2) of ONE channel:
 * d_samples[nsamples] -> d_blocks_re_im[nblocks][5184][2] and d_powers8[nblocks][8], nblocks = (nsamples - 5184) / 2592 + 1
 * (0 if nsamples < 5184); the same kernels and the same bits as msk144_frontend_host.  All three pointers are DEVICE
 * memory of the CURRENT device (cudaSetDevice is the caller's), supplied and sized by the caller; `stream` is a
 * cudaStream_t (NULL = the default stream); the call is asynchronous and returns once the kernels are queued.
 *
 * The one thing the pipeline allocates for this stage, 1 / rms per block between its two kernels, lives in a fixed
 * array inside the library's device image instead: 4 bytes x MSK144_NOALLOC_MAX_BLOCKS = 32 KB on every GPU that
 * loads the library, held until the process exits.  Consequences:
 *   - a call with more than MSK144_NOALLOC_MAX_BLOCKS blocks (nsamples > 2592 x 8193) is refused with MSK144_EINVAL;
 *     cut longer input into pieces that overlap by 2592 samples;
 *   - every call on a GPU uses the same array.  Calls that follow each other on one stream are fine (each call
 *     rewrites every entry it reads; nothing is carried over and nothing needs clearing), but two calls that can run
 *     on the GPU at the same time (different streams, different host threads) race on it and must be serialised by the caller;
 *   - the first call on each device looks up the array's device address and raises the kernel's shared-memory limit;
 *     make that first call before starting any stream capture on that device.
 * No cudaMalloc, cudaFree, cudaMallocHost or other allocating routine is called, on any path. */
#define MSK144_NOALLOC_MAX_BLOCKS 8192
int msk144_frontend_device_noalloc(const void* d_samples, size_t nsamples, int read_mode, float* d_blocks_re_im,
                                   float* d_powers8, void* stream, int64_t* nblocks_out);

/* ---- measurement helpers -------------------------------------------------------------------------------------- */
/* Sustained FP32 FMA rate of the decoder's device: a dependency-free FFMA stream (8 independent chains per thread,
 * every SM sub-partition saturated), timed with CUDA events; *tflops = 2 * FMAs / time.  This is the measured
 * denominator of bench.py's FP32 roofline (MEASURED_PEAKS.json has no FP32 number). */
int msk144_measure_fp32_peak(msk144_decoder* dec, double* tflops);
/* "nvcc <major>.<minor>.<build>; sm_100a; <date>": the LDPC kernel's tanh reproduces libdevice of this toolkit
 * (msk144_selftest_ldpc_math checks it on the device). */
const char* msk144_build_info(void);

#ifdef __cplusplus
}
#endif
#endif /* MSK144_B200_H */
