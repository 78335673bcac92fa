/*
 * tmkphy.h — C ABI of libtmkphy, the B200-native batched acoustic-modem PHY engine.
 *
 * Drop-in boundary for the PHY hot path of HeZeBang/TrackMaker-rs (reference paths are
 * relative to the reference repository root):
 *
 *   tmk_modulate        replaces PhyEncoder::encode_frames      src/phy/encoder.rs:70-92
 *                       (Frame::to_bytes src/phy/frame.rs:74-102,
 *                        LineCode::encode src/phy/line_coding.rs:74-91,202-240)
 *   tmk_decode_batch    replaces PhyDecoder::process_samples    src/phy/decoder.rs:68-112
 *                       (search_for_preamble :123-238, decode_frame :242-356,
 *                        LineCode::decode line_coding.rs:93-123,243-293,
 *                        Frame::from_bits frame.rs:147-183, crc8 crc.rs:7-22)
 *   tmk_correlate       parity probe for the normalised sliding correlation
 *                                                              src/phy/decoder.rs:130-147
 *   tmk_channel_busy    replaces mac::is_channel_busy           src/mac/mod.rs:18-27
 *   tmk_phy_config      mirrors PhyDecoder::new / PhyEncoder::new arguments
 *                                                              src/phy/decoder.rs:36-65
 *   tmk_frame_desc      mirrors struct Frame                    src/phy/frame.rs:34-41
 *
 * Conventions: every call returns 0 (TMK_OK) or a negative error code; nothing unwinds
 * across the boundary; all buffers are caller-allocated and caller-owned; sample / record /
 * payload pointers may be HOST or DEVICE pointers (detected with cudaPointerGetAttributes);
 * one engine per GPU, driven by one host thread at a time (like a PhyDecoder: Send, !Sync).
 * There is no CPU fallback: every entry point that computes runs CUDA kernels on sm_100a.
 */
#ifndef TMKPHY_H
#define TMKPHY_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TMK_ABI_VERSION 1

/* LineCodingKind (src/phy/line_coding.rs:26-30) */
#define TMK_LINE_MANCHESTER 0
#define TMK_LINE_4B5B 1

/* FrameType (src/phy/frame.rs:12-16) */
#define TMK_FRAME_TYPE_DATA 0x01
#define TMK_FRAME_TYPE_ACK 0x02

/* return codes */
#define TMK_OK 0
#define TMK_E_INVALID (-1)  /* bad argument (null pointer, bad config, misaligned stream) */
#define TMK_E_CUDA (-2)     /* CUDA runtime error; see tmk_last_error */
#define TMK_E_CAPACITY (-3) /* an output buffer was too small; counts are still returned */
#define TMK_E_NOMEM (-4)

/* per-candidate verdict: one record per preamble detection, in the order the reference
 * decoder would have met them (src/phy/decoder.rs:266-355). Only TMK_FRAME_OK records
 * correspond to Frames returned by PhyDecoder::process_samples. */
#define TMK_FRAME_OK 0
#define TMK_FRAME_HDR_PARSE 1    /* decoder.rs:266-273  consumes header_samples */
#define TMK_FRAME_BAD_LEN 2      /* decoder.rs:277-286  consumes 1 sample */
#define TMK_FRAME_SHORT_DECODE 3 /* decoder.rs:311-320  4B5B invalid symbol mid-frame */
#define TMK_FRAME_NOT_FOR_US 4   /* decoder.rs:322-329  dst != local_addr */
#define TMK_FRAME_CRC_FAIL 5     /* decoder.rs:347-354 */
#define TMK_FRAME_TRUNCATED 6    /* capture ended while the decoder was waiting for samples
                                    (decoder.rs:252-254,295-297 return None forever) */

/* Reference defaults: src/utils/consts.rs:24-48 and src/phy/decoder.rs:57,61 */
#define TMK_DEFAULT_SPL 3
#define TMK_DEFAULT_PREAMBLE_BYTES 2
#define TMK_DEFAULT_THRESHOLD 0.9f
#define TMK_DEFAULT_MAX_FRAME_BYTES 256
#define TMK_DEFAULT_GAP_SAMPLES 48
#define TMK_PHY_HEADER_BYTES 7

typedef struct tmk_engine tmk_engine;

typedef struct tmk_phy_config {
    int32_t line_code;        /* TMK_LINE_* */
    uint32_t spl;             /* samples per level (3) */
    uint32_t preamble_bytes;  /* 2 */
    int32_t local_addr;       /* 0..255; -1 accepts every dst (extension) */
    float threshold;          /* 0.9f */
    uint32_t max_frame_bytes; /* 256 */
    uint64_t chunk;           /* size of each simulated process_samples() call in samples;
                                 0 = the whole capture in one call (src/main.rs:540).
                                 The live receiver polls 1200 samples (src/mac/csma.rs:446). */
} tmk_phy_config;

/* one capture of a ragged batch; `offset` is in samples from the `samples` base pointer and
 * must be a multiple of 4 (16-byte aligned float4 loads). */
typedef struct tmk_stream {
    uint64_t offset;
    uint64_t n_samples;
    tmk_phy_config cfg;
} tmk_stream;

typedef struct tmk_frame_rec {
    uint32_t stream;         /* index into the streams array */
    uint32_t status;         /* TMK_FRAME_* */
    uint64_t preamble_start; /* absolute sample index of the first window with corr >= thr */
    uint64_t frame_start;    /* first header sample after sync-word refinement */
    uint64_t payload_off;    /* byte offset into the payload buffer (status OK / CRC_FAIL /
                                NOT_FOR_US: `len` decoded payload bytes are stored there) */
    float corr;              /* correlation of the detecting window, computed with the accurately summed window energy:
                                within 1e-5 relative of the reference's value (which uses its drifting rolling energy);
                                not part of the bit-exact contract — tmk_correlate is the bit-exact probe */
    uint32_t consumed;       /* samples consumed from preamble_start */
    uint16_t len;            /* header fields (zero when fewer than 7 header bytes decoded) */
    uint8_t crc;
    uint8_t type;
    uint8_t seq;
    uint8_t src;
    uint8_t dst;
    uint8_t hdr_valid;       /* 1 if the header fields above were decoded */
} tmk_frame_rec;

/* struct Frame (src/phy/frame.rs:34-41) for the modulator */
typedef struct tmk_frame_desc {
    uint64_t payload_off; /* into the payload buffer */
    uint32_t len;         /* payload bytes, <= 65535 */
    uint8_t type;
    uint8_t seq;
    uint8_t src;
    uint8_t dst;
} tmk_frame_desc;

/* counters returned by tmk_stats (accumulated since engine creation / tmk_stats_reset) */
enum {
    TMK_STAT_SAMPLES = 0,
    TMK_STAT_DETECTIONS = 1,
    TMK_STAT_FRAMES_OK = 2,
    TMK_STAT_CRC_FAIL = 3,
    TMK_STAT_HDR_FAIL = 4, /* HDR_PARSE + BAD_LEN + SHORT_DECODE */
    TMK_STAT_NOT_FOR_US = 5,
    TMK_STAT_EXACT_REPLAYS = 6, /* threshold decisions that needed the sequential replay */
    TMK_STAT_CANDIDATES = 7,
    TMK_STAT_COUNT = 8
};

int tmk_abi_version(void);
int tmk_engine_create(int device, tmk_engine** out);
void tmk_engine_destroy(tmk_engine* e);
const char* tmk_last_error(const tmk_engine* e);
void tmk_default_config(int32_t line_code, int32_t local_addr, tmk_phy_config* out);

/* LineCode::samples_for_bits / preamble length (line_coding.rs:125-127,295-300,13-21) */
uint64_t tmk_samples_for_bits(const tmk_phy_config* cfg, uint64_t n_bits);
uint64_t tmk_preamble_len(const tmk_phy_config* cfg);
/* number of samples tmk_modulate will produce */
uint64_t tmk_modulated_len(const tmk_phy_config* cfg, const tmk_frame_desc* frames, uint32_t n_frames,
                           uint32_t gap_samples);

int tmk_modulate(tmk_engine* e, const tmk_phy_config* cfg, const tmk_frame_desc* frames, uint32_t n_frames,
                 const uint8_t* payload, uint64_t payload_len, uint32_t gap_samples, float* out_samples,
                 uint64_t out_cap, uint64_t* out_len);

int tmk_decode_batch(tmk_engine* e, const float* samples, uint64_t total_samples, const tmk_stream* streams,
                     uint32_t n_streams, tmk_frame_rec* recs, uint64_t rec_cap, uint64_t* n_recs,
                     uint8_t* payload, uint64_t payload_cap, uint64_t* payload_len);

/* corr_out[offset + i] = correlation of window i of each stream (n_samples - P + 1 windows),
 * one-shot semantics: window energy seeded at window 0 and rolled (decoder.rs:133-136,224-233),
 * AVX-order dot product; bit-exact with the reference arithmetic (sequential per stream: a
 * parity probe, not the hot path). */
int tmk_correlate(tmk_engine* e, const float* samples, uint64_t total_samples, const tmk_stream* streams,
                  uint32_t n_streams, float* corr_out);

/* Diagnostic: the raw output of the preamble scan kernel (the candidate filter in front of the walk).
 * Blocks are TMK_SCAN_BLOCK windows; stream i owns blocks blk_base[i] .. blk_base[i] + ceil(nwin_i / block) - 1.
 * blk_info: 4 floats per block (min window energy, max window energy, max corr^2, #candidates);
 * cand_bits: tmk_scan_block() / 32 words per block, bit w = window w of the block passed
 * corr >= threshold * 0.85 / 0.9. Host pointers only. */
#define TMK_SCAN_BLOCK 256
uint32_t tmk_scan_block(void); /* the block size this library was built with (TMK_SCAN_BLOCK; 480 in the A/B build of the round-1 scan) */
int tmk_scan_probe(tmk_engine* e, const float* samples, uint64_t total_samples, const tmk_stream* streams,
                   uint32_t n_streams, float* blk_info, uint32_t* cand_bits, uint64_t block_cap, uint64_t* n_blocks,
                   uint64_t* blk_base);

/* Diagnostic: the correlation numerators D(w) = sum_t x[w + t] p[t] of the scan filter for the shipped shapes
 * (TF32 tensor-core products; within 2^-10 sum |x| of the exact value), dot_out[stream offset + w]. Host output. */
int tmk_scan_dot_probe(tmk_engine* e, const float* samples, uint64_t total_samples, const tmk_stream* streams,
                       uint32_t n_streams, float* dot_out);

/* Diagnostic: the exact rolling window energy E_ref (decoder.rs:133-136, 224-233) of the windows[i] (ascending,
 * >= seed) for a search_for_preamble call entered at `seed`, through the very replay the walk uses when a
 * decision is too close for its drift bound (lane-parallel rounds + literal rounds). Host pointers for windows /
 * energy. counts: [0] steps rolled literally, [1] steps covered by lane-parallel rounds, [2] segments of those
 * rounds that had to be rolled literally. */
int tmk_replay_probe(tmk_engine* e, const float* samples, uint64_t n_samples, const tmk_phy_config* cfg, uint64_t seed,
                     const uint64_t* windows, uint32_t n_windows, float* energy, uint64_t counts[3]);

/* busy[i] = -1 (fewer than 20 samples), 0 or 1 : mac::is_channel_busy per stream */
int tmk_channel_busy(tmk_engine* e, const float* samples, uint64_t total_samples, const tmk_stream* streams,
                     uint32_t n_streams, int8_t* busy);

/* The same result for callers that may not allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc). No counterpart in the reference.
 * All device scratch (stream descriptors, one flag per stream, a staging row for host results) lives in
 * one static __device__ array of the library, sized for TMK_NOALLOC_MAX_STREAMS streams (84 KB per GPU, held until
 * the process exits); the call issues no cudaMalloc / cudaFree / cudaMallocHost. `samples` must be DEVICE memory
 * (staging a host buffer would allocate); `streams` is a host array; `busy` may be host or device. More than
 * TMK_NOALLOC_MAX_STREAMS streams -> TMK_E_CAPACITY (nothing is computed). Every call on a GPU shares the array,
 * which keeps the last call's contents (the call clears what it needs at zero): calls that overlapped on the GPU
 * would race on it, so the entry point serialises its callers and returns only when its stream has drained. */
#define TMK_NOALLOC_MAX_STREAMS 4096
int tmk_channel_busy_noalloc(tmk_engine* e, const float* samples, uint64_t total_samples, const tmk_stream* streams,
                             uint32_t n_streams, int8_t* busy);

int tmk_stats(tmk_engine* e, uint64_t counters[TMK_STAT_COUNT]);
int tmk_stats_reset(tmk_engine* e);
/* Diagnostic counters of the walk kernel of the last tmk_decode_batch call: [0] blocks examined window by
 * window, [1] exact re-anchorings of the rolling-energy bound, [2] sequential replay steps,
 * [3] groups of 32 block summaries read, [4] device time of the cluster pass in microseconds (single-chunk
 * calls), [5] rolling steps covered by the lane-parallel exact replay, [6] its segments that had to be rolled
 * literally; [7] reserved. ([2] counts literally rolled steps only.) */
int tmk_debug_counters(tmk_engine* e, uint64_t out[8]);

/* plumbing for callers without a CUDA runtime of their own (tests, bench, the Rust shim) */
int tmk_device_alloc(tmk_engine* e, uint64_t bytes, void** out);
int tmk_device_free(tmk_engine* e, void* p);
int tmk_host_alloc_pinned(tmk_engine* e, uint64_t bytes, void** out);
int tmk_host_free_pinned(tmk_engine* e, void* p);
int tmk_memcpy_h2d(tmk_engine* e, void* dst, const void* src, uint64_t bytes);
int tmk_memcpy_d2h(tmk_engine* e, void* dst, const void* src, uint64_t bytes);
int tmk_synchronize(tmk_engine* e);
void* tmk_engine_stream(tmk_engine* e); /* the cudaStream_t every kernel is launched on */
/* device time of the kernels of the last tmk_decode_batch call (CUDA events), ms */
int tmk_last_kernel_ms(tmk_engine* e, float* ms);
/* device time of the stages of the last tmk_decode_batch call, ms:
 * [0] preamble scan kernel(s), [1] walk kernel, [2] record ordering, [3] host->device staging of
 * the samples (0 when they were already resident) */
int tmk_last_stage_ms(tmk_engine* e, float ms[4]);
/* CUDA-event stopwatch on the engine's stream (for callers without a CUDA runtime of their own) */
int tmk_timer_start(tmk_engine* e);
int tmk_timer_stop(tmk_engine* e, float* ms);
/* PCI bus id of the engine's GPU ("0000:1b:00.0"), so that a host thread can bind itself to the GPU's NUMA node before it
 * allocates pinned capture buffers (/sys/bus/pci/devices/<id>/local_cpulist) */
int tmk_device_pci_bus_id(tmk_engine* e, char* out, int cap);
/* how many kernels this engine has launched so far */
uint64_t tmk_launch_count(const tmk_engine* e);

/* Synthetic capture generator (measurement harness): fills `samples` (device or
 * host) with n_streams captures of n_per_stream samples each — random frames (payload
 * 1..128 B, 10 % ACKs, 10 % foreign dst), random gaps, AWGN of the given SNR — generated on
 * the GPU from counter-based hashes of (seed, stream). Ground truth (what was sent) is written
 * to truth records (status = OK, preamble_start, len, seq, dst ...) when truth != NULL. */
int tmk_synth_batch(tmk_engine* e, const tmk_phy_config* cfg, uint64_t seed, uint32_t n_streams,
                    uint64_t n_per_stream, float snr_db, float amplitude, float* samples,
                    tmk_frame_rec* truth, uint64_t truth_cap, uint64_t* n_truth);
/* The same generator with the impairments of SURVEY §8d (BASELINE configs 3 and 4):
 *   multipath   3-tap FIR h = [1, 0.35 @ t1, -0.2 @ t2] renormalised to unit energy, t1 in U{2..12}, t2 in U{13..40}
 *               samples per stream
 *   drift_ppm   clock drift: linear-interpolation resample by (1 + delta), delta in U[-drift_ppm, +drift_ppm] per
 *               stream (truth positions are then rounded capture positions)
 *   amp_min     > 0: the whole capture (frames and noise) of a stream is scaled by U[amp_min, 1]
 *   gap_min/max inter-frame gaps U[gap_min, gap_max) samples (both 0: U[48, 4800)); sparse plans for hour captures
 * All zero = tmk_synth_batch. */
typedef struct tmk_synth_params {
    float snr_db;      /* >= 100: no noise */
    float amplitude;
    uint32_t multipath;
    float drift_ppm;
    uint32_t gap_min, gap_max;
    float amp_min;
    uint32_t _pad;
} tmk_synth_params;
int tmk_synth_batch_ex(tmk_engine* e, const tmk_phy_config* cfg, uint64_t seed, uint32_t n_streams,
                       uint64_t n_per_stream, const tmk_synth_params* prm, float* samples,
                       tmk_frame_rec* truth, uint64_t truth_cap, uint64_t* n_truth);

/* ------------------------------------------------------------------------------------------------
 * Links: n persistent PhyDecoder instances fed with process_samples calls (the live cadence,
 * src/mac/csma.rs:446-465 polls 25 ms = 1200 samples; src/mac/acoustic_interface.rs:334-353). All links of
 * a call are decoded in ONE batch on the GPU. Carried per link, exactly as decoder.rs:16-33 does:
 * the sample buffer drained to P-1 samples before buffer_offset (:90-109), buffer_offset, and the state
 * Searching | Decoding(frame_start).
 *   tmk_links_create           replaces n x PhyDecoder::new        src/phy/decoder.rs:36-65 (cfg.chunk is ignored)
 *   tmk_links_process_samples  replaces n x process_samples        src/phy/decoder.rs:68-112
 *   tmk_links_reset            replaces reset                      src/phy/decoder.rs:114-119
 */
typedef struct tmk_links tmk_links;
int tmk_links_create(tmk_engine* e, const tmk_phy_config* cfgs, uint32_t n_links, tmk_links** out);
void tmk_links_destroy(tmk_links* l);
/* link i receives samples[offsets[i] .. offsets[i] + lens[i]) (host or device memory; 0 = empty slice).
 * Returns the records of the frames whose decode finished in this call (host memory), sorted by
 * (link, order of detection): `stream` is the link index, preamble_start / frame_start count from the
 * first sample ever fed to the link; status == TMK_FRAME_OK records are the Frames process_samples
 * returns. A frame still being gathered is carried state, not a record. On an error (TMK_E_CAPACITY
 * included) nothing was consumed and the call may be repeated. */
int tmk_links_process_samples(tmk_links* l, const float* samples, uint64_t total_samples, const uint64_t* offsets,
                              const uint64_t* lens, tmk_frame_rec* recs, uint64_t rec_cap, uint64_t* n_recs,
                              uint8_t* payload, uint64_t payload_cap, uint64_t* payload_len);
/* link == UINT32_MAX resets every link */
int tmk_links_reset(tmk_links* l, uint32_t link);
/* sample_buffer.len(), buffer_offset and (state == Decoding) of one link */
int tmk_links_state(const tmk_links* l, uint32_t link, uint64_t* buffered, uint64_t* buffer_offset, int* decoding);

/* ------------------------------------------------------------------------------------------------
 * Chirp + BPSK modem of examples/ask.rs (project 1): 440-sample chirp preamble 2 kHz -> 10 kHz -> 2 kHz,
 * BPSK on a 10 kHz carrier at 44 samples per bit, frames of 100 bits + 8 zero bits.
 *   tmk_ask_modulate      replaces the TX loop                  examples/ask.rs:608-637 (live :176-218)
 *   tmk_ask_decode_batch  replaces the RX loop                  examples/ask.rs:668-796 (live :360-496):
 *                         EMA power, 440-tap sliding chirp correlation / 200, peak pick with 200-sample
 *                         hold-off, carrier mixing, 11-tap moving average, integrate-and-dump, id check
 *   tmk_ask_tables        the preamble / carrier tables         examples/ask.rs:574-605
 */
#define TMK_ASK_PREAMBLE 440
#define TMK_ASK_SAMPLES_PER_BIT 44
#define TMK_ASK_FRAME_BITS 108
#define TMK_ASK_DATA_BITS 100
#define TMK_ASK_FRAME_SAMPLES (TMK_ASK_SAMPLES_PER_BIT * TMK_ASK_FRAME_BITS) /* 4752 */

typedef struct tmk_ask_rec {
    uint32_t stream;
    uint32_t accepted;    /* 1 <= id <= 100 (ask.rs:771) */
    uint64_t start_index; /* sample index of the sync peak that locked (ask.rs:716) */
    float sync_power;     /* its value */
    uint8_t id;           /* first 8 decided bits, MSB first */
    uint8_t bits[14];     /* the 108 decided bits, MSB first: bit j is bits[j / 8] >> (7 - j % 8) */
    uint8_t _pad[5];
} tmk_ask_rec;

/* the tables the engine uses (computed once on the host in the reference's f32 operation order) */
int tmk_ask_tables(tmk_engine* e, float* preamble440, float* carrier, uint64_t n_carrier);
/* bits: n_frames x 100 values 0/1; gap1/gap2: zeros before/after each frame. Returns the track. */
uint64_t tmk_ask_modulated_len(uint32_t n_frames, const uint32_t* gap1, const uint32_t* gap2);
int tmk_ask_modulate(tmk_engine* e, const uint8_t* bits, uint32_t n_frames, const uint32_t* gap1,
                     const uint32_t* gap2, float* out_samples, uint64_t out_cap, uint64_t* out_len);
/* one record per locked frame, sorted by (stream, start_index). Only offset / n_samples of the
 * streams are used. samples: host or device; recs: host. */
int tmk_ask_decode_batch(tmk_engine* e, const float* samples, uint64_t total_samples, const tmk_stream* streams,
                         uint32_t n_streams, tmk_ask_rec* recs, uint64_t rec_cap, uint64_t* n_recs);
/* Synthetic chirp + BPSK captures (measurement harness): n_streams x n_per_stream samples of AWGN (SNR against the
 * BPSK carrier power 0.5) plus frames laid out like the TX loop of examples/ask.rs:608-637 (gaps U[0,100) before and
 * after every frame, ids 1..100 cycling, hashed payload bits), generated on the GPU. samples: device or host. */
int tmk_ask_synth_batch(tmk_engine* e, uint64_t seed, uint32_t n_streams, uint64_t n_per_stream, float snr_db,
                        float* samples, uint64_t* n_frames);
/* device time of the stages of the last tmk_ask_decode_batch call, ms: [0] chirp scan, [1] walk */
int tmk_ask_last_stage_ms(tmk_engine* e, float ms[2]);

/* ------------------------------------------------------------------------------------------------
 * Multi-carrier modem: FSK and OFDM demodulation on a batched shared-memory FFT with symbol decision, bit packing
 * and CRC-8 (north_star stages 3-4, BASELINE config 3). THE REFERENCE HAS NO CODE FOR THIS PATH (no FFT, FSK or
 * OFDM in any .rs file; docs/proj/report.md:177-319 describes an amodem-style multi-carrier design in prose): the
 * specification is this build's own (oracle/mcm_modem.hpp) and every result of these entry points
 * is "parity unpinned by the reference". What IS the reference's: the 440-sample chirp preamble and its synchroniser
 * (examples/ask.rs:586-605, :693-723), the MSB-first bit packing and CRC-8 (src/phy/crc.rs:7-22,41-60).
 * A frame is the chirp followed by a 4752-sample body (the envelope of an ask.rs frame):
 *   TMK_MCM_OFDM  16 zeros | training symbol | 13 data symbols | 256 zeros; symbol = 64 (cyclic prefix) + 256 samples,
 *                 32 carriers on FFT bins 16..47, QPSK, one-tap equaliser from the training symbol; 104 bytes
 *   TMK_MCM_FSK   16 zeros | 56 symbols | 256 zeros; symbol = 16 + 64 samples of one of 16 tones (FFT bins 4..19),
 *                 decision = strongest bin; 28 bytes
 *   bytes = [seq][len][payload][crc8(payload)][zero padding]; status: TMK_FRAME_OK / TMK_FRAME_BAD_LEN / TMK_FRAME_CRC_FAIL */
#define TMK_MCM_FSK 0
#define TMK_MCM_OFDM 1
#define TMK_MCM_BODY_SAMPLES 4752
#define TMK_MCM_MAX_PAYLOAD_FSK 25
#define TMK_MCM_MAX_PAYLOAD_OFDM 100

typedef struct tmk_mcm_rec {
    uint32_t stream;
    uint32_t status;      /* TMK_FRAME_OK, TMK_FRAME_BAD_LEN, TMK_FRAME_CRC_FAIL */
    uint64_t start_index; /* sample index of the sync peak that locked (the body starts one sample later) */
    uint64_t payload_off; /* `len` bytes at payload + payload_off (status OK or CRC_FAIL) */
    float sync_power;
    uint8_t seq, len, crc, mode;
} tmk_mcm_rec;

uint64_t tmk_mcm_modulated_len(uint32_t n_frames, const uint32_t* gap1, const uint32_t* gap2);
/* payload_off: n_frames + 1 offsets into `payload`; seq: one byte per frame; gap1 / gap2: zeros before / after a frame */
int tmk_mcm_modulate(tmk_engine* e, int32_t mode, const uint8_t* payload, const uint32_t* payload_off, const uint8_t* seq,
                     uint32_t n_frames, const uint32_t* gap1, const uint32_t* gap2, float* out_samples, uint64_t out_cap,
                     uint64_t* out_len);
/* one record per locked frame, sorted by (stream, start_index). samples: host or device; recs / payload: host. */
int tmk_mcm_decode_batch(tmk_engine* e, int32_t mode, const float* samples, uint64_t total_samples, const tmk_stream* streams,
                         uint32_t n_streams, tmk_mcm_rec* recs, uint64_t rec_cap, uint64_t* n_recs, uint8_t* payload,
                         uint64_t payload_cap, uint64_t* payload_len);
/* Synthetic captures (measurement harness): n_streams x n_per_stream samples of AWGN (SNR against the body power 0.25)
 * plus frames with hashed payloads, optionally through the 3-tap multipath channel of SURVEY §8d. */
int tmk_mcm_synth_batch(tmk_engine* e, int32_t mode, uint64_t seed, uint32_t n_streams, uint64_t n_per_stream, float snr_db,
                        int32_t multipath, float* samples, uint64_t* n_frames);
/* parity probe of the FFT itself: n_ffts rows of n (64 or 256) real samples -> n_ffts x n x (re, im) */
int tmk_mcm_fft_probe(tmk_engine* e, uint32_t n, const float* rows, uint32_t n_ffts, float* spectra);
/* device time of the last tmk_mcm_decode_batch: chirp filter, sync walk, FFT demodulation (ms) */
int tmk_mcm_last_stage_ms(tmk_engine* e, float ms[3]);

/* ------------------------------------------------------------------------------------------------
 * 16-bit PCM captures. Recordings reach the reference as int16 WAV:
 *   tmk_pcm16_encode        replaces the sample loop of dump_to_wav   src/utils/dump.rs:37-40
 *                           amplitude = (sample * i16::MAX as f32) as i16  (Rust `as`: truncates toward zero,
 *                           saturates at -32768 / 32767, NaN -> 0)
 *   tmk_pcm16_to_f32        replaces the S16 arm of convert_audio_buffer_to_f32   src/audio/codec.rs:121-125
 *                           sample as f32 / 32768.0 (exact in f32)
 *   tmk_decode_batch_pcm16  = tmk_pcm16_to_f32 on the device followed by tmk_decode_batch: host captures cross
 *                           PCIe as 2 bytes per sample; records / payload exactly as tmk_decode_batch returns them
 *                           for the converted samples. stream offsets are multiples of 4 samples, the pcm base
 *                           pointer is 8-byte aligned.
 * All pointers host or device. */
int tmk_pcm16_encode(tmk_engine* e, const float* samples, uint64_t n, int16_t* pcm_out);
int tmk_pcm16_to_f32(tmk_engine* e, const int16_t* pcm, uint64_t n, float* samples_out);
int tmk_decode_batch_pcm16(tmk_engine* e, const int16_t* pcm, uint64_t total_samples, const tmk_stream* streams,
                           uint32_t n_streams, tmk_frame_rec* recs, uint64_t rec_cap, uint64_t* n_recs,
                           uint8_t* payload, uint64_t payload_cap, uint64_t* payload_len);

/* WAV container ingest: RIFF/WAVE files as dump_to_wav stores recordings (src/utils/dump.rs:23-46,
 * hound: a `fmt ` chunk of 16 bytes, or the 40-byte WAVE_FORMAT_EXTENSIBLE form, and a `data` chunk) turned into the
 * f32 samples the reference's reader produces (convert_audio_buffer_to_f32, src/audio/codec.rs:114-148): CHANNEL 0
 * only; U8 (s - 128) / 128, S16 s / 32768, S24 s / 8388608, S32 s / 2147483648, F32 as is.
 *   tmk_wav_parse         the RIFF chunk walk (host only; TMK_E_INVALID for anything that is not such a file)
 *   tmk_wav_to_f32        n files back to back in one blob (host or device; file i = blob[file_off[i] .. file_off[i+1]))
 *                         -> channel 0 of file i at samples_out[offsets_out[i] .. + lens_out[i]) (offsets multiples of 4;
 *                         samples_out device or host; *total_out = floats needed, TMK_E_CAPACITY if sample_cap is smaller)
 *   tmk_decode_batch_wav  the same unpack on the device in front of tmk_decode_batch: the files cross PCIe as raw
 *                         bytes; record.stream = file index; streams_out (optional) receives the stream table.
 * FLAC (src/audio/codec.rs:10-112, symphonia) and the JSON AudioData dump (dump.rs:12-21) are decoded on the host
 * (the Python mirror reads the JSON form); their samples enter through tmk_decode_batch. */
typedef struct tmk_wav_info {
    uint32_t format;          /* 1 = integer PCM, 3 = IEEE float */
    uint32_t channels;
    uint32_t sample_rate;
    uint32_t bits_per_sample;
    uint32_t block_align;     /* bytes per frame (all channels) */
    uint32_t _pad;
    uint64_t data_off;        /* first byte of the first frame, from the start of the file */
    uint64_t n_frames;        /* samples per channel */
} tmk_wav_info;
int tmk_wav_parse(const uint8_t* file, uint64_t len, tmk_wav_info* out);
int tmk_wav_to_f32(tmk_engine* e, const uint8_t* blob, uint64_t blob_len, const uint64_t* file_off, uint32_t n_files,
                   float* samples_out, uint64_t sample_cap, uint64_t* offsets_out, uint64_t* lens_out,
                   uint64_t* total_out);
int tmk_decode_batch_wav(tmk_engine* e, const uint8_t* blob, uint64_t blob_len, const uint64_t* file_off,
                         uint32_t n_files, const tmk_phy_config* cfgs, tmk_stream* streams_out, tmk_frame_rec* recs,
                         uint64_t rec_cap, uint64_t* n_recs, uint8_t* payload, uint64_t payload_cap,
                         uint64_t* payload_len);

/* ------------------------------------------------------------------------------------------------
 * Counter reduction over the GPUs of a box: streams shard over ranks with no data-path traffic; the
 * only collective is an NCCL all-reduce of the decode counters (and of a step time) after a batch. One process
 * per GPU, one communicator per engine; libnccl is opened at run time (libnccl.so.2 or $TMK_NCCL_LIB).
 *   tmk_comm_unique_id   ncclGetUniqueId: rank 0 creates the 128-byte id and hands it to the other ranks (the
 *                        launcher's business: a file, a socket, an environment variable)
 *   tmk_comm_init        ncclCommInitRank on the engine's device
 *   tmk_stats_allreduce  the engine's tmk_stats counters summed over all ranks (ncclAllReduce, ncclUint64, ncclSum)
 *   tmk_comm_allreduce_* the same for caller-provided counters / the maximum of doubles (device time of a step)
 * Without a communicator the reductions are the identity, so single-GPU callers need no special case. */
int tmk_comm_unique_id(void* out128);
int tmk_comm_init(tmk_engine* e, const void* unique_id128, int rank, int world);
int tmk_comm_size(const tmk_engine* e);
int tmk_stats_allreduce(tmk_engine* e, uint64_t counters[TMK_STAT_COUNT]);
int tmk_comm_allreduce_sum_u64(tmk_engine* e, uint64_t* counters, uint32_t n);
int tmk_comm_allreduce_max_f64(tmk_engine* e, double* values, uint32_t n);
/* all[r * n + i] = rank r's mine[i] (n <= 8): per-GPU record counts / step times after a batch (ncclAllGather) */
int tmk_comm_allgather_u64(tmk_engine* e, const uint64_t* mine, uint32_t n, uint64_t* all);
int tmk_comm_barrier(tmk_engine* e);
int tmk_comm_destroy(tmk_engine* e);

/* ------------------------------------------------------------------------------------------------
 * IP fragments: the payload layer directly above the PHY. Decoded frames carry IPv4 packets or
 * fragments of them in Frame::data.
 *   tmk_reassemble_batch  replaces, per link, the loop of AcousticInterface::receive_packet
 *                         (src/mac/acoustic_interface.rs:343-366: frames with type Data, or Ack with data, go to the
 *                         reassembler) around IpReassembler::process_fragment (src/net/fragmentation.rs:232-401):
 *                         key (identification, source ip); a packet with MF = 0 and offset 0 passes through
 *                         unchanged; fragments are kept in arrival order, and once the MF = 0 fragment of a key was
 *                         seen every arrival re-checks the stable-sorted offsets for a gap-free chain starting at 0
 *                         (next = offset + ceil(len / 8)); a complete datagram = header of the FIRST fragment that
 *                         arrived + payloads in offset order, total_length rewritten (as u16), bytes 6..8 zeroed,
 *                         the key forgotten. Every link starts with an empty reassembler and, unlike
 *                         receive_packet (which returns at the first packet), every packet of the capture is
 *                         returned; an Err of process_fragment (fragmentation.rs:236-249) is a record with
 *                         status != TMK_PKT_OK instead of an early return.
 *   tmk_fragment_batch    replaces IpFragmenter::fragment_packet (src/net/fragmentation.rs:85-203) for a list of
 *                         packets sent through ONE fragmenter: packets <= mtu pass unchanged, others are cut into
 *                         ((mtu - ihl) / 8) * 8-byte pieces under copies of the first 20 header bytes (+ options)
 *                         with identification = first_identification + (number of fragmented packets before it),
 *                         wrapping; flags/offset, total length rewritten, checksum copied as is.
 */
#define TMK_PKT_OK 0
#define TMK_PKT_E_SHORT 1      /* fragmentation.rs:236-240 / :111-113: fewer than 20 bytes */
#define TMK_PKT_E_IHL 2        /* fragmentation.rs:246-248 / :122-124: ihl < 20 or ihl > len */
#define TMK_PKT_E_TOO_MANY 3   /* more than TMK_MAX_FRAGMENTS fragments under one key (engine limit) */
#define TMK_PKT_E_MTU 4        /* fragmentation.rs:131-136: (mtu - ihl) / 8 == 0 */
#define TMK_MAX_FRAGMENTS 512

typedef struct tmk_packet_rec {
    uint32_t link;      /* reassembly: stream index of the frames; fragmentation: index of the input packet */
    uint32_t status;    /* TMK_PKT_* */
    uint32_t frame;     /* reassembly: index of the frame record that completed (or failed) the packet;
                           fragmentation: index of the fragment within its packet */
    uint32_t len;       /* bytes at data_off (0 when status != TMK_PKT_OK) */
    uint64_t data_off;  /* into the output byte buffer */
    uint16_t identification;
    uint16_t n_fragments; /* reassembly: fragments joined (1 for a pass-through packet) */
    uint32_t _pad;
} tmk_packet_rec;

/* recs / payload: what tmk_decode_batch returned (host or device), sorted by stream. pkts (host) sorted by
 * (link, frame); bytes (host). */
int tmk_reassemble_batch(tmk_engine* e, const tmk_frame_rec* recs, uint64_t n_recs, const uint8_t* payload,
                         uint64_t payload_len, uint32_t n_links, tmk_packet_rec* pkts, uint64_t pkt_cap,
                         uint64_t* n_pkts, uint8_t* bytes, uint64_t byte_cap, uint64_t* n_bytes);
/* packets: n_packets IPv4 packets, packet i = data[pkt_off[i] .. pkt_off[i + 1]) (host or device; pkt_off host).
 * frags (host): one record per output fragment sorted by (link = packet, frame = fragment index), or one record
 * with status != TMK_PKT_OK for a packet fragment_packet would reject. *next_identification (optional) receives
 * the fragmenter's counter after the batch. */
int tmk_fragment_batch(tmk_engine* e, const uint8_t* data, const uint64_t* pkt_off, uint32_t n_packets, uint32_t mtu,
                       uint16_t first_identification, tmk_packet_rec* frags, uint64_t frag_cap, uint64_t* n_frags,
                       uint8_t* bytes, uint64_t byte_cap, uint64_t* n_bytes, uint16_t* next_identification);

/* The reassembler of a link kept across calls: AcousticInterface owns ONE IpReassembler for all its receive polls
 * (src/mac/acoustic_interface.rs:316-378), so the fragments of a datagram may arrive in different process_samples calls.
 * tmk_reasm_push takes the records of one decode call (tmk_decode_batch / tmk_links_process_samples: grouped by link,
 * record.stream = link), puts the fragments still pending from earlier calls in front of them and returns the packets
 * that complete now (same records and rules as tmk_reassemble_batch; `frame` = index of the completing record in THIS
 * call). On an error nothing is consumed. tmk_reasm_pending: fragments / bytes the link still holds. */
typedef struct tmk_reasm tmk_reasm;
int tmk_reasm_create(tmk_engine* e, uint32_t n_links, tmk_reasm** out);
void tmk_reasm_destroy(tmk_reasm* r);
int tmk_reasm_push(tmk_reasm* r, const tmk_frame_rec* recs, uint64_t n_recs, const uint8_t* payload, uint64_t payload_len,
                   tmk_packet_rec* pkts, uint64_t pkt_cap, uint64_t* n_pkts, uint8_t* bytes, uint64_t byte_cap,
                   uint64_t* n_bytes);
int tmk_reasm_pending(const tmk_reasm* r, uint32_t link, uint32_t* n_fragments, uint64_t* n_bytes);
/* link == UINT32_MAX forgets every link's fragments */
int tmk_reasm_reset(tmk_reasm* r, uint32_t link);

#ifdef __cplusplus
}
#endif
#endif /* TMKPHY_H */
