/* brotligpu.h — C ABI of the B200 Brotli encoder hot path (libbrotligpu.so).
 *
 * The reference (countertype/brotli-lib) has no FFI layer; its only boundary is the TypeScript
 * API.  Every entry point below names the reference interface it stands in for (paths relative
 * to the reference tree).  node/addon.cc (Node-API; compiled only where <node_api.h> exists, which
 * is not the case in this repository's build image) binds these 1:1; the shipped, tested binding is
 * ctypes (brotli_lib_b200/_lib.py).  Plain C, no exceptions cross the boundary, every int-returning function yields
 * 0 or a negative BGPU_E_* code.  There is NO CPU fallback: without a usable CUDA device the
 * calls fail with BGPU_E_NO_DEVICE / BGPU_E_CUDA.
 */
#ifndef BROTLIGPU_H_
#define BROTLIGPU_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BGPU_OK 0
#define BGPU_E_INVALID (-1)          /* NULL pointer / bad argument */
#define BGPU_E_OUTPUT_TOO_SMALL (-2) /* *out_len is set to the size needed */
#define BGPU_E_UNSUPPORTED (-3)      /* a call the GPU path has no form for (bgpu_encode_part at quality 0) */
#define BGPU_E_NO_DEVICE (-4)
#define BGPU_E_CUDA (-5)             /* bgpu_last_cuda_error() has the text */
#define BGPU_E_NOMEM (-6)
#define BGPU_E_STATE (-7)            /* stream used after finish */
#define BGPU_E_VERIFY (-8)           /* BGPU_F_VERIFY: a shard did not decode back to its input (bgpu_last_counter(ctx, 4) = how many) */
#define BGPU_E_DECODE (-9)           /* bgpu_decode: corrupt / truncated stream, or a static-dictionary reference (not supported) */

/* src/encode/enc-constants.ts:56-60  const enum EncoderMode */
#define BGPU_MODE_GENERIC 0
#define BGPU_MODE_TEXT 1
#define BGPU_MODE_FONT 2

typedef struct bgpu_ctx bgpu_ctx;
typedef struct bgpu_stream bgpu_stream;

/* Tunables that have no counterpart in the reference (it is single-threaded and serial).
 * 0 selects the default.  segment_bytes = 0xFFFFFFFF selects the exact serial parse (one
 * parse unit per shard, bit-identical commands to src/encode/backward-references.ts:14-134). */
typedef struct bgpu_tuning {
  uint32_t shard_bytes;     /* independent window, <= 4 MiB (default 4 MiB)      */
  uint32_t segment_bytes;   /* parse unit inside a shard (default 8 KiB for q5-9, 32 KiB for q10/11)  */
  uint32_t metablock_bytes; /* metablock size inside a shard (default = shard)                    */
  uint32_t flags;           /* BGPU_F_*                                                           */
  uint32_t lanes;           /* CUDA streams that encode groups of shards of one large input at once
                               (default 4; 1 = one in-order stream; q10/q11 use lanes only when set)  */
  uint32_t group_shards;    /* shards per group (default: shards / lanes, clamped to 16..64)       */
} bgpu_tuning;
#define BGPU_F_KEEP_DEVICE_OUTPUT 1u /* bench: skip the device->host gather (device-resident timing) */
#define BGPU_F_RING_HANDOVER 4u      /* accepted and ignored (round-1 experiment; the second parse pass + ring reconciliation replaced it) */
#define BGPU_F_PROBE_WINDOW 64u      /* match finding q5-9: the round-1 probe kernel (k_probe_sorted) instead of the lane-per-entry k_probe_diag (A/B runs; same best[]) */
#define BGPU_F_PRIO_MODEL 128u       /* lanes: the modelling phase (block split .. header) of each lane runs on a high-priority side stream (A/B runs) */
#define BGPU_F_WRITER_TWO_PASS 256u   /* bit writer q5-11: the round-1 two-pass writer (k_tilebits2 + k_write2 recompute every token) instead of k_tile_pack + k_tile_copy (A/B runs; same bytes) */
#define BGPU_F_SPLIT_3ITER 1024u     /* block split q5-8: the reference's 3 refinement iterations instead of 2 (size -0.007 points, step +6 %) */
#define BGPU_F_SORT_UNPACKED 2048u   /* match finding q5-6: (position, key bytes) pairs between the sort passes, as q7-9 and round 1 (A/B runs; same best[]) */
#define BGPU_F_NO_REFINE 32u         /* greedy parse: skip the second pass over segments that start inside a running copy (A/B runs) */
#define BGPU_F_MODEL_STATS 8u        /* bench: after the call bgpu_last_counter(1..3) hold the table bytes of the clustering / Huffman stages (one extra tiny kernel + a blocking read-back) */
#define BGPU_F_VERIFY 16u            /* after encoding, every shard is decoded ON THE DEVICE (csrc/decode.cuh, one thread per shard) and compared with the
                                        staged input: no device->host copy, no CPU decode; a mismatch fails the call with BGPU_E_VERIFY */
#define BGPU_F_TRIVIAL_METABLOCKS 2u /* storeMetaBlockTrivial for every metablock (metablock.ts:290-389): no block split, no context map */

/* Context: owns the CUDA device, stream, scratch arenas.  Not thread-safe; one per thread.
 * device_id < 0 selects the current device. */
int bgpu_ctx_create(int device_id, bgpu_ctx** out);
/* One context over several GPUs of one box (stands in for the serial metablock loop of
 * src/encode/encode.ts:206-278 run by one thread per device).  bgpu_encode on such a context cuts an input of more
 * than one shard into n_dev contiguous parts of whole shards, encodes part g on device_ids[g] from its own host thread
 * and returns ONE stream (parts are byte-aligned; the first carries the window header, the last ISLAST);
 * bgpu_encode_batch deals the items to the devices by bytes.  device_ids may repeat (several contexts on one GPU).
 * Stage entry points, streaming and bgpu_encode_part / bgpu_encode_device run on device_ids[0]. */
int bgpu_ctx_create_multi(const int* device_ids, int n_dev, bgpu_ctx** out);
void bgpu_ctx_destroy(bgpu_ctx* ctx);
int bgpu_ctx_set_tuning(bgpu_ctx* ctx, const bgpu_tuning* t);
/* Runs all further work of the context on the caller's CUDA stream (a cudaStream_t passed as void*;
 * NULL restores the context's own stream).  Lets a harness bracket calls with its own CUDA events. */
int bgpu_ctx_set_stream(bgpu_ctx* ctx, void* cuda_stream);

/* Upper bound of the stream size for n input bytes (any quality / lgwin / mode, default tuning):
 * n + n / 32 + 70 (n / 1008 + 1) + 16.  A shard whose compressed form would exceed len + len / 32 + 64 bytes is stored
 * as one uncompressed metablock (<= len + 6).  With a bgpu_tuning.shard_bytes below 1008 the calls may need more and
 * then report it through BGPU_E_OUTPUT_TOO_SMALL / *out_len. */
size_t bgpu_max_compressed_size(size_t n);

/* brotliEncode(input, {quality, lgwin, mode})            src/encode/encode.ts:50-90
 * Options are clamped, never rejected (encode.ts:57-62); TEXT == GENERIC; FONT sets
 * npostfix=1/ndirect=12 at quality >= 4 (enc-constants.ts:121-126).  n == 0 -> A1 01
 * (encode.ts:92-103); quality 0 or n < 64 -> uncompressed metablocks (encode.ts:105-138); both
 * are produced on the host (no kernel).  quality 1 = encodeFast (encode.ts:140-179) and quality 2..4
 * (encode.ts:192-193, 238-244) use the one-slot SimpleHasher (src/encode/hash-simple.ts) and always store
 * trivial metablocks (metablock.ts:517-524); quality 5..9 the hash-chain hasher, 10..11 H10 + Zopfli.
 * `in`/`out` are host buffers owned by the caller. */
int bgpu_encode(bgpu_ctx* ctx, const uint8_t* in, size_t n, int quality, int lgwin, int mode,
                uint8_t* out, size_t out_cap, size_t* out_len);

/* Batch of independent streams (BASELINE config 5): item i is brotliEncode(in[i], ...).  Batches of more than 8192 streams (or 8 GiB) are
 * run as several device jobs; one device job holds at most 65535 shards / 21845 metablocks (BGPU_E_INVALID beyond: only
 * reachable with lanes switched off and tiny shards). */
int bgpu_encode_batch(bgpu_ctx* ctx, int count, const uint8_t* const* in, const size_t* n,
                      int quality, int lgwin, int mode, uint8_t* const* out, const size_t* out_cap,
                      size_t* out_len);

/* One contiguous PART of a brotliEncode stream, for inputs split over several contexts / GPUs
 * (north_star: "each chunk is compressed on its own GPU ... stitched on the host into a single valid
 * stream at byte-aligned metablock boundaries").  The reference has no such call: it stands in for the
 * metablock loop of src/encode/encode.ts:213-281 run over bytes [off, off + n) of the input.
 *   BGPU_PART_FIRST  the part opens the stream: window header (bit-writer.ts:172-194) and the
 *                    {4, 11, 15, 16} distance ring (encode.ts:202);
 *   BGPU_PART_LAST   the part's final metablock carries ISLAST.
 * A part without FIRST starts with an unknown distance ring and takes its first literal contexts from
 * prev2[0..1] = the two input bytes before `in` (in stream order); a part without LAST ends with an
 * empty metadata block (RFC 7932 section 9.2) so that it is byte-aligned.  The concatenation of the
 * parts of an input, in order, is one valid stream.  Parts should be multiples of the 4 MiB shard.
 * quality >= 1 (quality 0, the host-only uncompressed path, has no part form: BGPU_E_UNSUPPORTED). */
#define BGPU_PART_FIRST 1u
#define BGPU_PART_LAST 2u
int bgpu_encode_part(bgpu_ctx* ctx, const uint8_t* in, size_t n, const uint8_t* prev2, unsigned part_flags,
                     int quality, int lgwin, int mode, uint8_t* out, size_t out_cap, size_t* out_len);

/* Same pipeline with the input already resident in device memory (d_in: device pointer, 256-byte
 * aligned, n + 64 bytes readable).  The compressed stream stays on the device unless `out` is
 * non-NULL.  Used by bench.py for the device-resident `value`; the timed CUDA-event duration of
 * the kernels is returned in *kernel_ms when non-NULL. */
int bgpu_encode_device(bgpu_ctx* ctx, const void* d_in, size_t n, int quality, int lgwin, int mode,
                       uint8_t* out, size_t out_cap, size_t* out_len, float* kernel_ms);

/* On-device Brotli decoder (format: RFC 7932, the reference's decoder is src/decode/engine.ts:1012-1518).
 * Decodes one whole stream with ONE device thread — meant for self-verification and tests, not for throughput; streams made
 * by this library are verified shard-parallel with BGPU_F_VERIFY instead.  No static dictionary: a stream that uses it
 * fails with BGPU_E_DECODE.  out_cap bytes are available at out; *out_len receives the decoded size. */
int bgpu_decode(bgpu_ctx* ctx, const uint8_t* stream, size_t n, uint8_t* out, size_t out_cap, size_t* out_len);

/* Decoder entry point that allocates nothing.  It has NO counterpart in the reference and was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same decoder as bgpu_decode (same device code, one device thread per stream), but its scratch memory — job
 * descriptors, results and one table arena per stream — is a file-scope `static __device__` array of
 * BGPU_NOALLOC_SCRATCH_BYTES (10 MiB) that every GPU which loads the library holds until the process exits.  The call makes
 * no cudaMalloc / cudaFree / cudaMallocHost and uses no context: the caller passes inputs and outputs.
 *   d_stream[k], stream_len[k]  DEVICE pointer to whole Brotli stream k and its size in bytes (k < n_streams);
 *   d_out[k], out_cap[k]        DEVICE pointer that receives its bytes, and the room there;
 *   out_len[k]                  (host) decoded size of stream k;   status[k] (host, may be NULL) 0 or a BGPU_E_* code per stream;
 *   cuda_stream                 a cudaStream_t of the CURRENT device (NULL: the default stream); the call returns after the
 *                               work on it has finished.
 * The device used is the calling thread's current device (cudaGetDevice).  The arrays d_stream .. status are host arrays.
 * Limit: n_streams <= BGPU_NOALLOC_MAX_STREAMS; more is refused with BGPU_E_INVALID (bgpu_last_cuda_error() names the limit)
 * — split the batch over several calls.  Returns 0, or the first failing stream's code (BGPU_E_DECODE, BGPU_E_OUTPUT_TOO_SMALL).
 * All calls on one device share the array, which keeps the previous call's contents (the call clears what it needs zeroed):
 * calls that overlap on the same GPU — two host threads, or two streams — race on it and must be serialised by the caller. */
#define BGPU_NOALLOC_MAX_STREAMS 8
#define BGPU_NOALLOC_SCRATCH_BYTES (10u << 20)
int bgpu_decode_noalloc(int n_streams, const void* const* d_stream, const size_t* stream_len, void* const* d_out,
                        const size_t* out_cap, size_t* out_len, int* status, void* cuda_stream);
size_t bgpu_noalloc_scratch_bytes(void);   /* = BGPU_NOALLOC_SCRATCH_BYTES: the per-GPU footprint of the array */

/* WOFF2 glyf/loca transform front end: the step upstream of BASELINE config 5.  Not in the
 * reference (README.md:63 only points at WOFF2); format: W3C WOFF 2.0 section 5.1 "Transformed glyf table format" (36-byte
 * header, nContour / nPoints / flag / glyph / composite / bbox / instruction sub-streams, optional overlapSimpleBitmap);
 * the transformed loca table is empty (section 5.3).  One thread per glyph of every table of the batch; out[i] receives
 * table i's transformed glyf (out_len[i] bytes; BGPU_E_OUTPUT_TOO_SMALL sets every out_len to the size needed). */
typedef struct bgpu_glyf_table {
  const uint8_t* glyf; size_t glyf_len;   /* the font's glyf table */
  const uint8_t* loca; size_t loca_len;   /* its loca table */
  uint32_t num_glyphs;                    /* maxp.numGlyphs */
  int index_format;                       /* head.indexToLocFormat: 0 = u16 offsets / 2, 1 = u32 offsets */
} bgpu_glyf_table;
size_t bgpu_woff2_glyf_bound(size_t glyf_len, uint32_t num_glyphs);   /* enough for ordinary fonts; not a proven bound */
int bgpu_woff2_transform_glyf(bgpu_ctx* ctx, int count, const bgpu_glyf_table* tables, uint8_t* const* out, const size_t* out_cap,
                              size_t* out_len);

/* new BrotliEncoder(opts) / update(chunk) / finish()     src/encode/encode.ts:290-409
 * update copies the chunk (encode.ts:360-363) and returns the bytes completed so far (possibly
 * none); *out stays valid until the next call on the same stream.  Every quality 0..11 (the streaming
 * class has no uncompressed path: quality 0..4 run the SimpleHasher, encode.ts:343-344). */
int bgpu_stream_create(bgpu_ctx* ctx, int quality, int lgwin, int mode, bgpu_stream** s);
int bgpu_stream_update(bgpu_stream* s, const uint8_t* chunk, size_t n, const uint8_t** out, size_t* out_len);
int bgpu_stream_finish(bgpu_stream* s, const uint8_t** out, size_t* out_len);
void bgpu_stream_destroy(bgpu_stream* s);

/* ---- stage entry points for parity tests (device does the work; host buffers in/out) ---- */

/* HashChainHasher.findLongestMatch, chain part            src/encode/hash-chains.ts:117-152
 * For every position p with p + 4 <= n: best match over the <= 1<<(quality-1) most recent
 * positions with the same hashBytes4 key (match.ts:162-172), maxLength = min(n-p, 128),
 * maxBackward = min(p, 4 MiB - 16).  len[p] == 0 when none.  n <= 4 MiB. */
int bgpu_stage_match_candidates(bgpu_ctx* ctx, const uint8_t* in, size_t n, int quality, int lgwin,
                                uint32_t* len, uint32_t* dist);

/* createBackwardReferences + trailing-insert merge        src/encode/backward-references.ts:14-134,
 * encode.ts:250-267.  Commands use the reference's 16-byte layout (src/encode/command.ts:138-144:
 * insertLen u32, copyLen u32, distExtra u32, cmdPrefix u16, distPrefix u16).  Honors the context's
 * segment_bytes (exact serial parse when 0xFFFFFFFF). n <= 4 MiB. */
int bgpu_stage_commands(bgpu_ctx* ctx, const uint8_t* in, size_t n, int quality, int lgwin, int mode,
                        void* cmds, size_t cmd_cap, size_t* ncmd);

/* storeMetaBlockTrivial on caller-supplied commands       src/encode/metablock.ts:290-389
 * (histograms :305-325, createHuffmanTree entropy-encode.ts:24-131, convertBitDepthsToSymbols
 * :234-258, buildAndStoreHuffmanTree context-map.ts:215-347, emit loop :355-383).
 * start_nbits/start_value: bits already in the stream (the window header).  tables (optional,
 * 3 x (u32 hist, u8 depth, u16 bits) laid out as oracle's TrivialTables) receives the histograms
 * and codes. */
int bgpu_stage_store_trivial(bgpu_ctx* ctx, const uint8_t* in, size_t n, const void* cmds, size_t ncmd,
                             int is_last, int mode, unsigned start_nbits, unsigned start_value,
                             uint8_t* out, size_t out_cap, size_t* out_len, uint64_t* out_bits, void* tables);

/* createHuffmanTree + convertBitDepthsToSymbols           src/encode/entropy-encode.ts:24-131, 234-258
 * and buildAndStoreHuffmanTree                            src/encode/context-map.ts:215-347
 * count histograms of `alphabet` symbols each (<= 704).  tree_bits: 1 KiB per histogram of
 * serialised prefix-code bits, tree_nbits their lengths (either may be NULL). */
int bgpu_stage_huffman(bgpu_ctx* ctx, const uint32_t* hist, int count, int alphabet,
                       uint8_t* depths, uint16_t* bits, uint8_t* tree_bits, uint32_t* tree_nbits);

/* ---- full-metablock stages (storeMetaBlock, src/encode/metablock.ts:504-761), cut at the model boundary ----
 * The float-driven decisions (block splits block-splitter.ts:394-464, context maps metablock.ts:763-797) are INPUTS
 * here, so that the integer stages behind them can be compared bit for bit with the reference's on identical inputs. */
typedef struct bgpu_block_split {       /* block-splitter.ts BlockSplit */
  uint32_t num_types, num_blocks;       /* num_types == 1: num_blocks is 0 or 1 (one block of type 0, no split code) */
  const uint8_t* types;                 /* [num_blocks] */
  const uint32_t* lengths;              /* [num_blocks], summing to the category's symbol count */
} bgpu_block_split;
typedef struct bgpu_mb_model {
  bgpu_block_split split[3];            /* literal, command, distance */
  int context_mode;                     /* context.ts ContextMode 0..3; -1: chooseContextMode (context.ts:179-227) on the device */
  const uint32_t* literal_context_map;  /* [split[0].num_types * 64] cluster of every (block type, context) */
  uint32_t num_literal_clusters;        /* <= 64 (metablock.ts:789) */
  const uint32_t* distance_context_map; /* [split[2].num_types * 4] */
  uint32_t num_distance_clusters;
} bgpu_mb_model;

/* Per-context BuildHistograms of storeMetaBlock            src/encode/metablock.ts:558-637
 * One metablock = the whole input (n <= 4 MiB) under the caller's commands, block splits and context mode (the maps of
 * `model` are not read).  prev2 (may be NULL = stream start, contexts 0, 0) holds the two bytes before `in`.
 * lit_hist [split[0].num_types * 64][256], cmd_hist [split[1].num_types][704], dist_hist [split[2].num_types * 4][A]
 * with A = 64 (GENERIC/TEXT) or 124 (FONT); *context_mode_out (optional) the mode used. */
int bgpu_stage_histograms(bgpu_ctx* ctx, const uint8_t* in, size_t n, const void* cmds, size_t ncmd, int mode, const uint8_t* prev2,
                          const bgpu_mb_model* model, uint32_t* lit_hist, uint32_t* cmd_hist, uint32_t* dist_hist, int* context_mode_out);

/* encodeContextMap (moveToFront + zero-run RLE + prefix code)   src/encode/context-map.ts:114-170
 * map: `size` cluster ids < num_clusters (<= 256); bits: LSB-first output, cap_bytes large; *nbits the bit count. */
int bgpu_stage_context_map(bgpu_ctx* ctx, const uint32_t* map, size_t size, uint32_t num_clusters, uint8_t* bits, size_t cap_bytes,
                           uint64_t* nbits);

/* storeMetaBlock emit: header, block-split codes, context maps, prefix codes, commands with block switches
 * src/encode/metablock.ts:651-760 (BlockEncoder :392-501, buildAndStoreBlockSplitCode :161-202, storeBlockSwitch :204-220)
 * on caller-supplied commands + splits + context mode + context maps; histograms and prefix codes are derived on the
 * device exactly as the default bgpu_encode path derives them (same kernels).  start_nbits/start_value as in
 * bgpu_stage_store_trivial; prev2 as in bgpu_stage_histograms. */
int bgpu_stage_store_metablock(bgpu_ctx* ctx, const uint8_t* in, size_t n, const void* cmds, size_t ncmd, int is_last, int mode,
                               unsigned start_nbits, unsigned start_value, const uint8_t* prev2, const bgpu_mb_model* model,
                               uint8_t* out, size_t out_cap, size_t* out_len);

/* BinaryTreeHasher.findAllMatches for every position    src/encode/hash-binary-tree.ts:57-227
 * (17-bit buckets, <= 64 tree levels, 128-byte tree compare, 63-position short scan), called in
 * position order with maxLength = n - p and maxBackward = min(p, 4 MiB - 16) as pass 0 of
 * createHqZopfliBackwardReferences does (backward-references-hq.ts:520-527), every position stored.
 * offsets[p] .. offsets[p+1] index the (dist, len) entries of position p (strictly increasing
 * length).  offsets has n + 1 entries; n <= 4 MiB. */
int bgpu_stage_h10_matches(bgpu_ctx* ctx, const uint8_t* in, size_t n, int lgwin, uint32_t* offsets,
                           uint32_t* dist, uint32_t* len, size_t cap, size_t* total);

/* Per-kernel CUDA-event milliseconds of the last encode call on this context.  names/ms receive up
 * to cap entries; returns the number of stages. */
int bgpu_stage_timings(bgpu_ctx* ctx, const char** names, float* ms, int cap);
/* Number of kernel launches issued by the last encode call. */
int bgpu_last_launch_count(bgpu_ctx* ctx);
/* Work counters of the last encode call: which = 0 -> H10 match entries found (sum over all positions); with
 * BGPU_F_MODEL_STATS also 1 -> bytes of histogram rows entering the clustering problems (4 A H_in; clusterBlocks + context
 * maps), 2 -> bytes of rows leaving them (4 A H_out), 3 -> bytes of the prefix-code stage (7 A H: u32 count in, u8 depth +
 * u16 code out), SURVEY.md section 8d's per-stage algorithmic traffic. */
uint64_t bgpu_last_counter(bgpu_ctx* ctx, int which);

const char* bgpu_strerror(int code);
const char* bgpu_last_cuda_error(void);
int bgpu_version(void);

#ifdef __cplusplus
}
#endif
#endif /* BROTLIGPU_H_ */
