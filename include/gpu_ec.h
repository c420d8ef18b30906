/*
 * gpu_ec.h — C ABI of libgpu_ec.so: the B200 (sm_100a) implementation of the Reed-Solomon
 * GF(2^8) + CRC-32 byte path of getumen/rust-hadoop-generated-by-llm.
 *
 * The reference has no FFI for this path (it calls two Rust crates directly); each entry point
 * below names the reference call site it replaces (paths relative to the reference root).
 * A Rust `dfs/gpu_ec` crate binds these with `extern "C"`.
 *
 * Conventions
 *   - every function returns a gpu_ec_status (0 = OK, negative = error) unless noted;
 *   - the caller owns all buffers; nothing is retained past return; outputs are caller-allocated;
 *   - host-pointer entry points are synchronous, thread-safe and re-entrant (each call checks out
 *     a device + stream + staging slot); `_dev` entry points take device pointers and a
 *     cudaStream_t (as void*), only enqueue work, and never synchronise;
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails with
 *     GPU_EC_ERR_NO_DEVICE;
 *   - checksums are CRC-32/ISO-HDLC ("IEEE", what crc32fast computes; the reference's *_crc32c
 *     field names are a misnomer) unless poly = GPU_EC_POLY_CASTAGNOLI is passed.
 */
#ifndef GPU_EC_H
#define GPU_EC_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum gpu_ec_status {
    GPU_EC_OK = 0,
    GPU_EC_ERR_ZERO_SHARDS = -1,     /* erasure.rs:8-10  "data_shards and parity_shards must both be > 0" */
    GPU_EC_ERR_EMPTY_DATA = -2,      /* erasure.rs:11-13 "data must not be empty" */
    GPU_EC_ERR_TOO_MANY_SHARDS = -3, /* crate Error::TooManyShards (k+m > 256) -> "RS init error: TooManyShards" */
    GPU_EC_ERR_TOO_FEW_PRESENT = -4, /* crate Error::TooFewShardsPresent -> "RS reconstruct error: ..." */
    GPU_EC_ERR_EMPTY_SHARD = -5,     /* crate Error::EmptyShard */
    GPU_EC_ERR_SHARD_SIZE = -6,      /* crate Error::IncorrectShardSize */
    GPU_EC_ERR_COUNT_MISMATCH = -7,  /* chunkserver.rs:277-279 "Checksum count mismatch" */
    GPU_EC_ERR_CRC_MISMATCH = -8,    /* chunkserver.rs:281-289 "Checksum mismatch at chunk {i}" */
    GPU_EC_ERR_CUDA = -9,            /* a CUDA call failed; see gpu_ec_last_error() */
    GPU_EC_ERR_OOM = -10,
    GPU_EC_ERR_INVALID_ARG = -11,
    GPU_EC_ERR_NO_DEVICE = -12,
    GPU_EC_ERR_UNALIGNED = -13,      /* a _dev pointer/stride is not 16-byte aligned */
    GPU_EC_ERR_SINGULAR = -14
} gpu_ec_status;

enum { GPU_EC_POLY_IEEE = 0, GPU_EC_POLY_CASTAGNOLI = 1 };

/* ---- lifecycle ------------------------------------------------------------------------- */
int gpu_ec_init(void);                       /* optional; every entry point initialises lazily */
/* Frees streams and staging buffers. Call it only when no other thread is inside an entry point and none will
 * enter one concurrently (a call racing a shutdown is undefined); the next call after it initialises again. */
void gpu_ec_shutdown(void);
int gpu_ec_device_count(void);               /* CUDA devices visible (0 if none) */
/* Restrict host-pointer entry points to these CUDA ordinals (round-robin between them).
 * Default: all visible devices. One-process-per-GPU callers pass their own ordinal. */
int gpu_ec_set_devices(const int *devices, int n);
const char *gpu_ec_strerror(int status);     /* reference-compatible message text */
const char *gpu_ec_last_error(void);         /* thread-local detail of the last GPU_EC_ERR_CUDA */
int gpu_ec_alloc_pinned(void **ptr, size_t bytes);   /* page-locked host memory for the Rust Vec-equivalents */
int gpu_ec_free_pinned(void *ptr);

/* ---- pure host helpers (matrices are <= 256x256 bytes; no data-path work) --------------- */
/* erasure.rs:56-59 shard_len: ceil(len / k). Returns 0 when k == 0 (the Rust fn panics). */
size_t gpu_ec_shard_len(size_t len, size_t k);
/* crate ReedSolomon::new (call sites erasure.rs:14,36; chunkserver.rs:620): the (k+m) x k
 * systematic matrix, row-major. */
int gpu_ec_rs_matrix(int k, int m, uint8_t *out);
/* The one-pass decode rows the reconstruct kernels use: for every absent slot (index order) a
 * row of k coefficients over the FIRST k present shards (the crate's survivor choice).
 * rows_out: up to (k+m)*k bytes; missing_out: up to k+m ints; survivors_out: k ints. */
int gpu_ec_rs_decode_matrix(int k, int m, const uint8_t *present, uint8_t *rows_out,
                            int *missing_out, int *survivors_out, int *n_missing_out);
/* chunkserver.rs:196: checksums.flat_map(u32::to_be_bytes) */
void gpu_ec_meta_encode_be(const uint32_t *crcs, size_t n, uint8_t *out);
/* combine two finalised CRCs of adjacent buffers (4-byte arithmetic) */
uint32_t gpu_ec_crc32_combine(uint32_t crc1, uint32_t crc2, size_t len2, int poly);

/* ---- host-pointer entry points (H2D + kernels + D2H inside the call) --------------------- */
/* erasure.rs:7-24 encode(), fused with the client's per-shard and whole-buffer CRCs
 * (client/src/mod.rs:326-341, 600-611) and the chunkserver's per-512-byte .meta arrays
 * (chunkserver.rs:182-196) when the optional outputs are non-NULL.
 *   shards_out[k+m]  : each shard_len bytes; [0,k) receive the zero-padded input split (a NULL entry
 *                      skips that host copy: the shard is data[j*shard_len..] verbatim), [k,k+m) parity
 *   shard_crc_out    : NULL or k+m words, CRC-32 of each shard
 *   full_crc_out     : NULL or 1 word, CRC-32 of the original `len` bytes
 *   meta_out[k+m]    : NULL or, per shard, 4*ceil(shard_len/512) bytes of big-endian chunk CRCs */
int gpu_ec_encode(const uint8_t *data, size_t len, int k, int m, uint8_t *const *shards_out,
                  uint32_t *shard_crc_out, uint32_t *full_crc_out, uint8_t *const *meta_out);
/* crate ReedSolomon::reconstruct (erasure.rs:38; chunkserver.rs:620-623): fills EVERY absent
 * slot (data and parity) in place; shards[i] must point at shard_len writable bytes for absent
 * slots too. present[i] != 0 marks a surviving shard. */
int gpu_ec_reconstruct(uint8_t *const *shards, const uint8_t *present, size_t shard_len, int k, int m);
/* The repair path's variant (chunkserver.rs:503-640 keeps only opt_shards[shard_index] of what
 * reconstruct rebuilt): rebuild ONLY the absent slots listed in want[n_want] (one pass over the first
 * k present shards, n_want outputs); other absent slots are left untouched. Same errors as
 * gpu_ec_reconstruct; a wanted slot that is present or out of range is GPU_EC_ERR_INVALID_ARG.
 * shards[i] of an absent slot that is not wanted is never dereferenced and may be NULL. */
int gpu_ec_reconstruct_only(uint8_t *const *shards, const uint8_t *present, size_t shard_len, int k, int m,
                            const int *want, int n_want);
/* erasure.rs:30-49 decode(): reconstruct, then concatenate shards[0..k) truncated to original_len.
 * `present` is updated (all ones on success), mirroring the in-place Option fill. */
int gpu_ec_decode(uint8_t *const *shards, uint8_t *present, size_t shard_len, int k, int m,
                  size_t original_len, uint8_t *out);
/* crc32fast::Hasher::new/update/finalize over one buffer (chunkserver.rs:746-750,1006-1010;
 * client/src/mod.rs:329-341,426-428,601-611). len == 0 yields 0 like the crate. */
int gpu_ec_crc32(const uint8_t *data, size_t len, int poly, uint32_t *out);
/* chunkserver.rs:182-190 calculate_checksums (chunk = 512 there): out[ceil(len/chunk)] */
int gpu_ec_crc32_chunks(const uint8_t *data, size_t len, size_t chunk, int poly, uint32_t *out);
/* chunkserver.rs:265-289 verify_block compare (and :296-351 verify_partial_read when handed the
 * affected chunk range): parses meta_be like chunks_exact(4), recomputes, compares count then
 * element-wise. Returns OK, GPU_EC_ERR_COUNT_MISMATCH, or GPU_EC_ERR_CRC_MISMATCH with
 * *first_bad = index of the first differing chunk (else -1). */
int gpu_ec_verify_chunks(const uint8_t *data, size_t len, size_t chunk, int poly,
                         const uint8_t *meta_be, size_t meta_len, int64_t *first_bad);
/* chunkserver.rs:746-769 + 192-196 in one pass over the bytes:
 * whole-block CRC and the per-chunk array from a single H2D. Either output may be NULL. */
int gpu_ec_block_checksums(const uint8_t *data, size_t len, size_t chunk, int poly,
                           uint32_t *whole_crc_out, uint32_t *chunk_crc_out);
/* The same for n blocks of different lengths in one go (a chunkserver taking many sub-MiB WriteBlocks,
 * chunkserver.rs:746-769 + :195 per block): one H2D per block into one staging buffer, ONE CRC launch over
 * every chunk of every block, one small launch that repairs short last chunks and folds the whole-block
 * CRCs (per-block offsets in a device array), one synchronisation. whole_crc_out: n words or NULL;
 * chunk_crc_out: n pointers (ceil(lens[b]/chunk) words each; entries may be NULL) or NULL. chunk must be a
 * multiple of 128 up to 1024 for the one-launch path (512 in the reference); other sizes run block by block. */
int gpu_ec_block_checksums_batch(size_t n, const uint8_t *const *data, const size_t *lens, size_t chunk, int poly,
                                 uint32_t *whole_crc_out, uint32_t *const *chunk_crc_out);

/* Batches of equal-sized stripes laid out back to back in (ideally pinned) host memory;
 * stripes are independent and are pipelined over streams and spread round-robin over the
 * devices of gpu_ec_set_devices (stripe i -> device i mod G; no collective).
 *   data            : n_stripes * stripe_len bytes
 *   parity_out      : n_stripes * m * shard_len bytes (stripe-major, then parity row)
 *   shard_crc_out   : NULL or n_stripes * (k+m) words;  full_crc_out: NULL or n_stripes words */
int gpu_ec_encode_batch(const uint8_t *data, size_t stripe_len, size_t n_stripes, int k, int m,
                        uint8_t *parity_out, uint32_t *shard_crc_out, uint32_t *full_crc_out);
/* shards: n_stripes * (k+m) * shard_len bytes (stripe-major, then slot), same erasure pattern
 * in every stripe; absent slots are filled in place. */
int gpu_ec_reconstruct_batch(uint8_t *shards, const uint8_t *present, size_t shard_len,
                             size_t n_stripes, int k, int m);

/* run_background_scrubber (chunkserver.rs:642-718) as one batched call: verify n_blocks blocks of
 * different sizes against their .meta bytes, pipelined over streams (and devices). Per block i:
 * status_out[i] = GPU_EC_OK | GPU_EC_ERR_COUNT_MISMATCH | GPU_EC_ERR_CRC_MISMATCH and, when
 * first_bad_out is non-NULL, first_bad_out[i] = first differing chunk or -1 — the same verdicts
 * verify_block gives block by block. The return value reports call-level failures only. */
int gpu_ec_verify_chunks_batch(size_t n_blocks, const uint8_t *const *data, const size_t *lens, size_t chunk,
                               int poly, const uint8_t *const *meta_be, const size_t *meta_lens,
                               int *status_out, int64_t *first_bad_out);

/* ---- device-resident entry points (device pointers, stream = cudaStream_t) --------------- */
/* They only enqueue work on `stream` and never synchronise. The device that owns the pointer is used:
 * the call makes it current for its launches and restores the caller's current device before returning. */
/* n_stripes stripes of k+m shard slots in device memory:
 * slot s of stripe i starts at base + i*stripe_stride + s*shard_stride and holds shard_len bytes.
 * base and both strides must be multiples of 16 and shard_stride >= round_up(shard_len, 16):
 * kernels move whole 16-byte vectors and may overwrite the pad bytes of OUTPUT slots. Every slot a
 * call touches must lie within 64 GiB of its stripe's start (slot offsets are 32-bit vector indices). */
typedef struct gpu_ec_stripes {
    uint8_t *base;
    size_t shard_stride;
    size_t stripe_stride;
    size_t shard_len;
    size_t n_stripes;
} gpu_ec_stripes;

/* crate ReedSolomon::encode (erasure.rs:21): slots [0,k) -> slots [k,k+m) */
int gpu_ec_encode_dev(const gpu_ec_stripes *s, int k, int m, void *stream);
/* crate ReedSolomon::reconstruct with one erasure pattern for all stripes, one pass:
 * reads the first k present slots, writes every absent slot. */
int gpu_ec_reconstruct_dev(const gpu_ec_stripes *s, const uint8_t *present, int k, int m, void *stream);
/* generic: out slot out_idx[r] = sum_j rows[r*n_in + j] * in slot in_idx[j]  (GF(2^8)) */
int gpu_ec_apply_matrix_dev(const gpu_ec_stripes *s, const uint8_t *rows, const int *in_idx, int n_in,
                            const int *out_idx, int n_out, void *stream);
/* n_buf buffers of `len` bytes at base + b*buf_stride; per buffer ceil(len/chunk) CRCs written to
 * out + b*out_stride (in words). big_endian != 0 stores each word byte-swapped (.meta layout).
 * base and buf_stride should be multiples of 16 and chunk a multiple of 128 (otherwise a slower generic
 * kernel runs). Chunks up to 1 KiB and chunks of 16-128 KiB (multiples of 4 KiB) are hashed and, where
 * needed, folded inside one kernel; other sizes keep per-segment CRCs in stream-ordered scratch
 * (cudaMallocAsync on `stream`) and fold them with a second kernel. */
int gpu_ec_crc32_chunks_dev(const uint8_t *base, size_t n_buf, size_t buf_stride, size_t len,
                            size_t chunk, int poly, uint32_t *out, size_t out_stride,
                            int big_endian, void *stream);
/* whole-buffer CRC of each of n_buf buffers: out[b] */
int gpu_ec_crc32_dev(const uint8_t *base, size_t n_buf, size_t buf_stride, size_t len, int poly,
                     uint32_t *out, void *stream);
/* recompute chunk CRCs of ONE buffer and compare with the device copy of the big-endian .meta
 * words; *first_bad (device int64) receives the first differing chunk index or -1. */
int gpu_ec_verify_chunks_dev(const uint8_t *data, size_t len, size_t chunk, int poly,
                             const uint8_t *meta_be, long long *first_bad, void *stream);
/* The client write sequence dfs/client/src/mod.rs:600-611 (encode, a CRC per shard, a CRC of the
 * original) for every stripe: parity + CRC-32 of all k+m shards (+ CRC of the original original_len
 * bytes when full_crc is non-NULL; original_len = 0 means k*shard_len). shard_crc: n_stripes*(k+m) words;
 * full_crc: n_stripes. For 8 <= k <= 12, m <= 4, poly = IEEE and shards of a few hundred KiB or more this is
 * ONE pass over HBM: the shard CRCs are accumulated in registers inside the encode kernel (fused_kernels.cuh)
 * and finished by two small kernels over ~4 % of the data; it takes stream-ordered scratch
 * (cudaMallocAsync on `stream`) of about 4.5 % of the stripes' size. Other shapes run an encode pass and a
 * CRC pass. full_crc never costs a data pass: it is derived from the shard CRCs, which is why the bytes of
 * the data slots beyond original_len must be the zero padding of erasure.rs:18. */
int gpu_ec_encode_crc_dev(const gpu_ec_stripes *s, int k, int m, size_t original_len, int poly,
                          uint32_t *shard_crc, uint32_t *full_crc, void *stream);
/* No counterpart in the reference. Same result as gpu_ec_crc32_dev (whole-buffer CRC of each of
 * n_buf device buffers -> out[b]) for callers that must not allocate: processes whose device
 * memory is budgeted once at start-up, latency paths that cannot risk an allocator call, and
 * work that is recorded for replay. The per-segment CRCs that gpu_ec_crc32_dev keeps in
 * stream-ordered scratch live here in a 4 MiB array built into the library (held by every GPU
 * that loads it), so: at most 2^20 segments per call (n_buf * ceil(len / segment), segment grown
 * up to 64 KiB to fit; larger calls return GPU_EC_ERR_INVALID_ARG with an explanatory
 * gpu_ec_last_error()); calls that may overlap on one GPU race on that array and must be
 * serialised; the first call on a device performs a one-time set-up (address look-up, table
 * upload) and should be made before any stream capture begins. No cudaMalloc/cudaFree or other
 * allocating routine is called. */
int gpu_ec_crc32_dev_noalloc(const uint8_t *base, size_t n_buf, size_t buf_stride, size_t len, int poly,
                             uint32_t *out, void *stream);
/* synthetic bytes: byte i = byte (i%8) of splitmix64(seed ^ (i/8)), i counted from byte_offset */
int gpu_ec_fill_synthetic_dev(uint8_t *out, size_t len, uint64_t seed, uint64_t byte_offset, void *stream);

/* introspection for tests / bench: kernels launched by this library in this process */
uint64_t gpu_ec_kernel_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* GPU_EC_H */
