/*
 * ec_oracle.h — CPU ORACLE (test infrastructure, NOT product code).
 *
 * Plain-C restatement of the reference's Reed-Solomon + CRC-32 byte path. Only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * leg may load this library; the product (libgpu_ec.so) never links or calls it.
 *
 * PARITY UNPINNED: the arithmetic lives in two un-vendored third-party crates
 * (reed-solomon-erasure 6.0.0, crc32fast 1.5.0; Cargo.lock:2754-2763, 692-698)
 * and no reference test asserts a parity byte or CRC value. The
 * restatement is anchored on external fixed points instead: the Backblaze
 * JavaReedSolomon known-answer vector, CRC check values, zlib.
 */
#ifndef EC_ORACLE_H
#define EC_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* error codes mirror reed-solomon-erasure's Error enum + erasure.rs bail!()s */
enum {
    ORC_OK = 0,
    ORC_ERR_ZERO_SHARDS = -1,      /* erasure.rs:8-10 */
    ORC_ERR_EMPTY_DATA = -2,       /* erasure.rs:11-13 */
    ORC_ERR_TOO_MANY_SHARDS = -3,  /* crate: k+m > 256 */
    ORC_ERR_TOO_FEW_PRESENT = -4,  /* crate: TooFewShardsPresent */
    ORC_ERR_EMPTY_SHARD = -5,      /* crate: EmptyShard */
    ORC_ERR_SINGULAR = -6,
    ORC_ERR_COUNT_MISMATCH = -7,   /* chunkserver.rs:277-279 */
    ORC_ERR_CRC_MISMATCH = -8      /* chunkserver.rs:281-289 */
};

enum { ORC_POLY_IEEE = 0, ORC_POLY_CASTAGNOLI = 1 };

/* GF(2^8), poly 0x11D, generator 2 */
uint8_t orc_gf_mul(uint8_t a, uint8_t b);
uint8_t orc_gf_div(uint8_t a, uint8_t b);
uint8_t orc_gf_exp(uint8_t a, unsigned n);

/* (k+m) x k systematic matrix, row-major; crate ReedSolomon::new */
int orc_rs_matrix(int k, int m, uint8_t *out);
/* rows for the `n_missing` absent slots over the first-k present shards.
 * survivors_out[k] receives the shard indices used. rows_out is n_missing x k. */
int orc_rs_decode_matrix(int k, int m, const uint8_t *present, uint8_t *rows_out,
                         int *missing_out, int *survivors_out, int *n_missing_out);

size_t orc_shard_len(size_t len, size_t k);               /* erasure.rs:56-59 */
/* erasure.rs:7-24. shards_out: (k+m)*S bytes, shard i at i*S */
int orc_encode(const uint8_t *data, size_t len, int k, int m, uint8_t *shards_out);
/* crate reconstruct (erasure.rs:38, chunkserver.rs:620-623). shards: (k+m)*S, in place */
int orc_reconstruct(uint8_t *shards, const uint8_t *present, size_t S, int k, int m);
/* erasure.rs:30-49 */
int orc_decode(uint8_t *shards, const uint8_t *present, size_t S, int k, int m,
               size_t original_len, uint8_t *out);

uint32_t orc_crc32(const uint8_t *data, size_t len, int poly);            /* Hasher new/update/finalize */
uint32_t orc_crc32_bytewise(const uint8_t *data, size_t len, int poly);   /* independent slow path */
uint32_t orc_crc32_combine(uint32_t crc1, uint32_t crc2, size_t len2, int poly);
/* chunkserver.rs:182-190 (chunk=512 there) */
size_t orc_crc32_chunks(const uint8_t *data, size_t len, size_t chunk, int poly, uint32_t *out);
/* chunkserver.rs:196 */
void orc_meta_encode_be(const uint32_t *crcs, size_t n, uint8_t *out);
/* chunkserver.rs:265-289: 0 ok, ORC_ERR_COUNT_MISMATCH, ORC_ERR_CRC_MISMATCH (+first_bad) */
int orc_verify_chunks(const uint8_t *data, size_t len, size_t chunk, int poly,
                      const uint8_t *meta_be, size_t meta_len, int64_t *first_bad);

/* synthetic bytes: splitmix64(seed, 8-byte offset) */
void orc_fill_synthetic(uint8_t *out, size_t len, uint64_t seed, uint64_t byte_offset);

/* multi-threaded CPU baselines over independent stripes (one stripe per thread at a time) */
double orc_bench_encode(const uint8_t *data, size_t stripe_len, size_t n_stripes, int k, int m,
                        int threads, int with_crc);
double orc_bench_crc_chunks(const uint8_t *data, size_t len, size_t chunk, int poly, int threads);
double orc_bench_reconstruct(const uint8_t *shards, const uint8_t *present, size_t S,
                             size_t n_stripes, int k, int m, int threads);
#ifdef __cplusplus
}
#endif
#endif
