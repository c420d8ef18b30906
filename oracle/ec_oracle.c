/*
 * ec_oracle.c — CPU ORACLE (test infrastructure, NOT product code). See ec_oracle.h.
 *
 * Restates, in plain C:
 *   - dfs/common/src/erasure.rs:7-59 (encode / decode / shard_len wrapper semantics)
 *   - reed-solomon-erasure 6.0.0 galois_8 (un-vendored; Cargo.lock:2754-2763): GF(2^8) with
 *     polynomial 0x11D / generator 2, Vandermonde -> systematic matrix, encode in the crate's
 *     loop order (outer input shard, inner output row, whole-slice mul / mul_xor with one
 *     256-entry MUL_TABLE row per coefficient; simd-accel is OFF in the reference build),
 *     reconstruct (first k present shards, inverted sub-matrix, data rows then parity rows)
 *   - crc32fast 1.5.0 (un-vendored; Cargo.lock:692-698): CRC-32/ISO-HDLC, table path
 *     (slice-by-8 here) and the x86 PCLMULQDQ fold path it selects at run time for len >= 128
 *   - dfs/chunkserver/src/chunkserver.rs:16,182-196,238-292 (512-byte chunk CRCs, big-endian
 *     .meta words, count-then-elementwise verify)
 * PARITY UNPINNED at the crate boundary (no reference test holds a golden byte); pinned here
 * against external fixed points in tests/test_oracle_golden.py.
 */
#define _GNU_SOURCE
#include "ec_oracle.h"
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#if defined(__x86_64__)
#include <immintrin.h>
#include <cpuid.h>
#endif

/* ------------------------------------------------------------------ GF(2^8) */
static uint8_t g_exp[512];
static uint8_t g_log[256];
static uint8_t g_mul[256][256]; /* crate MUL_TABLE */
static pthread_once_t g_once = PTHREAD_ONCE_INIT;

static uint32_t g_crc_tab[2][8][256];
static const uint32_t k_poly_reflected[2] = {0xEDB88320u, 0x82F63B78u};

static void init_tables(void) {
    unsigned x = 1;
    for (int i = 0; i < 255; i++) {
        g_exp[i] = (uint8_t)x;
        g_log[x] = (uint8_t)i;
        x <<= 1;
        if (x & 0x100) x ^= 0x11D;
    }
    for (int i = 255; i < 512; i++) g_exp[i] = g_exp[i - 255];
    g_log[0] = 0;
    for (int a = 0; a < 256; a++)
        for (int b = 0; b < 256; b++)
            g_mul[a][b] = (a == 0 || b == 0) ? 0 : g_exp[g_log[a] + g_log[b]];
    for (int p = 0; p < 2; p++) {
        for (uint32_t i = 0; i < 256; i++) {
            uint32_t c = i;
            for (int j = 0; j < 8; j++) c = (c >> 1) ^ ((c & 1) ? k_poly_reflected[p] : 0);
            g_crc_tab[p][0][i] = c;
        }
        for (int t = 1; t < 8; t++)
            for (uint32_t i = 0; i < 256; i++) {
                uint32_t c = g_crc_tab[p][t - 1][i];
                g_crc_tab[p][t][i] = (c >> 8) ^ g_crc_tab[p][0][c & 0xff];
            }
    }
}
static inline void ensure_init(void) { pthread_once(&g_once, init_tables); }

uint8_t orc_gf_mul(uint8_t a, uint8_t b) { ensure_init(); return g_mul[a][b]; }
uint8_t orc_gf_div(uint8_t a, uint8_t b) {
    ensure_init();
    if (a == 0) return 0;
    if (b == 0) return 0; /* undefined in the crate (panics); never reached by callers here */
    int d = (int)g_log[a] - (int)g_log[b];
    if (d < 0) d += 255;
    return g_exp[d];
}
uint8_t orc_gf_exp(uint8_t a, unsigned n) {
    ensure_init();
    if (n == 0) return 1;
    if (a == 0) return 0;
    unsigned r = ((unsigned)g_log[a] * n) % 255;
    return g_exp[r];
}

/* Gauss-Jordan inverse of an n x n matrix over GF(2^8); returns 0 or ORC_ERR_SINGULAR */
static int mat_invert(const uint8_t *a, int n, uint8_t *inv) {
    uint8_t *w = (uint8_t *)malloc((size_t)n * 2 * n);
    if (!w) return ORC_ERR_SINGULAR;
    for (int r = 0; r < n; r++) {
        memcpy(w + (size_t)r * 2 * n, a + (size_t)r * n, n);
        memset(w + (size_t)r * 2 * n + n, 0, n);
        w[(size_t)r * 2 * n + n + r] = 1;
    }
    int W = 2 * n;
    for (int c = 0; c < n; c++) {
        int piv = -1;
        for (int r = c; r < n; r++) if (w[(size_t)r * W + c]) { piv = r; break; }
        if (piv < 0) { free(w); return ORC_ERR_SINGULAR; }
        if (piv != c)
            for (int j = 0; j < W; j++) {
                uint8_t t = w[(size_t)c * W + j]; w[(size_t)c * W + j] = w[(size_t)piv * W + j]; w[(size_t)piv * W + j] = t;
            }
        uint8_t d = w[(size_t)c * W + c];
        if (d != 1) {
            uint8_t s = orc_gf_div(1, d);
            for (int j = 0; j < W; j++) w[(size_t)c * W + j] = g_mul[s][w[(size_t)c * W + j]];
        }
        for (int r = 0; r < n; r++) {
            if (r == c) continue;
            uint8_t f = w[(size_t)r * W + c];
            if (!f) continue;
            for (int j = 0; j < W; j++) w[(size_t)r * W + j] ^= g_mul[f][w[(size_t)c * W + j]];
        }
    }
    for (int r = 0; r < n; r++) memcpy(inv + (size_t)r * n, w + (size_t)r * W + n, n);
    free(w);
    return 0;
}

int orc_rs_matrix(int k, int m, uint8_t *out) {
    ensure_init();
    if (k <= 0 || m <= 0) return ORC_ERR_ZERO_SHARDS;
    if (k + m > 256) return ORC_ERR_TOO_MANY_SHARDS;
    int n = k + m;
    uint8_t *v = (uint8_t *)malloc((size_t)n * k);
    uint8_t *top_inv = (uint8_t *)malloc((size_t)k * k);
    for (int r = 0; r < n; r++)
        for (int c = 0; c < k; c++) v[(size_t)r * k + c] = orc_gf_exp((uint8_t)r, (unsigned)c);
    int rc = mat_invert(v, k, top_inv);
    if (rc == 0) {
        for (int r = 0; r < n; r++)
            for (int c = 0; c < k; c++) {
                uint8_t acc = 0;
                for (int t = 0; t < k; t++) acc ^= g_mul[v[(size_t)r * k + t]][top_inv[(size_t)t * k + c]];
                out[(size_t)r * k + c] = acc;
            }
    }
    free(v); free(top_inv);
    return rc;
}

int orc_rs_decode_matrix(int k, int m, const uint8_t *present, uint8_t *rows_out,
                         int *missing_out, int *survivors_out, int *n_missing_out) {
    int n = k + m;
    uint8_t *M = (uint8_t *)malloc((size_t)n * k);
    int rc = orc_rs_matrix(k, m, M);
    if (rc) { free(M); return rc; }
    int np = 0, ns = 0, nm = 0;
    for (int i = 0; i < n; i++) {
        if (present[i]) { np++; if (ns < k) survivors_out[ns++] = i; }
        else missing_out[nm++] = i;
    }
    *n_missing_out = nm;
    if (np < k) { free(M); return ORC_ERR_TOO_FEW_PRESENT; }
    uint8_t *sub = (uint8_t *)malloc((size_t)k * k), *inv = (uint8_t *)malloc((size_t)k * k);
    for (int r = 0; r < k; r++) memcpy(sub + (size_t)r * k, M + (size_t)survivors_out[r] * k, k);
    rc = mat_invert(sub, k, inv);
    if (rc == 0) {
        for (int i = 0; i < nm; i++) {
            int idx = missing_out[i];
            if (idx < k) memcpy(rows_out + (size_t)i * k, inv + (size_t)idx * k, k);
            else
                for (int c = 0; c < k; c++) {
                    uint8_t acc = 0;
                    for (int t = 0; t < k; t++) acc ^= g_mul[M[(size_t)idx * k + t]][inv[(size_t)t * k + c]];
                    rows_out[(size_t)i * k + c] = acc;
                }
        }
    }
    free(M); free(sub); free(inv);
    return rc;
}

/* crate galois_8::mul_slice / mul_slice_xor, pure-Rust path: one MUL_TABLE row, byte loop */
static void mul_slice(uint8_t c, const uint8_t *in, uint8_t *out, size_t n) {
    const uint8_t *mt = g_mul[c];
    for (size_t i = 0; i < n; i++) out[i] = mt[in[i]];
}
static void mul_slice_xor(uint8_t c, const uint8_t *in, uint8_t *out, size_t n) {
    const uint8_t *mt = g_mul[c];
    for (size_t i = 0; i < n; i++) out[i] ^= mt[in[i]];
}
/* crate code_some_slices: outer input, inner output row */
static void code_some(const uint8_t *const *rows, int n_in, int n_out,
                      const uint8_t *const *in, uint8_t *const *out, size_t S) {
    for (int j = 0; j < n_in; j++)
        for (int r = 0; r < n_out; r++) {
            if (j == 0) mul_slice(rows[r][j], in[j], out[r], S);
            else mul_slice_xor(rows[r][j], in[j], out[r], S);
        }
}

size_t orc_shard_len(size_t len, size_t k) { return (len + k - 1) / k; }

int orc_encode(const uint8_t *data, size_t len, int k, int m, uint8_t *shards_out) {
    ensure_init();
    if (k <= 0 || m <= 0) return ORC_ERR_ZERO_SHARDS;
    if (len == 0) return ORC_ERR_EMPTY_DATA;
    if (k + m > 256) return ORC_ERR_TOO_MANY_SHARDS;
    size_t S = orc_shard_len(len, (size_t)k);
    /* erasure.rs:17-20: copy + zero pad, split, zeroed parity */
    memcpy(shards_out, data, len);
    memset(shards_out + len, 0, S * (size_t)k - len);
    memset(shards_out + S * (size_t)k, 0, S * (size_t)m);
    uint8_t *M = (uint8_t *)malloc((size_t)(k + m) * k);
    int rc = orc_rs_matrix(k, m, M);
    if (rc) { free(M); return rc; }
    const uint8_t *rows[256]; const uint8_t *in[256]; uint8_t *out[256];
    for (int r = 0; r < m; r++) { rows[r] = M + (size_t)(k + r) * k; out[r] = shards_out + S * (size_t)(k + r); }
    for (int j = 0; j < k; j++) in[j] = shards_out + S * (size_t)j;
    code_some(rows, k, m, in, out, S);
    free(M);
    return 0;
}

int orc_reconstruct(uint8_t *shards, const uint8_t *present, size_t S, int k, int m) {
    ensure_init();
    if (k <= 0 || m <= 0) return ORC_ERR_ZERO_SHARDS;
    if (k + m > 256) return ORC_ERR_TOO_MANY_SHARDS;
    int n = k + m, np = 0;
    for (int i = 0; i < n; i++) np += present[i] ? 1 : 0;
    if (np == n) return 0;
    if (np < k) return ORC_ERR_TOO_FEW_PRESENT;
    if (S == 0) return ORC_ERR_EMPTY_SHARD;
    uint8_t *M = (uint8_t *)malloc((size_t)n * k);
    int rc = orc_rs_matrix(k, m, M);
    if (rc) { free(M); return rc; }
    int surv[256], ns = 0;
    for (int i = 0; i < n && ns < k; i++) if (present[i]) surv[ns++] = i;
    uint8_t *sub = (uint8_t *)malloc((size_t)k * k), *inv = (uint8_t *)malloc((size_t)k * k);
    for (int r = 0; r < k; r++) memcpy(sub + (size_t)r * k, M + (size_t)surv[r] * k, k);
    rc = mat_invert(sub, k, inv);
    if (rc) { free(M); free(sub); free(inv); return rc; }
    /* missing data shards from the inverse rows over the survivors */
    const uint8_t *rows[256]; const uint8_t *in[256]; uint8_t *out[256];
    int no = 0;
    for (int j = 0; j < k; j++) in[j] = shards + S * (size_t)surv[j];
    for (int d = 0; d < k; d++) if (!present[d]) { rows[no] = inv + (size_t)d * k; out[no] = shards + S * (size_t)d; no++; }
    if (no) code_some(rows, k, no, in, out, S);
    /* missing parity shards from the now-complete data shards */
    no = 0;
    for (int j = 0; j < k; j++) in[j] = shards + S * (size_t)j;
    for (int p = k; p < n; p++) if (!present[p]) { rows[no] = M + (size_t)p * k; out[no] = shards + S * (size_t)p; no++; }
    if (no) code_some(rows, k, no, in, out, S);
    free(M); free(sub); free(inv);
    return 0;
}

int orc_decode(uint8_t *shards, const uint8_t *present, size_t S, int k, int m,
               size_t original_len, uint8_t *out) {
    int rc = orc_reconstruct(shards, present, S, k, m);
    if (rc) return rc;
    size_t total = S * (size_t)k;
    size_t n = original_len < total ? original_len : total; /* Vec::truncate never grows */
    memcpy(out, shards, n);
    return 0;
}

/* ------------------------------------------------------------------ CRC-32 */
uint32_t orc_crc32_bytewise(const uint8_t *d, size_t len, int poly) {
    ensure_init();
    uint32_t c = 0xFFFFFFFFu;
    for (size_t i = 0; i < len; i++) c = (c >> 8) ^ g_crc_tab[poly][0][(c ^ d[i]) & 0xff];
    return ~c;
}

static uint32_t crc_table_update(uint32_t c, const uint8_t *d, size_t len, int poly) {
    const uint32_t (*T)[256] = g_crc_tab[poly];
    while (len >= 8) {
        uint32_t lo, hi;
        memcpy(&lo, d, 4); memcpy(&hi, d + 4, 4);
        lo ^= c;
        c = T[7][lo & 0xff] ^ T[6][(lo >> 8) & 0xff] ^ T[5][(lo >> 16) & 0xff] ^ T[4][lo >> 24] ^
            T[3][hi & 0xff] ^ T[2][(hi >> 8) & 0xff] ^ T[1][(hi >> 16) & 0xff] ^ T[0][hi >> 24];
        d += 8; len -= 8;
    }
    while (len--) c = (c >> 8) ^ T[0][(c ^ *d++) & 0xff];
    return c;
}

#if defined(__x86_64__)
static int g_has_pclmul = -1;
static int has_pclmul(void) {
    if (g_has_pclmul < 0) {
        unsigned a, b, c, d;
        g_has_pclmul = (__get_cpuid(1, &a, &b, &c, &d) && (c & (1u << 1)) && (c & (1u << 19))) ? 1 : 0;
    }
    return g_has_pclmul;
}
/* Intel "Fast CRC Computation Using PCLMULQDQ" fold for the reflected IEEE polynomial, the
 * scheme crc32fast's specialized::pclmulqdq module uses for len >= 128. len must be >= 64. */
__attribute__((target("pclmul,sse4.1")))
static uint32_t crc_pclmul_ieee(uint32_t crc, const uint8_t *buf, size_t len, size_t *consumed) {
    const __m128i k1k2 = _mm_set_epi64x(0x01c6e41596LL, 0x0154442bd4LL);
    const __m128i k3k4 = _mm_set_epi64x(0x00ccaa009eLL, 0x01751997d0LL);
    const __m128i k5k0 = _mm_set_epi64x(0x0000000000LL, 0x0163cd6124LL);
    const __m128i poly = _mm_set_epi64x(0x01f7011641LL, 0x01db710641LL);
    const uint8_t *p = buf;
    size_t n = len;
    __m128i x0, x1, x2, x3, x4, x5, x6, x7, x8, y5, y6, y7, y8;
    x1 = _mm_loadu_si128((const __m128i *)(p + 0x00));
    x2 = _mm_loadu_si128((const __m128i *)(p + 0x10));
    x3 = _mm_loadu_si128((const __m128i *)(p + 0x20));
    x4 = _mm_loadu_si128((const __m128i *)(p + 0x30));
    x1 = _mm_xor_si128(x1, _mm_cvtsi32_si128((int)crc));
    x0 = k1k2;
    p += 64; n -= 64;
    while (n >= 64) {
        x5 = _mm_clmulepi64_si128(x1, x0, 0x00); x6 = _mm_clmulepi64_si128(x2, x0, 0x00);
        x7 = _mm_clmulepi64_si128(x3, x0, 0x00); x8 = _mm_clmulepi64_si128(x4, x0, 0x00);
        x1 = _mm_clmulepi64_si128(x1, x0, 0x11); x2 = _mm_clmulepi64_si128(x2, x0, 0x11);
        x3 = _mm_clmulepi64_si128(x3, x0, 0x11); x4 = _mm_clmulepi64_si128(x4, x0, 0x11);
        y5 = _mm_loadu_si128((const __m128i *)(p + 0x00)); y6 = _mm_loadu_si128((const __m128i *)(p + 0x10));
        y7 = _mm_loadu_si128((const __m128i *)(p + 0x20)); y8 = _mm_loadu_si128((const __m128i *)(p + 0x30));
        x1 = _mm_xor_si128(_mm_xor_si128(x1, x5), y5); x2 = _mm_xor_si128(_mm_xor_si128(x2, x6), y6);
        x3 = _mm_xor_si128(_mm_xor_si128(x3, x7), y7); x4 = _mm_xor_si128(_mm_xor_si128(x4, x8), y8);
        p += 64; n -= 64;
    }
    x0 = k3k4;
    x5 = _mm_clmulepi64_si128(x1, x0, 0x00); x1 = _mm_clmulepi64_si128(x1, x0, 0x11);
    x1 = _mm_xor_si128(_mm_xor_si128(x1, x2), x5);
    x5 = _mm_clmulepi64_si128(x1, x0, 0x00); x1 = _mm_clmulepi64_si128(x1, x0, 0x11);
    x1 = _mm_xor_si128(_mm_xor_si128(x1, x3), x5);
    x5 = _mm_clmulepi64_si128(x1, x0, 0x00); x1 = _mm_clmulepi64_si128(x1, x0, 0x11);
    x1 = _mm_xor_si128(_mm_xor_si128(x1, x4), x5);
    while (n >= 16) {
        x2 = _mm_loadu_si128((const __m128i *)p);
        x5 = _mm_clmulepi64_si128(x1, x0, 0x00); x1 = _mm_clmulepi64_si128(x1, x0, 0x11);
        x1 = _mm_xor_si128(_mm_xor_si128(x1, x2), x5);
        p += 16; n -= 16;
    }
    /* 128 -> 64 */
    x2 = _mm_clmulepi64_si128(x1, x0, 0x10);
    x3 = _mm_setr_epi32(~0, 0, ~0, 0);
    x1 = _mm_srli_si128(x1, 8);
    x1 = _mm_xor_si128(x1, x2);
    x0 = k5k0;
    x2 = _mm_srli_si128(x1, 4);
    x1 = _mm_and_si128(x1, x3);
    x1 = _mm_clmulepi64_si128(x1, x0, 0x00);
    x1 = _mm_xor_si128(x1, x2);
    /* Barrett 64 -> 32 */
    x0 = poly;
    x2 = _mm_and_si128(x1, x3);
    x2 = _mm_clmulepi64_si128(x2, x0, 0x10);
    x2 = _mm_and_si128(x2, x3);
    x2 = _mm_clmulepi64_si128(x2, x0, 0x00);
    x1 = _mm_xor_si128(x1, x2);
    *consumed = (size_t)(p - buf);
    return (uint32_t)_mm_extract_epi32(x1, 1);
}
#endif

uint32_t orc_crc32(const uint8_t *d, size_t len, int poly) {
    ensure_init();
    uint32_t c = 0xFFFFFFFFu;
#if defined(__x86_64__)
    if (poly == ORC_POLY_IEEE && len >= 128 && has_pclmul()) { /* crc32fast dispatch threshold */
        size_t used = 0;
        c = crc_pclmul_ieee(c, d, len, &used);
        d += used; len -= used;
    }
#endif
    c = crc_table_update(c, d, len, poly);
    return ~c;
}

/* GF(2)[x] mod P, reflected representation (zlib multmodp) */
static uint32_t multmodp(uint32_t a, uint32_t b, uint32_t poly) {
    uint32_t m = 1u << 31, p = 0;
    for (;;) {
        if (a & m) { p ^= b; if ((a & (m - 1)) == 0) break; }
        m >>= 1;
        b = (b & 1) ? (b >> 1) ^ poly : b >> 1;
    }
    return p;
}
static uint32_t x8nmodp(size_t n_bytes, uint32_t poly) { /* x^(8n) mod P */
    uint32_t p = 1u << 31;          /* x^0 */
    uint32_t sq = 1u << (31 - 8);   /* x^8 */
    while (n_bytes) {
        if (n_bytes & 1) p = multmodp(sq, p, poly);
        sq = multmodp(sq, sq, poly);
        n_bytes >>= 1;
    }
    return p;
}
uint32_t orc_crc32_combine(uint32_t crc1, uint32_t crc2, size_t len2, int poly) {
    uint32_t P = k_poly_reflected[poly];
    return multmodp(x8nmodp(len2, P), crc1, P) ^ crc2;
}

size_t orc_crc32_chunks(const uint8_t *d, size_t len, size_t chunk, int poly, uint32_t *out) {
    size_t n = 0;
    for (size_t off = 0; off < len; off += chunk) {
        size_t l = len - off < chunk ? len - off : chunk;
        out[n++] = orc_crc32(d + off, l, poly);
    }
    return n;
}

void orc_meta_encode_be(const uint32_t *crcs, size_t n, uint8_t *out) {
    for (size_t i = 0; i < n; i++) {
        out[4 * i + 0] = (uint8_t)(crcs[i] >> 24); out[4 * i + 1] = (uint8_t)(crcs[i] >> 16);
        out[4 * i + 2] = (uint8_t)(crcs[i] >> 8);  out[4 * i + 3] = (uint8_t)(crcs[i]);
    }
}

int orc_verify_chunks(const uint8_t *d, size_t len, size_t chunk, int poly,
                      const uint8_t *meta, size_t meta_len, int64_t *first_bad) {
    size_t n_exp = meta_len / 4; /* chunks_exact(4): trailing bytes ignored */
    size_t n_act = (len + chunk - 1) / chunk;
    if (first_bad) *first_bad = -1;
    if (n_exp != n_act) return ORC_ERR_COUNT_MISMATCH;
    for (size_t i = 0; i < n_act; i++) {
        size_t off = i * chunk, l = len - off < chunk ? len - off : chunk;
        uint32_t e = ((uint32_t)meta[4 * i] << 24) | ((uint32_t)meta[4 * i + 1] << 16) |
                     ((uint32_t)meta[4 * i + 2] << 8) | meta[4 * i + 3];
        if (orc_crc32(d + off, l, poly) != e) { if (first_bad) *first_bad = (int64_t)i; return ORC_ERR_CRC_MISMATCH; }
    }
    return 0;
}

/* ------------------------------------------------------------------ synthetic data */
static inline uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
void orc_fill_synthetic(uint8_t *out, size_t len, uint64_t seed, uint64_t byte_offset) {
    /* byte i of the stream = byte (i%8) (little-endian) of splitmix64(seed ^ (i/8)) */
    size_t i = 0;
    while (i < len) {
        uint64_t pos = byte_offset + i;
        uint64_t w = splitmix64(seed ^ (pos >> 3));
        unsigned sh = (unsigned)(pos & 7);
        while (sh < 8 && i < len) { out[i++] = (uint8_t)(w >> (8 * sh)); sh++; }
    }
}

/* ------------------------------------------------------------------ CPU baselines */
static double now_s(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }

typedef struct {
    int kind; const uint8_t *data; size_t stripe_len, n_stripes; int k, m, with_crc, poly;
    size_t chunk; const uint8_t *present; size_t S; int tid, nthreads; volatile uint32_t sink;
} job_t;

static void *job_run(void *arg) {
    job_t *j = (job_t *)arg;
    if (j->kind == 0) { /* encode (+ per-shard CRCs + full CRC: client/src/mod.rs:600-611) */
        size_t S = orc_shard_len(j->stripe_len, (size_t)j->k);
        uint8_t *sh = (uint8_t *)malloc(S * (size_t)(j->k + j->m));
        uint8_t *padded = (uint8_t *)malloc(S * (size_t)j->k);
        for (size_t s = (size_t)j->tid; s < j->n_stripes; s += (size_t)j->nthreads) {
            const uint8_t *d = j->data + s * j->stripe_len;
            /* erasure.rs:17-18: data.to_vec() + resize = input copy #1; orc_encode's own copy into the
             * shard buffers is erasure.rs:19 (chunks().map(to_vec)) = input copy #2 */
            memcpy(padded, d, j->stripe_len);
            memset(padded + j->stripe_len, 0, S * (size_t)j->k - j->stripe_len);
            j->sink ^= padded[j->stripe_len - 1];
            orc_encode(padded, j->stripe_len, j->k, j->m, sh);
            j->sink ^= sh[S * (size_t)j->k];
            if (j->with_crc) {
                for (int i = 0; i < j->k + j->m; i++) j->sink ^= orc_crc32(sh + S * (size_t)i, S, ORC_POLY_IEEE);
                j->sink ^= orc_crc32(d, j->stripe_len, ORC_POLY_IEEE);
            }
        }
        free(sh);
        free(padded);
    } else if (j->kind == 1) { /* chunk CRCs: each thread takes a contiguous slab of whole chunks */
        size_t nchunks = (j->stripe_len + j->chunk - 1) / j->chunk;
        size_t per = (nchunks + (size_t)j->nthreads - 1) / (size_t)j->nthreads;
        size_t c0 = per * (size_t)j->tid, c1 = c0 + per > nchunks ? nchunks : c0 + per;
        if (c0 < c1) {
            size_t off = c0 * j->chunk, end = c1 * j->chunk > j->stripe_len ? j->stripe_len : c1 * j->chunk;
            uint32_t *out = (uint32_t *)malloc(4 * (c1 - c0));
            orc_crc32_chunks(j->data + off, end - off, j->chunk, j->poly, out);
            j->sink ^= out[0];
            free(out);
        }
    } else { /* reconstruct */
        size_t n = (size_t)(j->k + j->m);
        uint8_t *sh = (uint8_t *)malloc(j->S * n);
        for (size_t s = (size_t)j->tid; s < j->n_stripes; s += (size_t)j->nthreads) {
            memcpy(sh, j->data + s * j->S * n, j->S * n);
            orc_reconstruct(sh, j->present, j->S, j->k, j->m);
            j->sink ^= sh[0];
        }
        free(sh);
    }
    return NULL;
}

static double run_jobs(job_t *proto, int threads) {
    if (threads < 1) threads = 1;
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)threads);
    job_t *jobs = (job_t *)malloc(sizeof(job_t) * (size_t)threads);
    ensure_init();
    double t0 = now_s();
    for (int i = 0; i < threads; i++) { jobs[i] = *proto; jobs[i].tid = i; jobs[i].nthreads = threads; pthread_create(&th[i], NULL, job_run, &jobs[i]); }
    for (int i = 0; i < threads; i++) pthread_join(th[i], NULL);
    double t1 = now_s();
    free(th); free(jobs);
    return t1 - t0;
}

double orc_bench_encode(const uint8_t *data, size_t stripe_len, size_t n_stripes, int k, int m, int threads, int with_crc) {
    job_t p; memset(&p, 0, sizeof p);
    p.kind = 0; p.data = data; p.stripe_len = stripe_len; p.n_stripes = n_stripes; p.k = k; p.m = m; p.with_crc = with_crc;
    return run_jobs(&p, threads);
}
double orc_bench_crc_chunks(const uint8_t *data, size_t len, size_t chunk, int poly, int threads) {
    job_t p; memset(&p, 0, sizeof p);
    p.kind = 1; p.data = data; p.stripe_len = len; p.chunk = chunk; p.poly = poly;
    return run_jobs(&p, threads);
}
double orc_bench_reconstruct(const uint8_t *shards, const uint8_t *present, size_t S, size_t n_stripes, int k, int m, int threads) {
    job_t p; memset(&p, 0, sizeof p);
    p.kind = 2; p.data = shards; p.present = present; p.S = S; p.n_stripes = n_stripes; p.k = k; p.m = m;
    return run_jobs(&p, threads);
}
