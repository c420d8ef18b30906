// sfold_search.c — offline search for "sparse fold" strides of a CRC-32 polynomial.
//
// The CRC of a long stream can be accumulated WITHOUT tables or carry-less multiply if the stream is
// walked with a stride of Dv 16-byte vectors chosen so that
//       x^(128*Dv)  ==  x^k1 + x^k2 + ... + x^kh   (mod P),   h small, all k < 128:
// then "multiply the accumulator by x^(128*Dv)" (one Horner step down the stream) is h funnel shifts
// and an XOR per 32-bit word. Such strides are rare (about h-subsets(128) / 2^32 per candidate), so
// they are searched offline and compiled into the library as a table (csrc/sfold_table.inc).
//
// Residues use zlib's reflected convention: bit 31 = x^0, bit 0 = x^31.
// build: gcc -O2 -o /tmp/sfold_search tools/sfold_search.c ; run: /tmp/sfold_search > csrc/sfold_table.inc
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static uint32_t multmodp(uint32_t a, uint32_t b, uint32_t poly) {
    uint32_t m = 1u << 31, p = 0;
    for (;;) {
        if (a & m) {
            p ^= b;
            if ((a & (m - 1)) == 0) break;
        }
        m >>= 1;
        b = (b & 1) ? (b >> 1) ^ poly : b >> 1;
    }
    return p;
}

#define KMAX 128
#define HBITS 16
typedef struct { uint32_t key; int16_t a, b; uint8_t used; } Ent;  // a,b = term exponents (-1 = none)
static Ent tab[1 << HBITS];
static inline uint32_t hsh(uint32_t v) { return (v * 2654435761u) >> (32 - HBITS); }
static void put(uint32_t key, int a, int b) {
    uint32_t h = hsh(key);
    while (tab[h].used) {
        if (tab[h].key == key) return;  // keep the first (cheapest inserted first)
        h = (h + 1) & ((1u << HBITS) - 1);
    }
    tab[h].used = 1; tab[h].key = key; tab[h].a = (int16_t)a; tab[h].b = (int16_t)b;
}
static const Ent *get(uint32_t key) {
    uint32_t h = hsh(key);
    while (tab[h].used) {
        if (tab[h].key == key) return &tab[h];
        h = (h + 1) & ((1u << HBITS) - 1);
    }
    return NULL;
}

typedef struct { uint32_t dv; int h; int k[4]; int cost; } Sol;
static int cmp_int(const void *x, const void *y) { return *(const int *)x - *(const int *)y; }

// cost model of one fold step of a 16-byte accumulator: one SHF per (term, word) unless the shift is a
// multiple of 32, plus one SHFL per neighbour word needed
static int cost_of(const int *k, int h) {
    int c = 0, top = 0;
    for (int i = 0; i < h; i++) {
        if (k[i] & 31) c += 4;
        int need = (k[i] + 31) / 32;  // neighbour words reached
        if (need > top) top = need;
    }
    return c + 2 * top;
}

static void search(const char *name, uint32_t poly, uint32_t dv_lo, uint32_t dv_hi, int kmax) {
    static uint32_t t[KMAX];
    t[0] = 1u << 31;
    for (int k = 1; k < KMAX; k++) t[k] = (t[k - 1] & 1) ? (t[k - 1] >> 1) ^ poly : t[k - 1] >> 1;
    memset(tab, 0, sizeof tab);
    put(0, -1, -1);
    for (int a = 0; a < kmax; a++) put(t[a], a, -1);
    for (int a = 0; a < kmax; a++)
        for (int b = a + 1; b < kmax; b++) put(t[a] ^ t[b], a, b);
    // x^128
    uint32_t x128 = 1u << 31;
    for (int i = 0; i < 128; i++) x128 = (x128 & 1) ? (x128 >> 1) ^ poly : x128 >> 1;
    uint32_t r = 1u << 31;
    for (uint32_t dv = 1; dv < dv_lo; dv++) r = multmodp(r, x128, poly);
    Sol *sols = malloc(sizeof(Sol) * 100000);
    int ns = 0;
    for (uint32_t dv = dv_lo; dv <= dv_hi; dv++) {
        r = multmodp(r, x128, poly);  // x^(128*dv)
        Sol best; best.cost = 1 << 30;
        for (uint32_t i = 0; i < (1u << HBITS); i++) {
            if (!tab[i].used) continue;
            const Ent *e = get(r ^ tab[i].key);
            if (!e) continue;
            int k[4], h = 0;
            if (tab[i].a >= 0) k[h++] = tab[i].a;
            if (tab[i].b >= 0) k[h++] = tab[i].b;
            if (e->a >= 0) k[h++] = e->a;
            if (e->b >= 0) k[h++] = e->b;
            qsort(k, h, sizeof(int), cmp_int);
            int dup = 0;
            for (int j = 1; j < h; j++) dup |= (k[j] == k[j - 1]);
            if (dup || h == 0) continue;
            int c = cost_of(k, h);
            if (c < best.cost) { best.cost = c; best.h = h; best.dv = dv; memcpy(best.k, k, sizeof k); }
        }
        if (best.cost < (1 << 30)) sols[ns++] = best;
    }
    printf("// %s (reflected 0x%08X): %d strides in [%u, %u] with x^(128*Dv) = sum of <= 4 terms x^k, k < %d\n", name, poly, ns, dv_lo, dv_hi, kmax);
    printf("// { Dv, h, {k...}, cost }\n");
    for (int i = 0; i < ns; i++) {
        // verify
        uint32_t chk = 0;
        for (int j = 0; j < sols[i].h; j++) chk ^= t[sols[i].k[j]];
        uint32_t want = 1u << 31;
        for (uint32_t d = 0; d < sols[i].dv; d++) want = multmodp(want, x128, poly);
        if (chk != want) { fprintf(stderr, "verify failed\n"); exit(1); }
        printf("    {%6u, %d, {%3d, %3d, %3d, %3d}, %2d},\n", sols[i].dv, sols[i].h, sols[i].k[0], sols[i].h > 1 ? sols[i].k[1] : -1,
               sols[i].h > 2 ? sols[i].k[2] : -1, sols[i].h > 3 ? sols[i].k[3] : -1, sols[i].cost);
    }
    free(sols);
}

int main(int argc, char **argv) {
    int kmax = argc > 1 ? atoi(argv[1]) : 128;
    uint32_t hi = argc > 2 ? (uint32_t)atoi(argv[2]) : 65536;
    search("IEEE", 0xEDB88320u, 256, hi, kmax);
    search("CASTAGNOLI", 0x82F63B78u, 256, hi, kmax);
    return 0;
}
