#!/usr/bin/env python3
"""CPU model of the sparse-fold CRC (csrc/sfold_kernels.cuh), word for word, checked against zlib.

The stream is walked as R rows of Dv 16-byte vectors (rows aligned to the END of the stream, leading
virtual zeros in front). Column c keeps a 128-bit accumulator; one Horner step down the stream
multiplies every accumulator by x^(128*Dv) == x^k1 + ... + x^kh (mod P): h funnel shifts per word,
bits carried from the next column of the same 32-column group (a warp). The bits that leave a group
at its low end are not reduced in the loop: the group's first column logs its two low words before
each step, and the finalizer replays them as virtual data words sitting just before the group in
that row. The accumulators of the last row are a Dv*16-byte message whose plain CRC finishes the job.

usage: python tools/sfold_model.py   (self-test, prints OK)
"""
import zlib

import numpy as np

POLY = 0xEDB88320


def multmodp(a, b, poly=POLY):
    if a == 0:
        return 0
    m, p = 1 << 31, 0
    while True:
        if a & m:
            p ^= b
            if a & (m - 1) == 0:
                break
        m >>= 1
        b = (b >> 1) ^ poly if b & 1 else b >> 1
    return p


def xpow(n, poly=POLY):
    """x^n mod P, n >= 0 (reflected: bit 31 = x^0)."""
    r, b = 1 << 31, 1 << 30
    while n:
        if n & 1:
            r = multmodp(r, b, poly)
        b = multmodp(b, b, poly)
        n >>= 1
    return r


def xinv_pow(n, poly=POLY):
    """x^(-n) mod P via the order of x: x^(2^32-1) = 1 for IEEE; in general use x^(ord - n)."""
    # the multiplicative order of x divides 2^32-1 only if P is primitive; be general: invert by Fermat on
    # the factor structure is overkill here -> x^-1 = (1 + P_low)/x exists because P(0) = 1:
    # x * inv = 1 (mod P)  =>  inv = (P - 1) / x  = P_low(x)/x with P = x^32 + P_low, P_low(0) = 1
    # reflected poly word holds coefficients of x^31..x^0 at bits 0..31 (bit 31 = x^0), plus the implicit x^32.
    # (P + 1)/x: coefficients shift down by one degree: x^32 -> x^31, x^k -> x^(k-1), constant term cancelled.
    inv1 = ((poly & 0x7FFFFFFF) << 1) | 1  # degree d>=1 terms move to d-1; x^32 -> x^31 (bit 0)
    r, b = 1 << 31, inv1
    while n:
        if n & 1:
            r = multmodp(r, b, poly)
        b = multmodp(b, b, poly)
        n >>= 1
    return r


def crc0_words(words, poly=POLY):
    """crc with init 0 / xorout 0 of a little-endian word array (bitwise, slow, model only)."""
    c = 0
    for w in words:
        c ^= int(w)
        for _ in range(32):
            c = (c >> 1) ^ poly if c & 1 else c >> 1
    return c


def funnel_r(lo, hi, s):
    return ((lo >> s) | (hi << (32 - s))) & 0xFFFFFFFF if s else lo


def sfold_crc(data: bytes, Dv, ks, poly=POLY):
    S = len(data)
    nvec = (S + 15) // 16
    padb = nvec * 16 - S
    buf = np.frombuffer(data + bytes(padb), dtype="<u4").reshape(nvec, 4)
    R = -(-nvec // Dv)
    lead = R * Dv - nvec
    U = np.zeros((R * Dv, 4), np.uint32)
    U[lead:] = buf
    U = U.reshape(R, Dv, 4)
    NG = -(-Dv // 32)
    acc = [[0, 0, 0, 0] for _ in range(Dv)]
    queue = np.zeros((R, NG, 2), np.uint32)
    qs = [(k >> 5, k & 31) for k in ks]
    assert max(ks) < 64
    for s in range(R):
        new = [None] * Dv
        for g in range(NG):
            c0, c1 = 32 * g, min(32 * g + 32, Dv)
            queue[s, g] = acc[c0][0], acc[c0][1]
            for c in range(c0, c1):
                nxt = acc[c + 1] if c + 1 < c1 else [0, 0, 0, 0]
                w = acc[c] + nxt
                out = []
                for i in range(4):
                    v = 0
                    for q, sh in qs:
                        v ^= funnel_r(w[i + q], w[i + q + 1], sh)
                    out.append(v ^ int(U[s, c, i]))
                new[c] = out
        acc = new
    # ---- finalize
    dump = [w for c in range(Dv) for w in acc[c]]
    total0 = crc0_words(dump, poly)
    xD = xpow(128 * Dv, poly)
    x32 = xpow(32, poly)
    x4096 = xpow(4096, poly)
    qsum = 0
    for g in range(NG):
        T = 0
        for s in range(1, R):
            w = [0, 0, int(queue[s, g, 0]), int(queue[s, g, 1]), 0]  # w[i+2] = word i of the first column; w[-1], w[-2] = 0
            hm1 = hm2 = 0
            for q, sh in qs:
                hm1 ^= funnel_r(w[-1 + q + 2], w[-1 + q + 3], sh)
                hm2 ^= funnel_r(w[-2 + q + 2], w[-2 + q + 3], sh)
            hc = hm1 ^ multmodp(hm2, x32, poly)
            T = multmodp(T, xD, poly) ^ hc
        qsum = multmodp(qsum, x4096, poly) ^ T
    tail_cols = Dv - 32 * (NG - 1)                     # weight of the last group's start: x^(128*tail_cols)
    qsum = multmodp(qsum, xpow(128 * tail_cols + 32, poly), poly)
    total0 ^= qsum
    real0 = multmodp(total0, xinv_pow(8 * padb, poly), poly)
    return real0 ^ (zlib.crc32(bytes(S)) if poly == POLY else None)


if __name__ == "__main__":
    rng = np.random.default_rng(1)
    assert multmodp(xpow(77), xinv_pow(77)) == 1 << 31
    # toy strides: any Dv works with the dense K = x^(128*Dv) mod P itself (terms = its set bits, all < 32)
    for S, Dv in [(16 * 40 * 7, 40), (10_001, 33), (50_000, 100), (12_345, 64), (777, 70), (15, 5), (16, 1), (4097, 32)]:
        d = rng.integers(0, 256, S, dtype=np.uint8).tobytes()
        r = xpow(128 * Dv)
        ks = [31 - b for b in range(32) if (r >> b) & 1]
        got = sfold_crc(d, Dv, ks)
        print(S, Dv, len(ks), hex(got), "OK" if got == zlib.crc32(d) else "MISMATCH")
        assert got == zlib.crc32(d)
    # real strides from tools/sfold_search.c
    for S, Dv, ks in [(16 * 11102 * 3 + 5, 11102, [3, 14, 33, 45]), (16 * 1651 * 5 - 3, 1651, [3, 22, 38, 49])]:
        d = rng.integers(0, 256, S, dtype=np.uint8).tobytes()
        got = sfold_crc(d, Dv, ks)
        print(S, Dv, hex(got), hex(zlib.crc32(d)), "OK" if got == zlib.crc32(d) else "MISMATCH")
