#!/usr/bin/env python3
"""Generate tests/golden/golden.json — an INDEPENDENT pure-Python restatement of the path.

The reference's own code cannot run here (Rust, no cargo, crates un-vendored),
so these vectors come from a second, deliberately different implementation of the published
algorithms (bitwise carry-less GF multiply instead of log/exp tables, Fraction-free Gauss
elimination written separately, bit-at-a-time CRC, zlib as a third opinion). The C oracle and
the CUDA path are both checked against this file. EXTERNAL fixed points (not produced by this
script, only re-checked by it) are marked "external": the Backblaze JavaReedSolomon
`testOneEncode` vector and the CRC catalogue check values.

Run:  python tests/golden/gen_golden.py   (rewrites golden.json; deterministic)
"""
import json
import os
import random
import zlib

POLY = 0x11D


def gmul(a, b):
    r = 0
    while b:
        if b & 1:
            r ^= a
        a <<= 1
        if a & 0x100:
            a ^= POLY
        b >>= 1
    return r


def gpow(a, n):
    r = 1
    for _ in range(n):
        r = gmul(r, a)
    return r


def ginv(a):
    assert a
    return gpow(a, 254)


def mat_mul(A, B):
    return [[_dot(row, [B[t][c] for t in range(len(B))]) for c in range(len(B[0]))] for row in A]


def _dot(u, v):
    r = 0
    for x, y in zip(u, v):
        r ^= gmul(x, y)
    return r


def mat_inv(A):
    n = len(A)
    W = [list(A[i]) + [1 if i == j else 0 for j in range(n)] for i in range(n)]
    for c in range(n):
        p = next(r for r in range(c, n) if W[r][c])
        W[c], W[p] = W[p], W[c]
        s = ginv(W[c][c])
        W[c] = [gmul(s, x) for x in W[c]]
        for r in range(n):
            if r != c and W[r][c]:
                f = W[r][c]
                W[r] = [x ^ gmul(f, y) for x, y in zip(W[r], W[c])]
    return [row[n:] for row in W]


def rs_matrix(k, m):
    V = [[gpow(r, c) for c in range(k)] for r in range(k + m)]
    return mat_mul(V, mat_inv(V[:k]))


def encode(data, k, m):
    S = -(-len(data) // k)
    padded = bytes(data) + bytes(S * k - len(data))
    shards = [padded[i * S:(i + 1) * S] for i in range(k)]
    M = rs_matrix(k, m)
    for p in range(m):
        row = M[k + p]
        shards.append(bytes(_dot(row, [shards[j][i] for j in range(k)]) for i in range(S)))
    return shards


def decode_rows(k, m, missing):
    M = rs_matrix(k, m)
    surv = [i for i in range(k + m) if i not in missing][:k]
    inv = mat_inv([M[i] for i in surv])
    rows = []
    for idx in missing:
        rows.append(inv[idx] if idx < k else mat_mul([M[idx]], inv)[0])
    return surv, rows


def crc_bitwise(data, poly_reflected):
    c = 0xFFFFFFFF
    for b in data:
        c ^= b
        for _ in range(8):
            c = (c >> 1) ^ (poly_reflected if c & 1 else 0)
    return c ^ 0xFFFFFFFF


def splitmix64(x):
    M = (1 << 64) - 1
    x = (x + 0x9E3779B97F4A7C15) & M
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & M
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & M
    return x ^ (x >> 31)


def synthetic(n, seed, off=0):
    out = bytearray()
    for i in range(off, off + n):
        out.append((splitmix64(seed ^ (i >> 3)) >> (8 * (i & 7))) & 0xFF)
    return bytes(out)


def main():
    g = {}
    g["field"] = {"2*128": gmul(2, 128), "3*7": gmul(3, 7), "0x53*0xCA": gmul(0x53, 0xCA)}
    g["parity_rows"] = {f"{k},{m}": [bytes(r).hex() for r in rs_matrix(k, m)[k:]]
                        for k, m in [(2, 2), (4, 2), (5, 5), (6, 3), (10, 4)]}
    # external: Backblaze JavaReedSolomon ReedSolomonTest.testOneEncode
    jrs_data = [[0, 1], [4, 5], [2, 3], [6, 7], [8, 9]]
    jrs_parity = [[12, 13], [10, 11], [14, 15], [90, 91], [94, 95]]
    flat = bytes(sum(jrs_data, []))
    got = encode(flat, 5, 5)
    assert [list(s) for s in got[5:]] == jrs_parity, "restatement disagrees with the external KAT"
    g["jrs_one_encode"] = {"external": True, "k": 5, "m": 5, "data": flat.hex(),
                           "parity": [bytes(p).hex() for p in jrs_parity]}
    hello = b"Hello, Erasure Coding World!"  # erasure.rs:67
    g["hello_4_2"] = {"data": hello.hex(), "shards": [s.hex() for s in encode(hello, 4, 2)]}
    sevens = encode(bytes([7]) * 4096, 4, 2)  # chunkserver.rs:1102
    g["sevens_4_2_parity_all_07"] = all(set(s) == {7} for s in sevens[4:])
    surv, rows = decode_rows(6, 3, [1, 4, 7])
    g["decode_6_3_erase_1_4_7"] = {"survivors": surv, "rows": [bytes(r).hex() for r in rows]}
    surv, rows = decode_rows(4, 2, [1, 4])  # erasure.rs:80-81
    g["decode_4_2_erase_1_4"] = {"survivors": surv, "rows": [bytes(r).hex() for r in rows]}

    # CRC: catalogue check values (external) + the 61-byte block of chunkserver.rs:1122
    chk = b"123456789"
    assert crc_bitwise(chk, 0xEDB88320) == 0xCBF43926 == zlib.crc32(chk)
    assert crc_bitwise(chk, 0x82F63B78) == 0xE3069283
    block61 = b"Hello, world! This is a test block for checksum verification."
    assert len(block61) == 61
    g["crc"] = {
        "external_check_ieee": 0xCBF43926, "external_check_castagnoli": 0xE3069283,
        "block61": block61.hex(), "block61_meta_be": crc_bitwise(block61, 0xEDB88320).to_bytes(4, "big").hex(),
    }

    rnd = random.Random(0xEC)
    cases = []
    for k, m, n in [(2, 2, 1), (4, 2, 28), (4, 2, 10000), (6, 3, 511), (6, 3, 513), (6, 3, 4096),
                    (10, 4, 1000), (5, 5, 77), (3, 1, 100), (1, 1, 17), (12, 4, 2500), (8, 8, 640)]:
        data = bytes(rnd.getrandbits(8) for _ in range(n)) if n != 10000 else bytes(i % 256 for i in range(n))
        sh = encode(data, k, m)
        cases.append({"k": k, "m": m, "data": data.hex(), "parity": [s.hex() for s in sh[k:]],
                      "shard_crc_ieee": [zlib.crc32(s) for s in sh], "full_crc_ieee": zlib.crc32(data)})
    g["encode_cases"] = cases

    crc_cases = []
    for n in [0, 1, 15, 16, 17, 171, 511, 512, 513, 1024, 1537, 4096, 65536 + 5]:
        data = bytes(rnd.getrandbits(8) for _ in range(n))
        chunks = [data[i:i + 512] for i in range(0, n, 512)]
        crc_cases.append({"data": data.hex(), "ieee": zlib.crc32(data),
                          "castagnoli": crc_bitwise(data, 0x82F63B78) if n <= 4096 else None,
                          "chunks512_ieee": [zlib.crc32(c) for c in chunks],
                          "meta_be": b"".join(zlib.crc32(c).to_bytes(4, "big") for c in chunks).hex()})
    g["crc_cases"] = crc_cases
    g["synthetic"] = {"seed": 0xEC000001, "first64": synthetic(64, 0xEC000001).hex(),
                      "at_1000003_len_21": synthetic(21, 0xEC000001, 1000003).hex()}

    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden.json")
    with open(out, "w") as f:
        json.dump(g, f, indent=1, sort_keys=True)
    print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
