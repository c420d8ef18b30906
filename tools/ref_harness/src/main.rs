//! ref_harness — the reference's own EC + checksum calls on caller-supplied bytes, results as hex.
//!
//! Mirrors dfs/common/src/erasure.rs:7-49 (encode: pad, split, ReedSolomon::encode; decode:
//! reconstruct, concat, truncate) and the crc32fast::Hasher sites (chunkserver.rs:182-190 per 512 B,
//! :746-750 whole buffer) with nothing but the two pinned crates, so its output IS the reference's.
//!
//!   ref_harness encode <k> <m> <hex data>            -> one line per shard: "<index> <hex> <crc32>"
//!   ref_harness decode <k> <m> <len> <hex|-> ...     -> the recovered data as hex ("-" = erased shard)
//!   ref_harness crc <chunk> <hex data>               -> whole-buffer CRC, then one CRC per chunk
//!   ref_harness bench <k> <m> <bytes> <iters>        -> seconds per encode (single thread, as the reference runs it)
//!
//! tools/ref_harness/make_ref_golden.py drives it over the inputs of tests/golden/golden.json and writes
//! tests/golden/ref_golden.json; tests/test_oracle_golden.py checks the oracle against that file when it exists.
use reed_solomon_erasure::galois_8::ReedSolomon;

fn unhex(s: &str) -> Vec<u8> {
    (0..s.len() / 2).map(|i| u8::from_str_radix(&s[2 * i..2 * i + 2], 16).unwrap()).collect()
}
fn hex(b: &[u8]) -> String {
    b.iter().map(|x| format!("{:02x}", x)).collect()
}
fn crc(b: &[u8]) -> u32 {
    let mut h = crc32fast::Hasher::new();
    h.update(b);
    h.finalize()
}
// dfs/common/src/erasure.rs:7-24
fn encode(data: &[u8], k: usize, m: usize) -> Vec<Vec<u8>> {
    let shard_size = (data.len() + k - 1) / k;
    let mut padded = data.to_vec();
    padded.resize(shard_size * k, 0);
    let mut shards: Vec<Vec<u8>> = padded.chunks(shard_size).map(|c| c.to_vec()).collect();
    for _ in 0..m {
        shards.push(vec![0u8; shard_size]);
    }
    ReedSolomon::new(k, m).unwrap().encode(&mut shards).unwrap();
    shards
}

fn main() {
    let a: Vec<String> = std::env::args().collect();
    match a[1].as_str() {
        "encode" => {
            let (k, m) = (a[2].parse().unwrap(), a[3].parse().unwrap());
            for (i, s) in encode(&unhex(&a[4]), k, m).iter().enumerate() {
                println!("{} {} {}", i, hex(s), crc(s));
            }
        }
        "decode" => {
            // dfs/common/src/erasure.rs:30-49
            let (k, m, len): (usize, usize, usize) = (a[2].parse().unwrap(), a[3].parse().unwrap(), a[4].parse().unwrap());
            let mut shards: Vec<Option<Vec<u8>>> = a[5..].iter().map(|s| if s == "-" { None } else { Some(unhex(s)) }).collect();
            ReedSolomon::new(k, m).unwrap().reconstruct(&mut shards).unwrap();
            let mut out = Vec::new();
            for s in shards.iter().take(k) {
                out.extend_from_slice(s.as_ref().unwrap());
            }
            out.truncate(len);
            println!("{}", hex(&out));
        }
        "crc" => {
            let chunk: usize = a[2].parse().unwrap();
            let d = unhex(&a[3]);
            println!("{}", crc(&d));
            for c in d.chunks(chunk) {
                println!("{}", crc(c));
            }
        }
        "bench" => {
            let (k, m, n, it): (usize, usize, usize, usize) = (a[2].parse().unwrap(), a[3].parse().unwrap(), a[4].parse().unwrap(), a[5].parse().unwrap());
            let d: Vec<u8> = (0..n).map(|i| (i * 2654435761usize >> 13) as u8).collect();
            let t = std::time::Instant::now();
            for _ in 0..it {
                std::hint::black_box(encode(&d, k, m));
            }
            println!("{}", t.elapsed().as_secs_f64() / it as f64);
        }
        _ => panic!("unknown command"),
    }
}
