//! dfs/gpu_ec — safe wrappers over libgpu_ec's C ABI with the signatures the DFS already uses.
//! SOURCE ONLY: never compiled here (no Rust toolchain in the build image).
//!
//! Drop-in points (reference file:line):
//!   dfs_common::erasure::{encode, decode, shard_len}      dfs/common/src/erasure.rs:7,30,56
//!   ReedSolomon::new(k,m)?.reconstruct(&mut opt_shards)   dfs/chunkserver/src/chunkserver.rs:620-623
//!   MyChunkServer::calculate_checksums / verify_block     dfs/chunkserver/src/chunkserver.rs:182,238
//!   crc32fast::Hasher::{new,update,finalize}              chunkserver.rs:746-750 and 7 other sites
use anyhow::{anyhow, bail, Result};
use std::ffi::CStr;
use std::os::raw::{c_char, c_int, c_void};

#[allow(non_camel_case_types)]
mod ffi {
    use super::*;
    extern "C" {
        pub fn gpu_ec_strerror(status: c_int) -> *const c_char;
        pub fn gpu_ec_last_error() -> *const c_char;
        pub fn gpu_ec_shard_len(len: usize, k: usize) -> usize;
        pub fn gpu_ec_encode(data: *const u8, len: usize, k: c_int, m: c_int, shards_out: *const *mut u8,
                             shard_crc_out: *mut u32, full_crc_out: *mut u32, meta_out: *const *mut u8) -> c_int;
        pub fn gpu_ec_reconstruct(shards: *const *mut u8, present: *const u8, shard_len: usize, k: c_int, m: c_int) -> c_int;
        pub fn gpu_ec_reconstruct_only(shards: *const *mut u8, present: *const u8, shard_len: usize, k: c_int, m: c_int,
                                       want: *const c_int, n_want: c_int) -> c_int;
        pub fn gpu_ec_verify_chunks_batch(n_blocks: usize, data: *const *const u8, lens: *const usize, chunk: usize, poly: c_int,
                                          meta_be: *const *const u8, meta_lens: *const usize, status_out: *mut c_int,
                                          first_bad_out: *mut i64) -> c_int;
        pub fn gpu_ec_crc32(data: *const u8, len: usize, poly: c_int, out: *mut u32) -> c_int;
        pub fn gpu_ec_crc32_chunks(data: *const u8, len: usize, chunk: usize, poly: c_int, out: *mut u32) -> c_int;
        pub fn gpu_ec_verify_chunks(data: *const u8, len: usize, chunk: usize, poly: c_int, meta_be: *const u8,
                                    meta_len: usize, first_bad: *mut i64) -> c_int;
        pub fn gpu_ec_alloc_pinned(ptr: *mut *mut c_void, bytes: usize) -> c_int;
        pub fn gpu_ec_free_pinned(ptr: *mut c_void) -> c_int;
    }
}

const ERR_COUNT_MISMATCH: c_int = -7;
const ERR_CRC_MISMATCH: c_int = -8;
pub const CHECKSUM_CHUNK_SIZE: usize = 512; // chunkserver.rs:16

fn msg(status: c_int) -> String {
    unsafe {
        let base = CStr::from_ptr(ffi::gpu_ec_strerror(status)).to_string_lossy().into_owned();
        let detail = CStr::from_ptr(ffi::gpu_ec_last_error()).to_string_lossy();
        if status == -9 && !detail.is_empty() { format!("{base}: {detail}") } else { base }
    }
}

/// dfs/common/src/erasure.rs:56
pub fn shard_len(data_len: usize, data_shards: usize) -> usize {
    assert!(data_shards > 0, "data_shards must be > 0");
    unsafe { ffi::gpu_ec_shard_len(data_len, data_shards) }
}

/// dfs/common/src/erasure.rs:7 — same signature, same error text.
pub fn encode(data: &[u8], data_shards: usize, parity_shards: usize) -> Result<Vec<Vec<u8>>> {
    if data_shards == 0 || parity_shards == 0 { bail!("data_shards and parity_shards must both be > 0"); }
    if data.is_empty() { bail!("data must not be empty"); }
    let s = shard_len(data.len(), data_shards);
    let mut shards: Vec<Vec<u8>> = (0..data_shards + parity_shards).map(|_| vec![0u8; s]).collect();
    let ptrs: Vec<*mut u8> = shards.iter_mut().map(|v| v.as_mut_ptr()).collect();
    let rc = unsafe {
        ffi::gpu_ec_encode(data.as_ptr(), data.len(), data_shards as c_int, parity_shards as c_int, ptrs.as_ptr(),
                           std::ptr::null_mut(), std::ptr::null_mut(), std::ptr::null())
    };
    if rc != 0 { return Err(anyhow!("{}", msg(rc))); }
    Ok(shards)
}

/// The client write path in one call (dfs/client/src/mod.rs:326-341, 600-611).
pub fn encode_with_checksums(data: &[u8], k: usize, m: usize) -> Result<(Vec<Vec<u8>>, Vec<u32>, u32)> {
    if k == 0 || m == 0 { bail!("data_shards and parity_shards must both be > 0"); }
    if data.is_empty() { bail!("data must not be empty"); }
    let s = shard_len(data.len(), k);
    let mut shards: Vec<Vec<u8>> = (0..k + m).map(|_| vec![0u8; s]).collect();
    let ptrs: Vec<*mut u8> = shards.iter_mut().map(|v| v.as_mut_ptr()).collect();
    let mut crcs = vec![0u32; k + m];
    let mut full = 0u32;
    let rc = unsafe {
        ffi::gpu_ec_encode(data.as_ptr(), data.len(), k as c_int, m as c_int, ptrs.as_ptr(), crcs.as_mut_ptr(),
                           &mut full, std::ptr::null())
    };
    if rc != 0 { return Err(anyhow!("{}", msg(rc))); }
    Ok((shards, crcs, full))
}

/// ReedSolomon::reconstruct as called at chunkserver.rs:620-623 and erasure.rs:38.
pub fn reconstruct(shards: &mut [Option<Vec<u8>>], k: usize, m: usize) -> Result<()> {
    if k == 0 || m == 0 { bail!("data_shards and parity_shards must both be > 0"); }
    if shards.len() != k + m { bail!("RS reconstruct error: TooFewShards"); }
    let s = shards.iter().flatten().map(|v| v.len()).next().unwrap_or(0);
    if shards.iter().flatten().any(|v| v.len() != s) { bail!("RS reconstruct error: IncorrectShardSize"); }
    let present: Vec<u8> = shards.iter().map(|o| o.is_some() as u8).collect();
    for o in shards.iter_mut() { if o.is_none() { *o = Some(vec![0u8; s]); } }
    let ptrs: Vec<*mut u8> = shards.iter_mut().map(|o| o.as_mut().unwrap().as_mut_ptr()).collect();
    let rc = unsafe { ffi::gpu_ec_reconstruct(ptrs.as_ptr(), present.as_ptr(), s, k as c_int, m as c_int) };
    if rc != 0 {
        for (o, p) in shards.iter_mut().zip(&present) { if *p == 0 { *o = None; } }
        return Err(anyhow!("{}", msg(rc)));
    }
    Ok(())
}

/// The repair path (chunkserver.rs:503-640) keeps only `opt_shards[shard_index]`: rebuild just that slot.
pub fn reconstruct_one(shards: &mut [Option<Vec<u8>>], k: usize, m: usize, shard_index: usize) -> Result<()> {
    if shards.len() != k + m { bail!("ec_shard_sources length {} != total shards {}", shards.len(), k + m); }
    shards[shard_index] = None; // chunkserver.rs:529
    let available = shards.iter().filter(|s| s.is_some()).count();
    if available < k { bail!("Only {} shards available, need at least {} for reconstruction", available, k); }
    let s = shards.iter().flatten().map(|v| v.len()).next().unwrap_or(0);
    let present: Vec<u8> = shards.iter().map(|o| o.is_some() as u8).collect();
    let mut target = vec![0u8; s];
    let ptrs: Vec<*mut u8> = shards.iter_mut().enumerate().map(|(i, o)| match o {
        Some(v) => v.as_mut_ptr(),
        None if i == shard_index => target.as_mut_ptr(),
        None => std::ptr::null_mut(), // absent and not wanted: gpu_ec_reconstruct_only neither reads nor writes it
    }).collect();
    let want = [shard_index as c_int];
    let rc = unsafe { ffi::gpu_ec_reconstruct_only(ptrs.as_ptr(), present.as_ptr(), s, k as c_int, m as c_int, want.as_ptr(), 1) };
    if rc != 0 { return Err(anyhow!("{}", msg(rc))); }
    shards[shard_index] = Some(target);
    Ok(())
}

/// One pass of run_background_scrubber (chunkserver.rs:652-701) as ONE batched verify.
/// Returns, per block, Ok(()) or the same Err text verify_block gives.
pub fn verify_blocks(blocks: &[&[u8]], metas: &[&[u8]]) -> Vec<std::result::Result<(), String>> {
    let n = blocks.len();
    let dp: Vec<*const u8> = blocks.iter().map(|b| b.as_ptr()).collect();
    let dl: Vec<usize> = blocks.iter().map(|b| b.len()).collect();
    let mp: Vec<*const u8> = metas.iter().map(|b| b.as_ptr()).collect();
    let ml: Vec<usize> = metas.iter().map(|b| b.len()).collect();
    let mut st = vec![0 as c_int; n];
    let mut bad = vec![-1i64; n];
    let rc = unsafe {
        ffi::gpu_ec_verify_chunks_batch(n, dp.as_ptr(), dl.as_ptr(), CHECKSUM_CHUNK_SIZE, 0, mp.as_ptr(), ml.as_ptr(),
                                        st.as_mut_ptr(), bad.as_mut_ptr())
    };
    (0..n).map(|i| match if rc != 0 { rc } else { st[i] } {
        0 => Ok(()),
        ERR_COUNT_MISMATCH => Err("Checksum count mismatch".to_string()),
        ERR_CRC_MISMATCH => Err(format!("Checksum mismatch at chunk {}", bad[i])),
        other => Err(msg(other)),
    }).collect()
}

/// dfs/common/src/erasure.rs:30
pub fn decode(shards: &mut [Option<Vec<u8>>], k: usize, m: usize, original_len: usize) -> Result<Vec<u8>> {
    reconstruct(shards, k, m)?;
    let mut result = Vec::new();
    for (i, s) in shards[..k].iter().enumerate() {
        let s = s.as_ref().ok_or_else(|| anyhow!("data shard {} missing after reconstruction", i))?;
        result.extend_from_slice(s);
    }
    result.truncate(original_len);
    Ok(result)
}

/// crc32fast::Hasher::new(); update(data); finalize()
pub fn crc32(data: &[u8]) -> u32 {
    let mut out = 0u32;
    let rc = unsafe { ffi::gpu_ec_crc32(data.as_ptr(), data.len(), 0, &mut out) };
    assert_eq!(rc, 0, "{}", msg(rc)); // the reference call is infallible; no CPU fallback exists
    out
}

/// MyChunkServer::calculate_checksums (chunkserver.rs:182)
pub fn calculate_checksums(data: &[u8]) -> Vec<u32> {
    let mut out = vec![0u32; data.len().div_ceil(CHECKSUM_CHUNK_SIZE)];
    let rc = unsafe { ffi::gpu_ec_crc32_chunks(data.as_ptr(), data.len(), CHECKSUM_CHUNK_SIZE, 0, out.as_mut_ptr()) };
    assert_eq!(rc, 0, "{}", msg(rc));
    out
}

/// The compare of MyChunkServer::verify_block (chunkserver.rs:265-289) given the .meta bytes.
pub fn verify_chunks(data: &[u8], meta: &[u8]) -> std::result::Result<(), String> {
    let mut bad: i64 = -1;
    let rc = unsafe {
        ffi::gpu_ec_verify_chunks(data.as_ptr(), data.len(), CHECKSUM_CHUNK_SIZE, 0, meta.as_ptr(), meta.len(), &mut bad)
    };
    match rc {
        0 => Ok(()),
        ERR_COUNT_MISMATCH => Err("Checksum count mismatch".to_string()),
        ERR_CRC_MISMATCH => Err(format!("Checksum mismatch at chunk {}", bad)),
        other => Err(msg(other)),
    }
}

/// Page-locked buffer for shard / block bytes so H2D/D2H run at full PCIe rate.
pub struct PinnedBuf { ptr: *mut u8, len: usize }
impl PinnedBuf {
    pub fn new(len: usize) -> Result<Self> {
        let mut p: *mut c_void = std::ptr::null_mut();
        let rc = unsafe { ffi::gpu_ec_alloc_pinned(&mut p, len) };
        if rc != 0 { return Err(anyhow!("{}", msg(rc))); }
        Ok(Self { ptr: p as *mut u8, len })
    }
    pub fn as_mut_slice(&mut self) -> &mut [u8] { unsafe { std::slice::from_raw_parts_mut(self.ptr, self.len) } }
    pub fn as_slice(&self) -> &[u8] { unsafe { std::slice::from_raw_parts(self.ptr, self.len) } }
}
impl Drop for PinnedBuf { fn drop(&mut self) { unsafe { ffi::gpu_ec_free_pinned(self.ptr as *mut c_void); } } }
unsafe impl Send for PinnedBuf {}
