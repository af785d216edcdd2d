//! Declarations of every entry point of `include/swb200.h` (same order as the header).
//! Not compiled in this repository's build image (no Rust toolchain); a test keeps the symbol list equal to the header's.
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_int, c_uint, c_void};

pub const SWB_GLOBAL: c_int = 0; // alignment_type, SA.rs:11-13
pub const SWB_GLOCAL: c_int = 1;
pub const SWB_LOCAL: c_int = 2;

pub const SWB_OK: c_int = 0;
pub const SWB_ERR_ARG: c_int = 1;
pub const SWB_ERR_SCALE: c_int = 2;
pub const SWB_ERR_LETTER: c_int = 3;
pub const SWB_ERR_CUDA: c_int = 4;
pub const SWB_ERR_EMPTY: c_int = 5;
pub const SWB_ERR_NOMEM: c_int = 6;
pub const SWB_ERR_RANGE: c_int = 7;
pub const SWB_RETAIN_ALL: c_uint = 1;
pub const SWB_MAX_TOPK: c_int = 4096;
pub const SWB_STATIC_MAX_PAIRS: i64 = 4096;
pub const SWB_STATIC_MAX_LEN: i64 = 1024;
pub const SWB_STATIC_MAX_RESIDUES: i64 = 1048576;

pub enum swb_engine {}
pub enum swb_batch {}
pub enum swb_db {}

#[repr(C)]
pub struct swb_matrix {
    pub n_letters: i32,
    pub letters: *const c_char,
    pub scores: *const f32,
}

#[repr(C)]
pub struct swb_alignment {
    pub score: f32,
    pub end_a: i64,
    pub end_b: i64,
    pub n_cols: i64,
    pub idx_a: *mut i64,
    pub idx_b: *mut i64,
    pub aln_len: i64,
    pub aln_a: *mut c_char,
    pub aln_b: *mut c_char,
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct swb_stats {
    pub cells: f64,
    pub kernel_ms: f64,
    pub h2d_bytes: f64,
    pub d2h_bytes: f64,
    pub launches: i64,
    pub i32_reruns: i64,
    pub scale: i32,
    pub exact_scale: i32,
    pub ckpt_pairs: i64,
    pub k3_pairs: i64,
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct swb_aln_stats {
    pub score: f32,
    pub status: i32,
    pub end_a: i64,
    pub end_b: i64,
    pub aln_len: i64,
    pub a_letters: i64,
    pub b_letters: i64,
    pub matches: i64,
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct swb_hit {
    pub score: f32,
    pub shard: i32,
    pub index: i64,
    pub end_q: i64,
    pub end_t: i64,
}

extern "C" {
    pub fn swb_matrix_blosum62(letters: *mut c_char, scores: *mut f32) -> i32;
    pub fn swb_matrix_match_mismatch(m: f32, mismatch: f32, letters: *mut c_char, scores: *mut f32) -> i32;
    pub fn swb_engine_create(m: *const swb_matrix, gap_open: f32, gap_extend: f32, mode: c_int, device: c_int, out: *mut *mut swb_engine) -> c_int;
    pub fn swb_engine_destroy(e: *mut swb_engine);
    pub fn swb_last_error(e: *const swb_engine) -> *const c_char;
    pub fn swb_get_stats(e: *const swb_engine, out: *mut swb_stats) -> c_int;
    pub fn swb_device_count() -> c_int;
    pub fn swb_align(e: *mut swb_engine, a: *const c_char, la: i64, b: *const c_char, lb: i64, retain_all: c_int, out: *mut swb_alignment) -> c_int;
    pub fn swb_align_partial(e: *mut swb_engine, a: *const c_char, la: i64, b: *const c_char, lb: i64, pa: i64, pb: i64, retain_all: c_int, score_only: c_int, out: *mut swb_alignment) -> c_int;
    pub fn swb_align_pssm(e: *mut swb_engine, scores: *const f32, a: *const c_char, la: i64, b: *const c_char, lb: i64, retain_all: c_int, out: *mut swb_alignment) -> c_int;
    pub fn swb_align_batch(e: *mut swb_engine, a_all: *const c_char, off_a: *const i64, b_all: *const c_char, off_b: *const i64, n: i64, flags: c_uint, outs: *mut swb_alignment) -> c_int;
    pub fn swb_alignment_free(a: *mut swb_alignment);
    pub fn swb_align_stats_batch(e: *mut swb_engine, a_all: *const c_char, off_a: *const i64, b_all: *const c_char, off_b: *const i64, n: i64, flags: c_uint, outs: *mut swb_aln_stats) -> c_int;
    pub fn swb_cluster(e: *mut swb_engine, seq_all: *const c_char, off: *const i64, n: i64, identity: f64, coverage_short: f64, coverage_long: f64, order: *mut i64, cluster_of: *mut i64, identical: *mut i64, n_alignments: *mut i64) -> c_int;
    pub fn swb_score_batch(e: *mut swb_engine, a_all: *const c_char, off_a: *const i64, b_all: *const c_char, off_b: *const i64, n: i64, scores: *mut f32, end_a: *mut i64, end_b: *mut i64) -> c_int;
    /// No counterpart in the reference: swb_score_batch on DEVICE buffers with
    /// the kernel's scratch in a static __device__ array; allocates nothing.  Limits: SWB_STATIC_MAX_PAIRS / _LEN / _RESIDUES.
    pub fn swb_score_batch_static(e: *mut swb_engine, d_a_all: *const c_char, d_off_a: *const i64, total_a: i64, d_b_all: *const c_char, d_off_b: *const i64, total_b: i64, n: i64, d_scores: *mut f32, d_end_a: *mut i64, d_end_b: *mut i64, d_status: *mut i32, stream: *mut c_void) -> c_int;
    pub fn swb_search_scores(e: *mut swb_engine, query: *const c_char, lq: i64, db_all: *const c_char, off_db: *const i64, n: i64, scores: *mut f32, end_q: *mut i64, end_t: *mut i64) -> c_int;
    pub fn swb_db_shard(off: *const i64, n: i64, rank: c_int, world: c_int, idx: *mut i64, n_idx: *mut i64) -> c_int;
    pub fn swb_db_upload(e: *mut swb_engine, letters: *const c_char, off: *const i64, n: i64, global_index: *const i64, out: *mut *mut swb_db) -> c_int;
    pub fn swb_db_upload_shard(e: *mut swb_engine, letters: *const c_char, off: *const i64, n_total: i64, rank: c_int, world: c_int, out: *mut *mut swb_db) -> c_int;
    pub fn swb_db_free(e: *mut swb_engine, db: *mut swb_db);
    pub fn swb_db_count(db: *const swb_db) -> i64;
    pub fn swb_db_residues(db: *const swb_db) -> i64;
    pub fn swb_search(e: *mut swb_engine, query: *const c_char, lq: i64, db: *mut swb_db, topk: c_int, hits: *mut swb_hit, n_hits: *mut i64) -> c_int;
    pub fn swb_db_scores(e: *mut swb_engine, db: *mut swb_db, scores: *mut f32, end_q: *mut i64, end_t: *mut i64) -> c_int;
    pub fn swb_search_merge(parts: *const swb_hit, n_parts: i64, topk: c_int, out: *mut swb_hit, n_out: *mut i64) -> c_int;
    pub fn swb_search_align(e: *mut swb_engine, query: *const c_char, lq: i64, db: *mut swb_db, hits: *const swb_hit, n_hits: i64, flags: c_uint, alns: *mut swb_alignment, owned: *mut i32) -> c_int;
    pub fn swb_search_sharded(engines: *mut *mut swb_engine, dbs: *mut *mut swb_db, n_shards: c_int, query: *const c_char, lq: i64, topk: c_int, flags: c_uint, hits: *mut swb_hit, n_hits: *mut i64, alns: *mut swb_alignment) -> c_int;
    pub fn swb_batch_upload(e: *mut swb_engine, a_all: *const c_char, off_a: *const i64, b_all: *const c_char, off_b: *const i64, n: i64, out: *mut *mut swb_batch) -> c_int;
    pub fn swb_batch_run(e: *mut swb_engine, b: *mut swb_batch) -> c_int;
    pub fn swb_batch_fetch(e: *mut swb_engine, b: *mut swb_batch, scores: *mut f32, end_a: *mut i64, end_b: *mut i64) -> c_int;
    pub fn swb_batch_free(e: *mut swb_engine, b: *mut swb_batch);
    pub fn swb_batch_last_run_ms(e: *mut swb_engine, b: *mut swb_batch, ms: *mut f32, launches: *mut i64) -> c_int;
    pub fn swb_batch_run_n(e: *mut swb_engine, b: *mut swb_batch, steps: c_int, total_ms: *mut f32, launches: *mut i64) -> c_int;
    pub fn swb_dpx_peak(device: c_int, which: c_int, thread_instr_per_s: *mut f64) -> c_int;
}
