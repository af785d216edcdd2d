//! The API of the reference's `sequence_alignment` crate (rust/sequence_alignment/src/sequence_alignment.rs) over the
//! swb200 CUDA engine.  Same type and method names, same argument meaning, same panics where the reference panics:
//!
//! | reference                                              | here                                   | C ABI call |
//! |--------------------------------------------------------|----------------------------------------|------------|
//! | `SequenceAlignment::new(Box<dyn ScoringMatrix>, go, ge, type)` SA.rs:26 | `SequenceAlignment::new` | `swb_engine_create` |
//! | `align(&mut self, &SeqData, &SeqData, retain_all)` SA.rs:389            | `align`                  | `swb_align` / `swb_align_pssm` |
//! | `align_partial(.., partial_region, score_only)` SA.rs:393               | `align_partial`          | `swb_align_partial` |
//! | the serial loops of the CLIs, SA-main.rs:35-45, OCL-main.rs:295-316     | `align_list`, `score_list` | `swb_align_batch`, `swb_score_batch` |
//! | the `-cluster` loop, OCL-main.rs:149-243                                | `cluster`                | `swb_cluster` |
//! | one query against many records (the same loops with a fixed `ss1`)      | `Database`, `search`     | `swb_db_upload`, `swb_search`, `swb_search_align` |
//!
//! `OpenCLSequenceAlignment::new(max_length, ..)` (OCL.rs:103) maps to the same type; `max_length` has no meaning any
//! more (buffers are sized per call, pairs beyond the direction-byte budget are traced through checkpointed recompute).
//!
//! NOT COMPILED IN THIS REPOSITORY'S BUILD IMAGE (no cargo / rustc); the C ABI it calls is exercised by the Python and
//! C++ mirrors (`smithwaterman_b200/engine.py`, `include/swb200.hpp`) in the test-suite.
use std::collections::HashMap;
use std::ffi::CStr;
use std::fs::File;
use std::io::{BufRead, BufReader};
use std::os::raw::{c_char, c_int};

use swb200_sys as sys;

pub const ALIGN_GLOBAL: usize = 0; // SA.rs:11-13
pub const ALIGN_GLOCAL: usize = 1;
pub const ALIGN_LOCAL: usize = 2;

// ---------------------------------------------------------------------------------------------------------------------
// SeqData (SA.rs:799-887): one String per residue, FASTA reader with the reference's header rules.
#[derive(Clone, Debug, Default)]
pub struct SeqData {
    pub name: String,
    pub desc: String,
    pub seq: Vec<String>,
}

impl SeqData {
    pub fn new() -> SeqData {
        SeqData::default()
    }
    pub fn create(n: String, d: String, s: String) -> SeqData {
        SeqData { name: n, desc: d, seq: SeqData::line_to_seq(s, true) }
    }
    /// SA.rs:811-821: `retainws` drops only CR / LF, otherwise every whitespace character goes.
    pub fn line_to_seq(s: String, retainws: bool) -> Vec<String> {
        s.chars()
            .filter(|c| if retainws { *c != '\r' && *c != '\n' } else { !c.is_whitespace() })
            .map(|c| c.to_string())
            .collect()
    }
    /// SA.rs:830-887: any line that contains `>` starts a record (a warning if it is not the first character); the name
    /// is the first token, the rest is the description; a record with neither name nor residues is dropped.
    pub fn load_fasta(filename: &str, retainws: bool) -> Vec<SeqData> {
        let reader = BufReader::new(File::open(filename).unwrap());
        let mut out: Vec<SeqData> = Vec::new();
        let mut cur: Option<SeqData> = None;
        let mut flush = |cur: &mut Option<SeqData>, out: &mut Vec<SeqData>| {
            if let Some(r) = cur.take() {
                if !r.name.is_empty() || !r.seq.is_empty() {
                    out.push(r);
                }
            }
        };
        for line in reader.lines() {
            let line = line.unwrap();
            match line.find('>') {
                Some(pos) => {
                    flush(&mut cur, &mut out);
                    if pos > 0 {
                        eprintln!("> was found at {}. This line was used as header anyway.", pos);
                    }
                    let t = line.trim();
                    let body = t.strip_prefix('>').unwrap_or(t);
                    // leading blanks are skipped, the name runs to the next blank, that blank is dropped and the rest
                    // of the line is the description verbatim (SA.rs:852-871)
                    let body = body.trim_start();
                    let (name, desc) = match body.char_indices().find(|(_, c)| c.is_whitespace()) {
                        Some((k, c)) => (&body[..k], &body[k + c.len_utf8()..]),
                        None => (body, ""),
                    };
                    cur = Some(SeqData { name: name.to_string(), desc: desc.to_string(), seq: Vec::new() });
                }
                None => {
                    let r = cur.get_or_insert_with(SeqData::default);
                    r.seq.extend(SeqData::line_to_seq(line, retainws));
                }
            }
        }
        flush(&mut cur, &mut out);
        out
    }
    fn bytes(&self) -> Vec<u8> {
        self.seq.iter().map(|s| s.as_bytes().first().copied().unwrap_or(b'X')).collect()
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// ScoringMatrix (SA.rs:574-580) and its two implementations.
pub trait ScoringMatrix {
    fn get_score(&self, a: usize, b: usize) -> f32;
    fn seq_to_index(&self, ss: &SeqData, partial_region: Option<usize>) -> Vec<usize>;
    fn set_score(&mut self, a: usize, b: usize, s: f32);
    fn prepare(&mut self, s1: &SeqData, s2: &SeqData);
    /// What the engine needs: the letters and the row-major score table, or `None` for a position-specific matrix.
    fn as_table(&self) -> Option<(Vec<u8>, Vec<f32>)>;
    fn position_scores(&self) -> Option<(&[f32], usize, usize)> {
        None
    }
}

/// SA.rs:626-795.
pub struct SubstitutionMatrix {
    pub string_to_index: HashMap<String, usize>,
    pub index_to_string: Vec<String>,
    pub scores: Vec<Vec<f32>>,
}

impl SubstitutionMatrix {
    fn from_table(letters: &[u8], flat: &[f32]) -> SubstitutionMatrix {
        let n = letters.len();
        let index_to_string: Vec<String> = letters.iter().map(|c| (*c as char).to_string()).collect();
        let string_to_index = index_to_string.iter().cloned().enumerate().map(|(i, s)| (s, i)).collect();
        let scores = (0..n).map(|i| flat[i * n..(i + 1) * n].to_vec()).collect();
        SubstitutionMatrix { string_to_index, index_to_string, scores }
    }
    /// SA.rs:697-733 (24 letters, `A R N D C Q E G H I L K M F P S T W Y V B Z X *`).
    pub fn get_blosum62_matrix() -> SubstitutionMatrix {
        let mut letters = vec![0u8; 32];
        let mut flat = vec![0f32; 32 * 32];
        let n = unsafe { sys::swb_matrix_blosum62(letters.as_mut_ptr() as *mut c_char, flat.as_mut_ptr()) } as usize;
        SubstitutionMatrix::from_table(&letters[..n], &flat[..n * n])
    }
    /// SA.rs:681-695 (letters A..Z).
    pub fn get_mat_matrix(match_score: f32, mismatch_score: f32) -> SubstitutionMatrix {
        let mut letters = vec![0u8; 32];
        let mut flat = vec![0f32; 32 * 32];
        let n = unsafe { sys::swb_matrix_match_mismatch(match_score, mismatch_score, letters.as_mut_ptr() as *mut c_char, flat.as_mut_ptr()) } as usize;
        SubstitutionMatrix::from_table(&letters[..n], &flat[..n * n])
    }
    /// SA.rs:735-794: NCBI-style text; `#` comments, a header line of letters, then one row per letter.
    pub fn load_from_lines(lines: Vec<String>) -> SubstitutionMatrix {
        let mut header: Option<Vec<String>> = None;
        let mut rows: Vec<(String, Vec<f32>)> = Vec::new();
        for l in lines {
            let t = l.trim();
            if t.is_empty() || t.starts_with('#') {
                continue;
            }
            let tok: Vec<&str> = t.split_whitespace().collect();
            match &header {
                None => {
                    let mut seen = HashMap::new();
                    for s in &tok {
                        if seen.insert(s.to_string(), ()).is_some() {
                            panic!("{} was already found.", s);
                        }
                    }
                    header = Some(tok.iter().map(|s| s.to_string()).collect());
                }
                Some(h) => {
                    if tok.len() != h.len() + 1 {
                        eprintln!("{} was not parsed.", t);
                        continue;
                    }
                    rows.push((tok[0].to_string(), tok[1..].iter().map(|v| v.parse::<f32>().unwrap_or_else(|e| panic!("{:?}", e))).collect()));
                }
            }
        }
        let h = header.expect("no header line");
        let n = h.len();
        let string_to_index: HashMap<String, usize> = h.iter().cloned().enumerate().map(|(i, s)| (s, i)).collect();
        let mut scores = vec![vec![0f32; n]; n];
        for (name, vals) in rows {
            let i = *string_to_index.get(&name).unwrap_or_else(|| panic!("{} is not in the header.", name));
            scores[i] = vals;
        }
        SubstitutionMatrix { string_to_index, index_to_string: h, scores }
    }
    /// SA.rs:669-679: unknown letters map to `X`; a matrix without `X` panics.
    pub fn get_string_index(&self, c: &str) -> usize {
        if let Some(i) = self.string_to_index.get(c) {
            return *i;
        }
        match self.string_to_index.get("X") {
            Some(i) => *i,
            None => panic!("unknown letter {}. please set X to allow scoring for undefined letter pair.", c),
        }
    }
}

impl ScoringMatrix for SubstitutionMatrix {
    fn get_score(&self, a: usize, b: usize) -> f32 {
        self.scores[a][b]
    }
    fn seq_to_index(&self, ss: &SeqData, partial_region: Option<usize>) -> Vec<usize> {
        let n = partial_region.unwrap_or(ss.seq.len());
        ss.seq[..n].iter().map(|s| self.get_string_index(s)).collect()
    }
    fn set_score(&mut self, a: usize, b: usize, s: f32) {
        self.scores[a][b] = s;
    }
    fn prepare(&mut self, _s1: &SeqData, _s2: &SeqData) {}
    fn as_table(&self) -> Option<(Vec<u8>, Vec<f32>)> {
        let letters = self.index_to_string.iter().map(|s| s.as_bytes()[0]).collect();
        Some((letters, self.scores.iter().flatten().cloned().collect()))
    }
}

/// SA.rs:583-623: scores[a + b * a_length], sized by `prepare`.
#[derive(Default)]
pub struct PositionSpecificMatrix {
    pub scores: Vec<f32>,
    pub a_length: usize,
    pub b_length: usize,
}

impl PositionSpecificMatrix {
    pub fn new() -> PositionSpecificMatrix {
        PositionSpecificMatrix::default()
    }
}

impl ScoringMatrix for PositionSpecificMatrix {
    fn get_score(&self, a: usize, b: usize) -> f32 {
        self.scores[a + b * self.a_length]
    }
    fn seq_to_index(&self, ss: &SeqData, partial_region: Option<usize>) -> Vec<usize> {
        (0..partial_region.unwrap_or(ss.seq.len())).collect()
    }
    fn set_score(&mut self, a: usize, b: usize, s: f32) {
        let k = a + b * self.a_length;
        self.scores[k] = s;
    }
    fn prepare(&mut self, s1: &SeqData, s2: &SeqData) {
        self.a_length = s1.seq.len();
        self.b_length = s2.seq.len();
        if self.scores.len() < self.a_length * self.b_length {
            self.scores.resize(self.a_length * self.b_length, 0.0);
        }
    }
    fn as_table(&self) -> Option<(Vec<u8>, Vec<f32>)> {
        None
    }
    fn position_scores(&self) -> Option<(&[f32], usize, usize)> {
        Some((&self.scores, self.a_length, self.b_length))
    }
}

// ---------------------------------------------------------------------------------------------------------------------
/// SA.rs:15-552 / OCL.rs:71-918.  `&mut self` everywhere as in the reference: one engine per thread.
pub struct SequenceAlignment {
    h: *mut sys::swb_engine,
    pub scoring_matrix: Box<dyn ScoringMatrix>,
    pub alignment_type: usize,
}

fn err_text(h: *const sys::swb_engine) -> String {
    unsafe { CStr::from_ptr(sys::swb_last_error(h)) }.to_string_lossy().into_owned()
}

fn take(out: &mut sys::swb_alignment) -> (Vec<String>, Vec<String>, f32) {
    let n = out.aln_len as usize;
    let col = |p: *const c_char| -> Vec<String> {
        if n == 0 {
            return Vec::new();
        }
        unsafe { std::slice::from_raw_parts(p as *const u8, n) }.iter().map(|c| (*c as char).to_string()).collect()
    };
    let r = (col(out.aln_a), col(out.aln_b), out.score);
    unsafe { sys::swb_alignment_free(out) };
    r
}

impl SequenceAlignment {
    /// SA.rs:26-45: penalties are sign-insensitive, `alignment_type` is 0 / 1 / 2.
    pub fn new(ss: Box<dyn ScoringMatrix>, go: f32, ge: f32, alignment_type: usize) -> SequenceAlignment {
        SequenceAlignment::new_on_device(ss, go, ge, alignment_type, 0)
    }
    pub fn new_on_device(ss: Box<dyn ScoringMatrix>, go: f32, ge: f32, alignment_type: usize, device: i32) -> SequenceAlignment {
        let (letters, flat) = ss.as_table().unwrap_or((vec![b'X'], vec![0.0]));
        let m = sys::swb_matrix { n_letters: letters.len() as i32, letters: letters.as_ptr() as *const c_char, scores: flat.as_ptr() };
        let mut h = std::ptr::null_mut();
        let rc = unsafe { sys::swb_engine_create(&m, go, ge, alignment_type as c_int, device, &mut h) };
        if rc != sys::SWB_OK {
            panic!("{}", err_text(std::ptr::null()));
        }
        SequenceAlignment { h, scoring_matrix: ss, alignment_type }
    }
    fn check(&self, rc: c_int) {
        if rc != sys::SWB_OK {
            panic!("{}", err_text(self.h)); // the reference panics on the same inputs (SA.rs:360-371, 676)
        }
    }
    pub fn prepare(&mut self, s1: &SeqData, s2: &SeqData) {
        self.scoring_matrix.prepare(s1, s2);
    }
    /// SA.rs:389-392.
    pub fn align(&mut self, s1: &SeqData, s2: &SeqData, retain_all: bool) -> (Vec<String>, Vec<String>, f32) {
        self.align_partial(s1, s2, retain_all, None, false)
    }
    /// SA.rs:393-551.
    pub fn align_partial(&mut self, s1: &SeqData, s2: &SeqData, retain_all: bool, partial_region: Option<(usize, usize)>, score_only: bool) -> (Vec<String>, Vec<String>, f32) {
        let (a, b) = (s1.bytes(), s2.bytes());
        let mut out: sys::swb_alignment = unsafe { std::mem::zeroed() };
        if let Some((ps, la, lb)) = self.scoring_matrix.position_scores() {
            assert!(la == a.len() && lb == b.len(), "PositionSpecificMatrix was prepared for other lengths");
            let rc = unsafe { sys::swb_align_pssm(self.h, ps.as_ptr(), a.as_ptr() as _, la as i64, b.as_ptr() as _, lb as i64, retain_all as c_int, &mut out) };
            self.check(rc);
            let r = take(&mut out);
            return if score_only { (vec![], vec![], r.2) } else { r };
        }
        let (pa, pb) = partial_region.unwrap_or((a.len(), b.len()));
        let rc = unsafe {
            sys::swb_align_partial(self.h, a.as_ptr() as _, a.len() as i64, b.as_ptr() as _, b.len() as i64, pa as i64, pb as i64, retain_all as c_int, score_only as c_int, &mut out)
        };
        self.check(rc);
        take(&mut out)
    }
    /// The serial loop of `sa_opencl -list` / `seqalign` (OCL-main.rs:295-316, SA-main.rs:35-45) as ONE device batch.
    pub fn align_list(&mut self, pairs: &[(&SeqData, &SeqData)], retain_all: bool) -> Vec<(Vec<String>, Vec<String>, f32)> {
        let (mut a, mut b, mut oa, mut ob) = (Vec::new(), Vec::new(), vec![0i64], vec![0i64]);
        for (s1, s2) in pairs {
            a.extend(s1.bytes());
            b.extend(s2.bytes());
            oa.push(a.len() as i64);
            ob.push(b.len() as i64);
        }
        let mut outs: Vec<sys::swb_alignment> = (0..pairs.len().max(1)).map(|_| unsafe { std::mem::zeroed() }).collect();
        let flags = if retain_all { sys::SWB_RETAIN_ALL } else { 0 };
        let rc = unsafe { sys::swb_align_batch(self.h, a.as_ptr() as _, oa.as_ptr(), b.as_ptr() as _, ob.as_ptr(), pairs.len() as i64, flags, outs.as_mut_ptr()) };
        self.check(rc);
        outs.iter_mut().take(pairs.len()).map(take).collect()
    }
    /// `align_partial(.., score_only = true)` of n pairs: (score, end cell in a, end cell in b).
    pub fn score_list(&mut self, pairs: &[(&SeqData, &SeqData)]) -> Vec<(f32, i64, i64)> {
        let (mut a, mut b, mut oa, mut ob) = (Vec::new(), Vec::new(), vec![0i64], vec![0i64]);
        for (s1, s2) in pairs {
            a.extend(s1.bytes());
            b.extend(s2.bytes());
            oa.push(a.len() as i64);
            ob.push(b.len() as i64);
        }
        let n = pairs.len();
        let (mut sc, mut ea, mut eb) = (vec![0f32; n], vec![0i64; n], vec![0i64; n]);
        let rc = unsafe { sys::swb_score_batch(self.h, a.as_ptr() as _, oa.as_ptr(), b.as_ptr() as _, ob.as_ptr(), n as i64, sc.as_mut_ptr(), ea.as_mut_ptr(), eb.as_mut_ptr()) };
        self.check(rc);
        (0..n).map(|k| (sc[k], ea[k], eb[k])).collect()
    }
    /// The clustering loop of `sa_opencl -cluster` (OCL-main.rs:149-243): (order, cluster_of, identical), indexed by
    /// sorted position as `swb_cluster` documents.
    pub fn cluster(&mut self, seqs: &[SeqData], identity: f64, coverage_short: f64, coverage_long: f64) -> (Vec<i64>, Vec<i64>, Vec<i64>) {
        let (mut all, mut off) = (Vec::new(), vec![0i64]);
        for s in seqs {
            all.extend(s.bytes());
            off.push(all.len() as i64);
        }
        let n = seqs.len();
        let (mut order, mut cl, mut ident) = (vec![0i64; n.max(1)], vec![0i64; n.max(1)], vec![0i64; n.max(1)]);
        let rc = unsafe {
            sys::swb_cluster(self.h, all.as_ptr() as _, off.as_ptr(), n as i64, identity, coverage_short, coverage_long, order.as_mut_ptr(), cl.as_mut_ptr(), ident.as_mut_ptr(), std::ptr::null_mut())
        };
        self.check(rc);
        order.truncate(n);
        cl.truncate(n);
        ident.truncate(n);
        (order, cl, ident)
    }
    /// Shard `rank` of `world` of a database, resident on this engine's GPU (`swb_db_upload_shard`).
    pub fn upload_database(&mut self, seqs: &[SeqData], rank: i32, world: i32) -> Database {
        let (mut all, mut off) = (Vec::new(), vec![0i64]);
        for s in seqs {
            all.extend(s.bytes());
            off.push(all.len() as i64);
        }
        let mut h = std::ptr::null_mut();
        let rc = unsafe { sys::swb_db_upload_shard(self.h, all.as_ptr() as _, off.as_ptr(), seqs.len() as i64, rank, world, &mut h) };
        self.check(rc);
        Database { h, e: self.h }
    }
    /// One query against a resident database: top-k hits by (score desc, index asc), selected on the device.
    pub fn search(&mut self, query: &SeqData, db: &mut Database, topk: usize) -> Vec<sys::swb_hit> {
        let q = query.bytes();
        let mut hits = vec![sys::swb_hit::default(); topk.max(1)];
        let mut n = 0i64;
        let rc = unsafe { sys::swb_search(self.h, q.as_ptr() as _, q.len() as i64, db.h, topk as c_int, hits.as_mut_ptr(), &mut n) };
        self.check(rc);
        hits.truncate(n as usize);
        hits
    }
    /// `align(query, subject, retain_all)` of the hits this shard owns; `None` for hits of other shards.
    pub fn align_hits(&mut self, query: &SeqData, db: &mut Database, hits: &[sys::swb_hit], retain_all: bool) -> Vec<Option<(Vec<String>, Vec<String>, f32)>> {
        let q = query.bytes();
        let mut outs: Vec<sys::swb_alignment> = (0..hits.len().max(1)).map(|_| unsafe { std::mem::zeroed() }).collect();
        let mut owned = vec![0i32; hits.len().max(1)];
        let flags = if retain_all { sys::SWB_RETAIN_ALL } else { 0 };
        let rc = unsafe { sys::swb_search_align(self.h, q.as_ptr() as _, q.len() as i64, db.h, hits.as_ptr(), hits.len() as i64, flags, outs.as_mut_ptr(), owned.as_mut_ptr()) };
        self.check(rc);
        (0..hits.len()).map(|k| if owned[k] != 0 { Some(take(&mut outs[k])) } else { None }).collect()
    }
}

impl Drop for SequenceAlignment {
    fn drop(&mut self) {
        unsafe { sys::swb_engine_destroy(self.h) }
    }
}

/// A database shard resident on the device.  Must be dropped before the engine that uploaded it.
pub struct Database {
    h: *mut sys::swb_db,
    e: *mut sys::swb_engine,
}

impl Drop for Database {
    fn drop(&mut self) {
        unsafe { sys::swb_db_free(self.e, self.h) }
    }
}

/// Global top-k of the per-shard lists (`swb_search_merge`): the one exchange of a sharded search.
pub fn merge_hits(parts: &[sys::swb_hit], topk: usize) -> Vec<sys::swb_hit> {
    let mut out = vec![sys::swb_hit::default(); topk.max(1)];
    let mut n = 0i64;
    let rc = unsafe { sys::swb_search_merge(parts.as_ptr(), parts.len() as i64, topk as c_int, out.as_mut_ptr(), &mut n) };
    assert!(rc == sys::SWB_OK);
    out.truncate(n as usize);
    out
}

/// `OpenCLSequenceAlignment` of rust/sa_opencl (OCL.rs:71-918): the same engine; `max_length` is accepted and ignored.
pub type OpenCLSequenceAlignment = SequenceAlignment;
pub fn opencl_new(_max_length: usize, ss: Box<dyn ScoringMatrix>, go: f32, ge: f32, alignment_type: usize) -> SequenceAlignment {
    SequenceAlignment::new(ss, go, ge, alignment_type)
}
