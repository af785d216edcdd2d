#!/usr/bin/env python3
"""Generate the committed golden fixtures from the (read-only) reference tree.

Run in the build container only (needs /root/reference; the GPU box has no such path):

    python tests/golden/make_golden.py [/root/reference]

Writes, next to this file:
  rust_kats.json      every known-answer alignment asserted by the reference's Rust #[test]s
                      (rust/sequence_alignment/src/sequence_alignment.rs:921-1221 and
                       rust/sa_opencl/src/opencl_sequence_alignment.rs:1267-1450): inputs, matrix,
                       penalties, mode, retain_all and the asserted score / padded strings.
  emboss.json.gz      the 100 test/inputs pairs with the EMBOSS water / needle results of
                      test/emboss_results (aligned s1/s2 strings and `# Score:`) for
                      local (res*), glocal (needle_glocal_res*) and global (needle_res*).
  fasta_test1.fas     copy of the FASTA edge-case DATA fixture rust/sequence_alignment/test/test1.fas
                      with the expectations of sw_fastaloadtest (SA.rs:1090-1130) in fasta_test1.json.
Only data (sequences, expected answers) is extracted; no reference code is copied.
"""
import gzip
import json
import os
import re
import sys

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))

MODES = {"ALIGN_GLOBAL": 0, "ALIGN_GLOCAL": 1, "ALIGN_LOCAL": 2}


def rust_kats(path, origin):
    src = open(path, encoding="utf-8").read()
    cases = []
    for m in re.finditer(r"#\[test\]\s*fn\s+(\w+)\s*\(\)\s*\{", src):
        name = m.group(1)
        nxt = src.find("#[test]", m.end())
        body = src[m.end(): nxt if nxt > 0 else len(src)]
        body = "\n".join(l for l in body.split("\n") if not l.lstrip().startswith("//"))   # skip commented-out code
        st = {"matrix": None, "matrix_scale": 1.0, "pssm": False}
        cur = None
        # walk statements in textual order
        tok = re.compile(
            r'let\s+(?:mut\s+)?seq1_\s*=\s*"([^"]*)"|let\s+(?:mut\s+)?seq2_\s*=\s*"([^"]*)"'
            r"|get_mat_matrix\(\s*([-0-9.]+)_f32\s*,\s*([-0-9.]+)_f32\s*\)"
            r"|(get_blosum62_matrix)\(\)|(PositionSpecificMatrix::new)\(\)"
            r"|\*vvv\s*\*=\s*([0-9.]+)"
            r"|SequenceAlignment::new\(\s*(?:\d+\s*,)?\s*Box::new\((\w+)\)\s*,\s*([-0-9.]+)\s*,\s*([-0-9.]+)\s*,\s*(ALIGN_\w+)\)"
            r"|sw\.align\(&seq1,&seq2,(true|false)\)"
            r'|assert_eq!\(res\.2,([-0-9.]+)\)|assert_eq!\(r1,"([^"]*)"\)|assert_eq!\(r2,"([^"]*)"\)')
        pending_matrix = None
        for t in tok.finditer(body):
            g = t.groups()
            if g[0] is not None:
                st["seq1"] = g[0]
            elif g[1] is not None:
                st["seq2"] = g[1]
            elif g[2] is not None:
                pending_matrix = {"kind": "match_mismatch", "match": float(g[2]), "mismatch": float(g[3])}
                st["matrix_scale"] = 1.0
            elif g[4] is not None:
                pending_matrix = {"kind": "blosum62"}
                st["matrix_scale"] = 1.0
            elif g[5] is not None:
                st["pssm_next"] = True
            elif g[6] is not None:
                st["matrix_scale"] = float(g[6])
            elif g[7] is not None:
                var = g[7]
                st["pssm"] = (var == "pm")
                st["matrix"] = dict(pending_matrix, scale=st["matrix_scale"])
                st["go"], st["ge"], st["mode"] = float(g[8]), float(g[9]), MODES[g[10]]
            elif g[11] is not None:
                cur = {"test": name, "origin": origin, "seq1": st["seq1"], "seq2": st["seq2"],
                       "matrix": st["matrix"], "go": st["go"], "ge": st["ge"], "mode": st["mode"],
                       "retain_all": g[11] == "true", "pssm": st["pssm"]}
                cases.append(cur)
            elif g[12] is not None and cur is not None:
                cur["score"] = float(g[12])
            elif g[13] is not None and cur is not None:
                cur["r1"] = g[13]
            elif g[14] is not None and cur is not None:
                cur["r2"] = g[14]
    return [c for c in cases if "score" in c or "r1" in c]


def read_fasta_simple(path):
    seq = []
    for line in open(path):
        if ">" in line:
            continue
        seq.append("".join(line.split()))
    return "".join(seq)


def emboss_parse(path):
    s1, s2, score = [], [], None
    for line in open(path):
        m = re.match(r"^# Score:\s*([-0-9.]+)", line)
        if m:
            score = float(m.group(1))
        m = re.match(r"^\s*s1\s+[0-9]+\s*(\S+)", line)      # same regex shape as check_results.pl:523-528
        if m:
            s1.append(m.group(1))
        m = re.match(r"^\s*s2\s+[0-9]+\s*(\S+)", line)
        if m:
            s2.append(m.group(1))
    return "".join(s1), "".join(s2), score


def main():
    kats = rust_kats(os.path.join(REF, "rust/sequence_alignment/src/sequence_alignment.rs"), "SA.rs")
    kats += rust_kats(os.path.join(REF, "rust/sa_opencl/src/opencl_sequence_alignment.rs"), "OCL.rs")
    with open(os.path.join(OUT, "rust_kats.json"), "w") as f:
        json.dump(kats, f, indent=0)
    print("rust_kats.json: %d cases" % len(kats))

    pairs = []
    for line in open(os.path.join(REF, "test/file_list.txt")):
        parts = line.rstrip("\n").split("\t") if "\t" in line else line.split()
        if len(parts) < 2:
            continue
        k = int(re.search(r"seq(\d+)\.1", parts[0]).group(1))
        rec = {"id": k, "file1": parts[0], "file2": parts[1],
               "seq1": read_fasta_simple(os.path.join(REF, "test", parts[0])),
               "seq2": read_fasta_simple(os.path.join(REF, "test", parts[1]))}
        for key, pre in (("local", "res"), ("glocal", "needle_glocal_res"), ("global", "needle_res")):
            a, b, sc = emboss_parse(os.path.join(REF, "test/emboss_results/%s%d.dat" % (pre, k)))
            rec[key] = {"s1": a, "s2": b, "score": sc}
        pairs.append(rec)
    with gzip.GzipFile(os.path.join(OUT, "emboss.json.gz"), "w", mtime=0) as f:
        f.write(json.dumps(pairs).encode())
    print("emboss.json.gz: %d pairs" % len(pairs))

    with open(os.path.join(REF, "rust/sequence_alignment/test/test1.fas"), "rb") as f:
        data = f.read()
    with open(os.path.join(OUT, "fasta_test1.fas"), "wb") as f:
        f.write(data)
    expect = [["seqA", "", "A" * 12], ["seqB", "b desu", "B" * 12], ["seqC", "c desu", "C" * 12],
              ["seqD", "d desu", ""], ["seqE", "e desu", "E" * 12], ["F", "", "F" * 12],
              ["G", "", "G" * 12], ["H", "h", "H" * 12], ["I", "", ""]]     # SA.rs:1090-1130
    with open(os.path.join(OUT, "fasta_test1.json"), "w") as f:
        json.dump(expect, f)


if __name__ == "__main__":
    main()
