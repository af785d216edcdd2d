"""CPU-side checks of the product: the C-ABI library builds/loads and exports every symbol that
include/swb200.h declares; host-side mirrors (matrix parser, FASTA, formatting) follow the reference.
No compute entry point is called here (no GPU)."""
import json
import os
import re

import numpy as np

from conftest import GOLDEN, ROOT
import smithwaterman_b200 as swb
from smithwaterman_b200 import _lib


def test_library_exports_every_declared_symbol():
    L = _lib.lib()
    header = open(os.path.join(ROOT, "include", "swb200.h")).read()
    declared = sorted(set(re.findall(r"\b(swb_[a-z0-9_]+)\s*\(", header)))
    assert len(declared) >= 17
    for name in declared:
        assert hasattr(L, name), name
    assert sorted(L._swb_symbols) == declared


def test_builtin_matrices_match_reference_tables():
    from oracle import oracle as O
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    ref = O.Matrix.blosum62()
    assert m.index_to_string == ref.letters
    assert np.array_equal(m.scores, ref.scores)
    mm = swb.SubstitutionMatrix.get_mat_matrix(5.0, -4.0)
    ref = O.Matrix.match_mismatch(5.0, -4.0)
    assert mm.index_to_string == ref.letters and np.array_equal(mm.scores, ref.scores)


def test_matrix_parser():                                   # SA.rs:897-919
    sm = swb.SubstitutionMatrix.load_from_lines(["  A T G C \n", "A 1 -3 -3 -3 ", "T -3  1 -3 -3 ",
                                                 "G -3  -3  1 -3 ", "C -3  -3  -3  1"])
    for x in "ATGC":
        assert sm.get_score_str(x, x) == 1.0
    for x, y in ("AC", "TG", "GA", "CT"):
        assert sm.get_score_str(x, y) == -3.0


def test_fasta_edge_cases():                                # SA.rs:1090-1130
    expect = json.load(open(os.path.join(GOLDEN, "fasta_test1.json")))
    fas = swb.SeqData.load_fasta(os.path.join(GOLDEN, "fasta_test1.fas"))
    assert [(f.name, f.desc, "".join(f.seq)) for f in fas] == [tuple(e) for e in expect]


def test_score_formatting():
    assert swb.rust_f32_display(485.0) == "485"
    assert swb.rust_f32_display(54.5) == "54.5"
    assert swb.rust_f32_display(-1770.5) == "-1770.5"
    assert swb.cli_record("s1", "s2", "AB", "A-", 4.0, swb.ALIGN_LOCAL) == "#score:4\n#type:Local alignment\n>s1\nAB\n\n>s2\nA-\n\n"


def test_no_oracle_in_product():
    """The product tree must not reference the oracle (no CPU fallback)."""
    pkg = os.path.join(ROOT, "smithwaterman_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".h", ".cpp")):
                txt = open(os.path.join(dp, f), errors="replace").read()
                assert "import oracle" not in txt and "from oracle" not in txt and "liboracle" not in txt, f


def test_rust_sys_crate_declares_every_symbol():
    """rust/swb200-sys (source only: no cargo in the image) declares exactly the entry points of include/swb200.h, and
    its #[repr(C)] structs have the fields of the C structs in the same order."""
    header = open(os.path.join(ROOT, "include", "swb200.h")).read()
    rs = open(os.path.join(ROOT, "rust", "swb200-sys", "src", "lib.rs")).read()
    declared = sorted(set(re.findall(r"\b(swb_[a-z0-9_]+)\s*\(", header)))
    bound = sorted(set(re.findall(r"pub fn (swb_[a-z0-9_]+)\s*\(", rs)))
    assert bound == declared
    for name in ("swb_alignment", "swb_stats", "swb_aln_stats", "swb_hit", "swb_matrix"):
        c_body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (name, name), header, re.S).group(1)
        c_body = re.sub(r"/\*.*?\*/", "", c_body, flags=re.S)
        c_fields = [f for decl in c_body.split(";") for f in re.findall(r"\*?\s*([a-z_0-9]+)\s*(?:,|$)", decl.split(None, 1)[1] if len(decl.split(None, 1)) > 1 else "")]
        c_fields = [f for f in c_fields if f not in ("const", "char", "float")]
        rs_body = re.search(r"pub struct %s \{(.*?)\n\}" % name, rs, re.S).group(1)
        rs_fields = re.findall(r"pub ([a-z_0-9]+):", rs_body)
        assert rs_fields == c_fields, (name, rs_fields, c_fields)
    wrapper = open(os.path.join(ROOT, "rust", "swb200", "src", "lib.rs")).read()
    for needed in ("pub fn new(", "pub fn align(", "pub fn align_partial(", "pub fn load_fasta(", "pub fn get_blosum62_matrix(", "pub fn get_mat_matrix(", "pub fn load_from_lines("):
        assert needed in wrapper
