"""EMBOSS `srspair` report writer: prints an alignment in the format of the reference's golden files
(`test/emboss_results/*.dat`, layout visible in res0.dat:17-45), so that the engine's output can be diffed against them
directly instead of through the string extraction of test/check_results.pl:523-528.

Only formatting happens here; the alignment itself comes from the engine (`SequenceAlignment.align_raw` / `align_list`).
The rules below are pinned against all 300 golden reports by tests/test_report_cpu.py:
  * markup line: `|` identical letters, `:` substitution score > 0, `.` any other residue pair, blank for a gap column;
  * Identity = `|` columns, Similarity = `|` + `:` columns, Gaps = columns with a `-` in either row;
  * residue numbers: a block that holds residues shows the number of its first and of its last residue, a block that
    holds only gaps repeats the number of residues printed so far (0 before the first one).
"""
import io

WIDTH = 50


def _markup(a, b, score_of):
    out = []
    for x, y in zip(a, b):
        if x == "-" or y == "-":
            out.append(" ")
        elif x == y:
            out.append("|")
        elif score_of(x, y) > 0:
            out.append(":")
        else:
            out.append(".")
    return "".join(out)


def srspair_body(aln_a, aln_b, score, matrix, gap_open, gap_extend, start_a=1, start_b=1, name_a="s1", name_b="s2",
                 matrix_name="EBLOSUM62"):
    """The report from the first `#====` rule to the closing `#----` rules.  `start_a` / `start_b` are the 1-based
    numbers of the first residue of each row (1 for needle-style outputs, the alignment's start for water-style
    ones); `matrix` is the SubstitutionMatrix the alignment was scored with (used for the `:` / `.` markup)."""
    if len(aln_a) != len(aln_b):
        raise ValueError("aligned strings differ in length")
    s2i = matrix.string_to_index
    x = s2i.get("X")

    def score_of(p, q):
        i, j = s2i.get(p, x), s2i.get(q, x)
        return matrix.scores[i, j] if i is not None and j is not None else 0.0

    mk = _markup(aln_a, aln_b, score_of)
    n = len(aln_a)
    ident = mk.count("|")
    simil = ident + mk.count(":")
    gaps = sum(1 for p, q in zip(aln_a, aln_b) if p == "-" or q == "-")
    pct = (lambda k: 100.0 * k / n) if n else (lambda k: 0.0)
    f = io.StringIO()
    f.write("#=======================================\n#\n# Aligned_sequences: 2\n")
    f.write("# 1: %s\n# 2: %s\n# Matrix: %s\n" % (name_a, name_b, matrix_name))
    f.write("# Gap_penalty: %.1f\n# Extend_penalty: %.1f\n#\n" % (abs(gap_open), abs(gap_extend)))
    f.write("# Length: %d\n" % n)
    f.write("# Identity:   %5d/%d (%4.1f%%)\n" % (ident, n, pct(ident)))
    f.write("# Similarity: %5d/%d (%4.1f%%)\n" % (simil, n, pct(simil)))
    f.write("# Gaps:       %5d/%d (%4.1f%%)\n" % (gaps, n, pct(gaps)))
    f.write("# Score: %.1f\n# \n#\n#=======================================\n\n" % score)
    done_a, done_b = start_a - 1, start_b - 1                  # residues printed so far (as sequence positions)
    for lo in range(0, n, WIDTH):
        ra, rb, rm = aln_a[lo:lo + WIDTH], aln_b[lo:lo + WIDTH], mk[lo:lo + WIDTH]
        na, nb = len(ra) - ra.count("-"), len(rb) - rb.count("-")
        f.write("%-13s %6d %s %6d\n" % (name_a[:13], done_a + 1 if na else done_a, ra, done_a + na))
        f.write("%-13s %6s %s\n" % ("", "", rm))
        f.write("%-13s %6d %s %6d\n\n" % (name_b[:13], done_b + 1 if nb else done_b, rb, done_b + nb))
        done_a += na
        done_b += nb
    f.write("\n#---------------------------------------\n#---------------------------------------\n")
    return f.getvalue()


def srspair_report(program, file_a, file_b, outfile, **kw):
    """Body preceded by the `########` program header of the golden files (without the `# Rundate:` line's date)."""
    go, ge = abs(kw["gap_open"]), abs(kw["gap_extend"])
    head = ("########################################\n# Program: %s\n# Commandline: %s\n#    -asequence %s\n"
            "#    -bsequence %s\n#    -gapopen %.1f\n#    -gapextend %.1f\n#    -outfile %s\n#    -datafile %s\n"
            "#    -sprotein1\n#    -sprotein2\n# Align_format: srspair\n# Report_file: %s\n"
            "########################################\n\n") % (program, program, file_a, file_b, go, ge, outfile,
                                                                 kw.get("matrix_name", "EBLOSUM62"), outfile)
    return head + srspair_body(**kw)


def starts_of(idx_a, idx_b):
    """1-based number of the first aligned residue of each sequence from the engine's index lists (-1 = gap)."""
    sa = next((int(v) + 1 for v in idx_a if v >= 0), 1)
    sb = next((int(v) + 1 for v in idx_b if v >= 0), 1)
    return sa, sb
