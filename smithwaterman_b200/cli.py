"""Drop-in command line for the reference's two CLIs over the CUDA engine.

  python -m smithwaterman_b200.cli [-global|-glocal|-local] <fasta1> <fasta2>            seqalign   (SA-main.rs:10-46)
  python -m smithwaterman_b200.cli [-mode] -list <list file> [-out F]                     sa_opencl  (OCL-main.rs:267-316)
  python -m smithwaterman_b200.cli -cluster [-mode] [-identity x] [-coverage_short x] [-coverage_long x] -out F <fasta>
                                                                                          sa_opencl  (OCL-main.rs:149-265)

Same flags, same record format (`#score:` / `#type:` / `>name` / aligned string, SA-main.rs:40-43), BLOSUM62 with
open 10 / extend 0.5 as both reference binaries hard-code (SA-main.rs:34, OCL-main.rs:147).  Where the reference aligns
one pair per call, all pairs of a run go to the device in ONE batch (`swb_align_batch`); the clustering loop sends the
candidates of one representative as one batch -- the decisions of a representative's inner loop do not depend on each
other (OCL-main.rs:187-243), so the result is the same as the serial loop's.
"""
import sys

from .engine import (ALIGN_GLOBAL, ALIGN_GLOCAL, ALIGN_LOCAL, SeqData, SequenceAlignment, SubstitutionMatrix, TYPE_NAMES,
                     cli_record, rust_f32_display)

USAGE = """usage: sa_opencl [(-global|-glocal|-local(default))] <infile1 (fasta file)>  <infile2 (fasta file)>
usage: sa_opencl [(-global|-glocal|-local(default))] [-list] <list file>
usage: sa_opencl -cluster[ing] [(-global|-glocal|-local(default))] [-identity 0.0-1.0] [-coverage_short 0.0-1.0] [-coverage_long 0.0-1.0] -out <output file> <fasta file>
"""


class Options:
    """AlignmentOptions::parse, OCL-main.rs:33-116 (errors are exceptions instead of panics)."""

    def __init__(self, args):
        self.alignment_type, self.list, self.clustering = ALIGN_LOCAL, False, False
        self.c_identity = self.c_coverage_short = self.c_coverage_long = None
        self.outfilename, self.file1, self.file2 = "", "", ""
        files, skip = [], set()
        for k, a in enumerate(args):
            if k in skip:
                continue
            if a in ("-glocal", "-global", "-local"):
                self.alignment_type = {"-glocal": ALIGN_GLOCAL, "-global": ALIGN_GLOBAL, "-local": ALIGN_LOCAL}[a]
            elif a == "-list":
                self.list = True
            elif a in ("-cluster", "-clustering"):
                self.clustering = True
            elif a in ("-coverage_short", "-coverage_long", "-identity", "-out"):
                if k + 1 >= len(args):
                    raise ValueError("option %s needs a value" % a)
                v = args[k + 1]
                skip.add(k + 1)
                if a == "-out":
                    self.outfilename = v
                else:
                    setattr(self, {"-coverage_short": "c_coverage_short", "-coverage_long": "c_coverage_long",
                                   "-identity": "c_identity"}[a], float(v))
            elif a.startswith("-"):
                raise ValueError("Unknown option %s" % a)
            else:
                files.append(a)
        if not self.clustering and not self.list:
            if len(files) != 2:
                raise ValueError("2 files must be provided %r." % files)
            self.file1, self.file2 = files
        else:
            if self.clustering and self.list:
                raise ValueError("Incompatible option -list & -cluster(ing)")
            if len(files) != 1:
                raise ValueError("1 file must be provided %r." % files)
            self.file1 = files[0]
            if self.clustering and not self.outfilename:
                raise ValueError("Clustering must have -out.")


def read_list(path, out=sys.stdout):
    """The list-file reader of OCL-main.rs:268-288: tab (else blank) separated pairs of FASTA paths."""
    pairs = []
    with open(path) as f:
        for line in f.read().splitlines():
            spp = line.split("\t")
            if len(spp) == 1:
                spp = line.split(" ")
            if len(spp) > 2:
                out.write("%s \n^ Only %s %s are used.\n" % (line, spp[0], spp[1]))    # (and, like the reference, not queued)
            elif len(spp) < 2:
                out.write("%s \n is ignoed.\n" % line)
            else:
                pairs.append((spp[0], spp[1]))
    return pairs


def align_files(engine, filelist, out=sys.stdout, outfile=None):
    """All-vs-all of every file pair (SA-main.rs:35-45 / OCL-main.rs:295-316), ONE device batch for the whole run."""
    names, sa, sb = [], [], []
    for f1, f2 in filelist:
        s1, s2 = SeqData.load_fasta(f1, False), SeqData.load_fasta(f2, False)
        for x in s1:
            for y in s2:
                names.append((x.name, y.name)); sa.append(x); sb.append(y)
    res = engine.align_list(sa, sb, True)
    mess = TYPE_NAMES[engine.alignment_type]
    for (n1, n2), r in zip(names, res):
        if outfile is not None:                               # OCL-main.rs:303-307 (no newline after #score / #type there)
            outfile.write("#score:%s#type:%s>%s\n%s\n>%s\n%s\n" % (rust_f32_display(r["score"]), mess, n1, r["aln_a"], n2, r["aln_b"]))
        else:
            out.write(cli_record(n1, n2, r["aln_a"], r["aln_b"], r["score"], engine.alignment_type))
    return len(res)


def cluster(seqs, align_many, identity=0.8, coverage_short=0.8, coverage_long=0.8, out=sys.stdout, stats_many=None):
    """Greedy length-sorted clustering of OCL-main.rs:149-243.

    seqs: list of SeqData; align_many(rep, candidates) -> list of (aln_a, aln_b) strings of align(rep, cand, false).
    stats_many(rep, candidates) -> list of (a_letters, b_letters, matches, aln_len), the four numbers the loop reads off
    those strings (OCL-main.rs:205-233); when given it is used instead of align_many (the device counts them during
    the traceback walk, `swb_align_stats_batch`) and receives the sequences as bytes, converted once.
    Returns (sorted sequences, cluster_of, members) exactly as the reference's loop leaves them.
    """
    seq1 = sorted(seqs, key=lambda s: -len(s.seq))                      # stable, longest first (OCL-main.rs:151)
    n = len(seq1)
    strs = ["".join(s.seq) for s in seq1]
    blen = [len(s.encode("utf-8")) for s in strs]                         # String::len() is a byte length
    raw = [s.as_bytes() for s in seq1] if stats_many is not None else None
    cluster_of, identical = list(range(n)), list(range(n))
    members = [[] for _ in range(n)]
    identical_members = [[] for _ in range(n)]
    for ii in range(n):                                                 # OCL-main.rs:160-180: exact duplicates
        if identical[ii] != ii:
            continue
        identical_members[ii].append(ii)
        for jj in range(ii + 1, n):
            if identical[jj] != jj:
                continue
            if blen[ii] != blen[jj]:
                break
            if strs[ii] == strs[jj]:
                identical[jj] = ii
                identical_members[ii].append(jj)
    processed = 0
    for ii in range(n):                                                 # OCL-main.rs:181-243
        if cluster_of[ii] != ii or identical[ii] != ii:
            continue
        members[ii].extend(identical_members[ii]); identical_members[ii] = []
        cand = [jj for jj in range(ii + 1, n) if cluster_of[jj] == jj and identical[jj] == jj]
        need = [jj for jj in cand if strs[jj] not in strs[ii]]
        if not need:
            alns = {}
        elif stats_many is not None:
            alns = dict(zip(need, stats_many(raw[ii], [raw[jj] for jj in need])))
        else:
            alns = {}
            for jj, (ra, rb) in zip(need, align_many(seq1[ii], [seq1[jj] for jj in need])):
                alns[jj] = (sum(1 for c in ra if c != "-"), sum(1 for c in rb if c != "-"),
                            sum(1 for x, y in zip(ra, rb) if x != "-" and y != "-" and x == y), len(ra))
        for jj in range(ii + 1, n):
            if cluster_of[jj] == jj and identical[jj] == jj:
                if jj not in alns:                                      # si.contains(sj)
                    if blen[ii] and blen[jj] / blen[ii] >= coverage_long:
                        cluster_of[jj] = ii
                        members[ii].extend(identical_members[jj]); identical_members[jj] = []
                else:
                    alen, blen_, matchnum, ncol = alns[jj]
                    lcov = alen / len(seq1[ii].seq) if len(seq1[ii].seq) else float("nan")
                    scov = blen_ / len(seq1[jj].seq) if len(seq1[jj].seq) else float("nan")
                    ident = matchnum / ncol if ncol else float("nan")
                    if lcov >= coverage_long and scov >= coverage_short and ident >= identity:
                        cluster_of[jj] = ii
                        members[ii].extend(identical_members[jj]); identical_members[jj] = []
                if (jj + 1) % 1000 == 0:
                    out.write("%d alignments were done.\n" % (jj + 1))
        processed = ii + 1
        if processed % 10 == 0:
            out.write("%d sequences were processed.\n" % processed)
    return seq1, cluster_of, members


def cluster_native(engine, seqs, identity=0.8, coverage_short=0.8, coverage_long=0.8, out=sys.stdout):
    """The same loop run by the library (`swb_cluster`: sequences resident on the device, one batch per
    representative).  Returns what cluster() returns; the member lists are rebuilt in the order the reference appends
    them (a representative and its exact copies first, then every sequence that joined, each followed by its copies)."""
    order, cl, ident, _ = engine.cluster(seqs, identity, coverage_short, coverage_long)
    n = len(order)
    seq1 = [seqs[k] for k in order.tolist()]
    cluster_of, identical = cl.tolist(), ident.tolist()
    copies = [[] for _ in range(n)]
    for p in range(n):
        copies[identical[p]].append(p)                                   # the first copy heads its own list
    members = [[] for _ in range(n)]
    for p in range(n):
        if identical[p] == p:
            members[cluster_of[p]].extend(copies[p])
    for p in range(n):                                                   # progress lines of OCL-main.rs:240-242
        if cluster_of[p] == p and identical[p] == p and (p + 1) % 10 == 0:
            out.write("%d sequences were processed.\n" % (p + 1))
    return seq1, cluster_of, members


def write_clusters(outfilename, seq1, cluster_of, members):
    """Representative FASTA + .clstr member lists, OCL-main.rs:244-265."""
    with open(outfilename, "w") as f:
        for cc, s in enumerate(seq1):
            if cluster_of[cc] == cc:
                f.write(">%s %s\n%s\n" % (s.name, s.desc, "".join(s.seq)))
    with open(outfilename + ".clstr", "w") as f:
        for cc in range(len(seq1)):
            if members[cc]:
                f.write(" ".join(seq1[m].name for m in members[cc]) + "\n")


def main(argv=None, out=sys.stdout):
    args = list(sys.argv[1:] if argv is None else argv)
    if len(args) < 2:
        sys.stderr.write(USAGE)
        return -1
    opp = Options(args)
    engine = SequenceAlignment(SubstitutionMatrix.get_blosum62_matrix(), 10.0, 0.5, opp.alignment_type)
    try:
        if opp.clustering:
            seqs = SeqData.load_fasta(opp.file1, False)

            seq1, cluster_of, members = cluster_native(engine, seqs,
                                                       opp.c_identity if opp.c_identity is not None else 0.8,
                                                       opp.c_coverage_short if opp.c_coverage_short is not None else 0.8,
                                                       opp.c_coverage_long if opp.c_coverage_long is not None else 0.8, out)
            write_clusters(opp.outfilename, seq1, cluster_of, members)
        else:
            filelist = read_list(opp.file1, out) if opp.list else [(opp.file1, opp.file2)]
            if opp.outfilename:
                with open(opp.outfilename, "w") as f:
                    align_files(engine, filelist, out, f)
            else:
                align_files(engine, filelist, out)
    finally:
        engine.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
