#!/usr/bin/env python3
"""Writes the reference's test data layout (inputs/seqN.{1,2}.fas, file_list.txt, emboss_results/*.dat) from the committed
fixtures tests/golden/emboss.json.gz and emboss_reports.json.gz into the directory given (default: this one)."""
import gzip
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(os.path.dirname(HERE), "tests", "golden")


def main(out):
    pairs = json.load(gzip.open(os.path.join(GOLDEN, "emboss.json.gz")))
    reports = json.load(gzip.open(os.path.join(GOLDEN, "emboss_reports.json.gz")))
    os.makedirs(os.path.join(out, "inputs"), exist_ok=True)
    os.makedirs(os.path.join(out, "emboss_results"), exist_ok=True)
    lines = []
    for p in pairs:
        for tag, key in ((1, "seq1"), (2, "seq2")):
            with open(os.path.join(out, "inputs", "seq%d.%d.fas" % (p["id"], tag)), "w") as f:
                f.write(">s%d\n%s\n" % (tag, p[key]))
        lines.append("inputs/seq%d.1.fas\tinputs/seq%d.2.fas" % (p["id"], p["id"]))
    with open(os.path.join(out, "file_list.txt"), "w") as f:
        f.write("\n".join(lines) + "\n")
    prefix = {"local": "res", "glocal": "needle_glocal_res", "global": "needle_res"}
    for r in reports:
        with open(os.path.join(out, "emboss_results", "%s%d.dat" % (prefix[r["kind"]], r["id"])), "w") as f:
            f.write(r["body"])
    print("materialized %d pairs, %d reports in %s" % (len(pairs), len(reports), out))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else HERE)
