#!/usr/bin/env python3
"""Generates tests/golden/emboss_reports.json.gz from /root/reference/test/emboss_results/*.dat: for each of the 300
golden reports the text from the first `#====` rule to the end of the file (the part an srspair writer must reproduce;
the `########` program header above it carries the run date) plus the residue numbers of the first block.  Run in the
build container only (the reference tree does not travel to the GPU box); the output is committed."""
import gzip
import json
import os
import re

REF = os.environ.get("SWB_REFERENCE", "/root/reference")
OUT = os.path.dirname(os.path.abspath(__file__))


def main():
    recs = []
    for key, pre in (("local", "res"), ("glocal", "needle_glocal_res"), ("global", "needle_res")):
        for k in range(100):
            text = open(os.path.join(REF, "test/emboss_results/%s%d.dat" % (pre, k))).read()
            body = text[text.index("#======================================="):]
            m1 = re.search(r"^s1\s+([0-9]+) ", body, re.M)
            m2 = re.search(r"^s2\s+([0-9]+) ", body, re.M)
            recs.append({"id": k, "kind": key, "body": body, "first1": int(m1.group(1)), "first2": int(m2.group(1))})
    with gzip.GzipFile(os.path.join(OUT, "emboss_reports.json.gz"), "w", mtime=0) as f:
        f.write(json.dumps(recs).encode())
    print("emboss_reports.json.gz: %d reports" % len(recs))


if __name__ == "__main__":
    main()
