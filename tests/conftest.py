import gzip
import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def pytest_collection_modifyitems(config, items):
    # GPU tests are skipped (not failed) when collected on a box without a device and -m gpu was not asked for.
    try:
        import torch
        has = torch.cuda.is_available()
    except Exception:
        has = False
    if has:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def rust_kats():
    with open(os.path.join(GOLDEN, "rust_kats.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def emboss():
    with gzip.open(os.path.join(GOLDEN, "emboss.json.gz")) as f:
        return json.loads(f.read().decode())


def trim_terminal(s1, s2):
    """Strip unaligned overhang columns at both ends (test/check_results.pl:486-508 of the reference)."""
    n = len(s1)
    lo = 0
    while lo < n and (s1[lo] == "-" or s2[lo] == "-"):
        lo += 1
    hi = n
    while hi > lo and (s1[hi - 1] == "-" or s2[hi - 1] == "-"):
        hi -= 1
    return s1[lo:hi], s2[lo:hi]
