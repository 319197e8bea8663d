"""include/mpb200.h must declare exactly the extern "C" entry points the CUDA sources define."""
import glob
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_declares_every_entry_point_and_nothing_else():
    defined = set()
    for path in glob.glob(os.path.join(ROOT, "mpi_pytorch_b200", "csrc", "*.cu")):
        src = open(path).read()
        defined |= set(re.findall(r"MPB_API\s+[A-Za-z_ ]*?\**\s*(mpb_[a-z0-9_]+)\s*\(", src))
    header = open(os.path.join(ROOT, "include", "mpb200.h")).read()
    declared = set(re.findall(r"\b(mpb_[a-z0-9_]+)\s*\(", header))
    assert defined, "no entry points found"
    assert defined - declared == set(), "missing from the header: {}".format(sorted(defined - declared))
    assert declared - defined == set(), "declared but not defined: {}".format(sorted(declared - defined))
