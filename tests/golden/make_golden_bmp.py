#!/usr/bin/env python3
"""Generates tests/golden/bmp_7x5.bmp by running the REFERENCE's own Write_To_File (engine/tools/texture.cpp:66-115, built
into oracle/_ref/libatray_ref.so by oracle/ref/build_ref.sh) on the deterministic 7x5 BGRX bitmap of tests/helpers.py:
bmp_test_bitmap(). Re-run only in the container that has /root/reference:  python tests/golden/make_golden_bmp.py"""
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import bind as B  # noqa: E402
from helpers import bmp_test_bitmap  # noqa: E402

with tempfile.TemporaryDirectory() as d:
    base = os.path.join(d, "resultings")
    assert B.write_bmp(B.ref(), bmp_test_bitmap(), base) == 1
    data = open(base + "_0.bmp", "rb").read()
open(os.path.join(ROOT, "tests", "golden", "bmp_7x5.bmp"), "wb").write(data)
print("wrote %d bytes" % len(data))
