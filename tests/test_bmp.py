"""Write_To_File (engine/tools/texture.cpp:66-115, 147-158): the product's BMP writer (C ABI: atray_write_bmp, host only)
and the oracle's restatement against the file the reference's own writer produced (tests/golden/bmp_7x5.bmp, generator
tests/golden/make_golden_bmp.py), and against the reference build itself where it exists."""
import os
import struct

import numpy as np

from helpers import GOLDEN, bmp_test_bitmap


def _golden():
    return open(os.path.join(GOLDEN, "bmp_7x5.bmp"), "rb").read()


def test_golden_file_is_the_documented_layout():
    g = _golden()
    assert len(g) == 14 + 56 + 7 * 5 * 4 and g[:2] == b"BM"
    total, r1, r2, off = struct.unpack_from("<ihhI", g, 2)
    assert (total, r1, r2, off) == (len(g), 0, 0, 70)
    dib = struct.unpack_from("<IIIhhIIIIIIIIII", g, 14)
    assert dib == (56, 7, 5, 1, 32, 3, 140, 197, 39, 0, 0, 0x00FF0000, 0x0000FF00, 0x000000FF, 0)
    assert g[70:] == bmp_test_bitmap().astype("<u4").tobytes()


def test_oracle_writer_equals_golden(port, tmp_path):
    from oracle import bind as B
    base = str(tmp_path / "resultings")
    assert B.write_bmp(port, bmp_test_bitmap(), base) == 1
    assert open(base + "_0.bmp", "rb").read() == _golden()


def test_product_writer_equals_golden_and_never_overwrites(built, tmp_path):
    import atray_b200.api as api
    base = str(tmp_path / "resultings")
    bm = bmp_test_bitmap()
    names = [api.Write_To_File(bm, base) for _ in range(3)]   # CREATE_NEW: _0, _1, _2
    assert names == [base + "_%d.bmp" % i for i in range(3)]
    for n in names:
        assert open(n, "rb").read() == _golden()
    big = np.arange(64 * 33, dtype=np.uint32).reshape(33, 64) & 0x00FFFFFF
    n = api.Write_To_File(big, str(tmp_path / "big"))
    data = open(n, "rb").read()
    assert len(data) == 70 + big.size * 4 and data[70:] == big.astype("<u4").tobytes()
    assert struct.unpack_from("<II", data, 18) == (64, 33)


def test_writers_equal_reference_build(ref, port, built, tmp_path):
    from oracle import bind as B
    import atray_b200.api as api
    bm = (np.random.default_rng(5).integers(0, 1 << 24, (19, 23))).astype(np.uint32)
    assert B.write_bmp(ref, bm, str(tmp_path / "r")) == 1
    assert B.write_bmp(port, bm, str(tmp_path / "o")) == 1
    n = api.Write_To_File(bm, str(tmp_path / "p"))
    r = open(str(tmp_path / "r_0.bmp"), "rb").read()
    assert r == open(str(tmp_path / "o_0.bmp"), "rb").read() == open(n, "rb").read()
