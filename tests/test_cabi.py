"""CPU: the C-ABI library loads and exports every symbol include/atray_b200.h declares; the product never touches
the oracle; device entry points fail loudly (no CPU fallback) when no GPU is present."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "atray_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(atray_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported(built):
    lib = ctypes.CDLL(os.path.join(ROOT, "atray_b200", "libatray_b200.so"))
    names = declared_symbols()
    assert len(names) >= 35
    for n in names:
        assert hasattr(lib, n), "declared in include/atray_b200.h but not exported: " + n
    from atray_b200 import api
    assert sorted(api.EXPORTED_SYMBOLS) == names


def test_header_compiles_as_c(built, tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "atray_b200.h"\nint main(void){AtrayCamera c; AtrayFrame f; (void)c; (void)f; return sizeof(AtrayMaterial)==28?0:1;}\n')
    exe = tmp_path / "t"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    assert subprocess.call([str(exe)]) == 0


def test_kernels_are_sm100a(built):
    out = subprocess.run(["cuobjdump", "-lelf", os.path.join(ROOT, "atray_b200", "libatray_b200.so")],
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True).stdout
    assert "sm_100a" in out, out


def test_product_does_not_reference_the_oracle():
    for top in ("atray_b200", "include", "Source"):
        for dp, _, files in os.walk(os.path.join(ROOT, top)):
            for f in files:
                if f.endswith((".py", ".cpp", ".cu", ".h", ".hpp", ".cuh")) or f == "Makefile":
                    text = open(os.path.join(dp, f), errors="ignore").read()
                    assert "oracle" not in text.lower().replace("no cpu fallback", ""), os.path.join(dp, f)


def test_no_cpu_fallback_without_gpu(built):
    import atray_b200 as A
    if A.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(A.AtrayError):
        A.Context(0)
    with pytest.raises(A.AtrayError):
        A.MultiContext([0, 1])
    # the C++ host layer fails loudly too: prep_scene leaves scene.gpu null and app_headless exits non-zero
    exe = os.path.join(ROOT, "Source", "engine", "app_headless")
    if os.path.exists(exe):
        from tools import gen_standin_obj
        path = gen_standin_obj.ensure("/tmp/atray_meshes/knot_48x32.obj", 48, 32)
        out = subprocess.run([exe, "--obj", path, "--res", "64x64"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        assert out.returncode != 0 and ("CUDA" in out.stdout or "device" in out.stdout), out.stdout
