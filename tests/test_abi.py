"""CPU tests of the boundary: the shared library loads, exports every symbol include/snakeorb.h declares, and
fails loudly (no fallback) when no GPU is present."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "snakeorb.h")).read()
    return sorted(set(re.findall(r"SNK_API\s+[\w\s\*]+?\b(snk_\w+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    import snakeorb

    if not os.path.exists(snakeorb.LIB_PATH):
        import __graft_entry__

        __graft_entry__.build()
    out = subprocess.run(["nm", "-D", "--defined-only", snakeorb.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (snk_\w+)", out))
    decl = declared_symbols()
    assert len(decl) >= 25
    missing = [s for s in decl if s not in exported]
    assert not missing, missing
    lib = snakeorb.lib()
    for s in decl:
        assert hasattr(lib, s)
    assert b"sm_100a" in lib.snk_version()


def test_sass_is_sm100_only():
    import snakeorb

    out = subprocess.run(["cuobjdump", "-lelf", snakeorb.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+\w*)", out))
    assert archs == {"100a"}, archs


def test_fails_loudly_without_gpu():
    import snakeorb

    try:
        import torch

        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("GPU present")
    with pytest.raises(snakeorb.SnakeOrbError):
        snakeorb.ORBExtractor()
    with pytest.raises(snakeorb.SnakeOrbError):
        snakeorb.Context()


def test_struct_layouts_match_header():
    import ctypes as C

    import snakeorb

    assert C.sizeof(snakeorb.OrbParams) == 28
    assert C.sizeof(snakeorb.StereoParams) == 8 + 8 + 16 * 8
    assert C.sizeof(snakeorb.Camera) == 48
    assert C.sizeof(snakeorb.GridView) == 6 * 8 + 8 + 8
    assert C.sizeof(snakeorb.ScalePyramidC) == 8 + 8 + 128
    assert snakeorb.POINT_FINE.itemsize == 96 and snakeorb.POINT_COARSE.itemsize == 96 and snakeorb.POINT_KF.itemsize == 64


def test_adapter_header_compiles_standalone():
    hdr = os.path.join(ROOT, "snake-slam_b200", "host", "snakeorb_saiga_adapter.h")
    src = '#include "%s"\nint main(){ return 0; }\n' % hdr
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    r = subprocess.run([cxx, "-std=c++17", "-fsyntax-only", "-x", "c++", "-"], input=src, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
