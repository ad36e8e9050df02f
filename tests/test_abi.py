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


def test_adapter_feature_detector_cache_path(tmp_path):
    """FeatureDetector::Detect glue of the C++ adapter (snakeorb_saiga_adapter.h: DetectFrameB200), compiled and run:
    with fd_bufferToFile on and both cache files present nothing is extracted (no GPU needed) and the frame gets
    exactly the cached features; the files are the reference's BinaryFile layout written by snk_features_write."""
    import numpy as np

    import snakeorb

    rng = np.random.default_rng(3)
    for name, n in (("17.features", 37), ("17_right.features", 0)):
        k = np.zeros(n, snakeorb.KP64)
        k["x"], k["y"], k["octave"] = rng.uniform(0, 700, n), rng.uniform(0, 400, n), rng.integers(0, 8, n)
        d = rng.integers(0, 2**63, (n, 4)).astype(np.uint64)
        snakeorb.features_write(str(tmp_path / name), k, d)
    src = tmp_path / "t.cpp"
    src.write_text('''
#include "snakeorb_saiga_adapter.h"
#include <cstring>
int main(int argc, char** argv)
{
    using namespace snkadapter;
    // the extractor is never used on the cache path; an unusable handle must not matter
    alignas(ORBExtractorB200) unsigned char raw[sizeof(ORBExtractorB200)];
    std::memset(raw, 0, sizeof(raw));
    ORBExtractorB200& ex = *reinterpret_cast<ORBExtractorB200*>(raw);
    ImageView<unsigned char> img, right;
    FrameFeaturesB200 f;
    int computed = DetectFrameB200(ex, true, argv[1], 17, img, &right, f);
    std::printf("%d %zu %zu %zu %zu %.3f\\n", computed, f.keypoints.size(), f.descriptors.size(), f.keypoints_right.size(),
                f.descriptors_right.size(), f.keypoints.empty() ? 0.0 : f.keypoints[5].x);
    return 0;
}
''')
    exe = tmp_path / "t"
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    libdir = os.path.join(ROOT, "snake-slam_b200")
    r = subprocess.run([cxx, "-std=c++17", "-I", os.path.join(libdir, "host"), "-I", os.path.join(ROOT, "include"), str(src),
                        "-o", str(exe), "-L", libdir, "-lsnakeorb", "-Wl,-rpath," + libdir], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    out = subprocess.run([str(exe), str(tmp_path)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    k, d = snakeorb.features_read(str(tmp_path / "17.features"))
    computed, nk, nd, nkr, ndr, x5 = out.stdout.split()
    assert (int(computed), int(nk), int(nd), int(nkr), int(ndr)) == (0, 37, 37, 0, 0)
    assert abs(float(x5) - k["x"][5]) < 1e-3
