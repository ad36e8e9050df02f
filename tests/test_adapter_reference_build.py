"""Drop-in check at the source level: the reference's own Snake/Preprocess/FeatureDetector.cpp — the only caller of
Saiga::ORBExtractor / ORBExtractorGPU — compiles UNCHANGED when snake-slam_b200/host/shadow is put in front of Saiga's
include directory, with and without SNAKE_CUDA, and then calls libsnakeorb's C ABI. A second translation unit uses the
rest of the original's public interface (GetImage, getPyramid). Needs /root/reference (this container); no GPU."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
SAIGA = REF + "/External/saiga"
INC = ["-I", ROOT + "/snake-slam_b200/host/shadow", "-I", ROOT + "/oracle/ref_stubs_snake", "-I", ROOT + "/oracle/ref_stubs",
       "-I", REF + "/Snake", "-I", REF + "/Snake/System", "-I", SAIGA + "/src", "-I", SAIGA + "/submodules/eigen"]
CXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"

pytestmark = pytest.mark.skipif(not os.path.isdir(REF + "/Snake"), reason="the reference tree is not on this machine")


def undefined_symbols(obj):
    return subprocess.run(["nm", "-u", obj], capture_output=True, text=True, check=True).stdout


@pytest.mark.parametrize("flags", [[], ["-DSNAKE_CUDA"]])
def test_reference_feature_detector_compiles_unchanged_against_the_shadow_headers(tmp_path, flags):
    obj = str(tmp_path / "FeatureDetector.o")
    r = subprocess.run([CXX, "-std=c++17", "-O1", "-fPIC", "-fopenmp", "-w", "-c", *flags, *INC,
                        REF + "/Snake/Preprocess/FeatureDetector.cpp", "-o", obj], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    syms = undefined_symbols(obj)
    for s in ("snk_orb_create", "snk_orb_detect", "snk_orb_destroy"):
        assert s in syms, s
    assert "ORBExtractor" not in syms  # nothing of Saiga's own extractor is left to link


def test_shadow_extractor_offers_the_originals_public_interface(tmp_path):
    src = tmp_path / "use.cpp"
    src.write_text('''
#include "saiga/vision/features/ORBExtractor.h"
#include "saiga/cuda/imageProcessing/ORBExtractorGPU.h"
int use(Saiga::ImageView<unsigned char> img)
{
    std::vector<Saiga::KeyPoint<float>> kps;
    std::vector<Saiga::DescriptorORB> desc;
    Saiga::ORBExtractor ex(1000, 1.2f, 8, 20, 7, 2);            // ORBExtractor.h:49
    ex.Detect(img, kps, desc);                                   // :52
    Saiga::ImageView<unsigned char> lvl = ex.GetImage(3);        // :57
    Saiga::ScalePyramid pyr = ex.getPyramid();                   // :61
    Saiga::ORBExtractorGPU gpu(1000, 1.2f, 8, 20, 7);            // ORBExtractorGPU.h:41
    int n = gpu.Detect(img, kps, desc);                          // :43, returns the count
    return n + lvl.rows + pyr.Features(0);
}
''')
    obj = str(tmp_path / "use.o")
    r = subprocess.run([CXX, "-std=c++17", "-O1", "-fPIC", "-w", "-c", *INC, str(src), "-o", obj], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    syms = undefined_symbols(obj)
    assert "snk_orb_read_level" in syms and "snk_orb_get_level" in syms
