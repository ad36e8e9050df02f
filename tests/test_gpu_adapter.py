"""The C++ host side above the C ABI (snake-slam_b200/host/snakeorb_saiga_adapter.h), compiled with g++ on the box and
run as its own process: ORBExtractorB200::Detect, BruteForceMatcherB200::matchKnn2 / filterMatches, StereoMatchingB200,
VocabularyB200::transform and undistortKeypointsB200 must return exactly what the oracle returns for the same inputs
(the oracle is the checker; the program under test links libsnakeorb.so only).

The program reads its inputs from raw files the test writes and dumps every result as raw little-endian arrays."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIBDIR = os.path.join(ROOT, "snake-slam_b200")
VOC = os.path.join(HERE, "golden", "voc_k9_L3.bin")

PROGRAM_HEAD = r'''
#include "snakeorb_saiga_adapter.h"
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>
using namespace snkadapter;

template <typename T>
static std::vector<T> slurp(const std::string& path)
{
    std::vector<T> v;
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) { std::perror(path.c_str()); std::exit(3); }
    std::fseek(f, 0, SEEK_END);
    long bytes = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    v.resize(bytes / sizeof(T));
    if (bytes && std::fread(v.data(), 1, bytes, f) != (size_t)bytes) std::exit(3);
    std::fclose(f);
    return v;
}
template <typename T>
static void dump(const std::string& path, const T* p, size_t n)
{
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) { std::perror(path.c_str()); std::exit(3); }
    if (n) std::fwrite(p, sizeof(T), n, f);
    std::fclose(f);
}
'''

PROGRAM_MAIN = r'''
int main(int argc, char** argv)
{
    if (argc < 6) return 2;
    const std::string dir = argv[1];
    const int rows = std::atoi(argv[2]), cols = std::atoi(argv[3]), pitch = std::atoi(argv[4]);
    const std::string voc_file = argv[5];

    // --- ORBExtractor::Detect (ORBExtractor.cpp:178) on a pitched view, twice on one extractor, plus the empty image
    auto left = slurp<unsigned char>(dir + "/left.raw"), right = slurp<unsigned char>(dir + "/right.raw");
    ORBExtractorB200 ex(1500, 1.2f, 8, 20, 7, /*num_threads=*/4);
    ImageView<unsigned char> li, ri, none;
    li.rows = ri.rows = rows;
    li.cols = ri.cols = cols;
    li.pitchBytes = ri.pitchBytes = pitch;
    li.data = left.data();
    ri.data = right.data();
    std::vector<KeyPointF> kl, kr, k0(3);
    std::vector<DescriptorORB> dl, dr, d0(3);
    if (ex.Detect(none, k0, d0) != 0 || !d0.empty()) return 4;
    ex.DetectStereo(li, ri, kl, dl, kr, dr);
    dump(dir + "/kl.bin", kl.data(), kl.size());
    dump(dir + "/dl.bin", dl.data(), dl.size());
    dump(dir + "/kr.bin", kr.data(), kr.size());
    dump(dir + "/dr.bin", dr.data(), dr.size());
    auto scales = ex.Scales();
    auto feats  = ex.Features();
    dump(dir + "/scales.bin", scales.data(), scales.size());
    dump(dir + "/nfeat.bin", feats.data(), feats.size());

    // --- BruteForceMatcher::matchKnn2 / filterMatches (Features.h:299-423)
    BruteForceMatcherB200 bf;
    bf.matchKnn2_omp(dl, dr, 8);
    dump(dir + "/knn2.bin", bf.knn2.data(), bf.knn2.size());
    const int nm = bf.filterMatches(50, 0.8f);
    std::vector<int32_t> flat;
    for (auto& m : bf.matches) { flat.push_back(m.first); flat.push_back(m.second); }
    if ((int)bf.matches.size() != nm) return 5;
    dump(dir + "/matches.bin", flat.data(), flat.size());

    // --- FeatureDetector's widening (:129-132) then Preprocess::StereoMatching (Preprocess.cpp:122-242)
    std::vector<KeyPointD> kl64(kl.size()), kr64(kr.size());
    check(snk_keypoints_widen(reinterpret_cast<const snk_keypoint_f32*>(kl.data()), (int)kl.size(),
                              reinterpret_cast<snk_keypoint_f64*>(kl64.data())), "widen");
    check(snk_keypoints_widen(reinterpret_cast<const snk_keypoint_f32*>(kr.data()), (int)kr.size(),
                              reinterpret_cast<snk_keypoint_f64*>(kr64.data())), "widen");
    std::vector<float> rp, depth;
    const int ns = StereoMatchingB200(bf.context(), 386.1, true, scales, kl64, dl, kr64, dr, rp, depth);
    dump(dir + "/rp.bin", rp.data(), rp.size());
    dump(dir + "/depth.bin", depth.data(), depth.size());

    // --- MiniBow2 vocabulary: loadRaw + transform (MiniBow2.h:883-931, :1020)
    VocabularyB200 voc(voc_file, bf.context());
    BowVectorB200 bv;
    FeatureVectorB200 fv;
    voc.transform(dl, bv, fv, 2, 4);
    std::vector<int32_t> ids, nodes, offs{0}, idx;
    std::vector<float> vals;
    for (auto& e : bv) { ids.push_back(e.first); vals.push_back(e.second); }
    for (auto& e : fv)
    {
        nodes.push_back(e.first);
        idx.insert(idx.end(), e.second.begin(), e.second.end());
        offs.push_back((int32_t)idx.size());
    }
    dump(dir + "/bow_ids.bin", ids.data(), ids.size());
    dump(dir + "/bow_vals.bin", vals.data(), vals.size());
    dump(dir + "/fv_nodes.bin", nodes.data(), nodes.size());
    dump(dir + "/fv_offs.bin", offs.data(), offs.size());
    dump(dir + "/fv_idx.bin", idx.data(), idx.size());

    // --- Preprocess::undistortKeypoints (Preprocess.cpp:55-77)
    auto rect_raw = slurp<snk_rectify>(dir + "/rect.bin");
    if (rect_raw.size() != 1) return 6;
    std::vector<KeyPointD> und;
    std::vector<double> nrm;
    undistortKeypointsB200(bf.context(), rect_raw[0], kl64, und, nrm);
    dump(dir + "/und.bin", und.data(), und.size());
    dump(dir + "/nrm.bin", nrm.data(), nrm.size());

    std::printf("%zu %zu %d %d %u %d %d\n", kl.size(), kr.size(), nm, ns, voc.size(), voc.getBranchingFactor(),
                voc.getDepthLevels());
    return 0;
}
'''


def _compile(tmp_path):
    src = tmp_path / "adapter_main.cpp"
    src.write_text(PROGRAM_HEAD + PROGRAM_MAIN)
    exe = tmp_path / "adapter_main"
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    r = subprocess.run([cxx, "-std=c++17", "-O1", "-Wall", "-I", os.path.join(LIBDIR, "host"), "-I", os.path.join(ROOT, "include"),
                        str(src), "-o", str(exe), "-L", LIBDIR, "-lsnakeorb", "-Wl,-rpath," + LIBDIR],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_cpp_adapter_equals_oracle(tmp_path, oracle):
    import snakeorb as so
    from snakeorb import synthetic as S

    rows, cols, pitch = 376, 1241, 1280
    left = S.image(9100, cols, rows)
    right = S.stereo_right(left, 9101, 4.0, 40.0)
    for name, img in (("left.raw", left), ("right.raw", right)):
        buf = np.full((rows, pitch), 0xAB, np.uint8)  # padding bytes must never be read as pixels
        buf[:, :cols] = img
        buf.tofile(tmp_path / name)
    K = [718.856, 718.856, 607.19, 185.22, 0.0]
    q = np.array([0.001, -0.002, 0.0005, 1.0])
    q /= np.sqrt((q * q).sum())
    rect = so.Rectify.from_vector(np.array(K + [-0.12, 0.03, 0, 0, 0, 0, 0.0004, -0.0002] + list(q) + K, np.float64))
    (tmp_path / "rect.bin").write_bytes(bytes(rect))

    exe = _compile(tmp_path)
    out = subprocess.run([str(exe), str(tmp_path), str(rows), str(cols), str(pitch), VOC], capture_output=True, text=True,
                         timeout=300)
    assert out.returncode == 0, (out.returncode, out.stdout, out.stderr)
    nl, nr, nm, ns, words, k, L = [int(x) for x in out.stdout.split()]
    assert (words, k, L) == (729, 9, 3)

    def load(name, dtype):
        return np.fromfile(tmp_path / name, dtype)

    # extraction
    ref = oracle.Extractor(1500, 1.2, 8, 20, 7, 4)
    kl, dl = ref.detect(left)
    kr, dr = ref.detect(right)
    assert nl == len(kl) > 1000 and nr == len(kr) > 1000
    assert load("kl.bin", so.KP32).tobytes() == np.ascontiguousarray(kl, so.KP32).tobytes()
    assert load("kr.bin", so.KP32).tobytes() == np.ascontiguousarray(kr, so.KP32).tobytes()
    assert np.array_equal(load("dl.bin", np.uint64).reshape(-1, 4), dl)
    assert np.array_equal(load("dr.bin", np.uint64).reshape(-1, 4), dr)
    scales, nfeat = oracle.scale_pyramid(8, 1.2, 1500)
    assert load("scales.bin", np.float64).tobytes() == np.asarray(scales, np.float64).tobytes()
    assert np.array_equal(load("nfeat.bin", np.int32), nfeat)

    # brute-force matcher
    knn = oracle.match_knn2(dl, dr)
    assert load("knn2.bin", np.uint8).tobytes() == knn.tobytes()
    want = oracle.filter_matches(knn, 50, 0.8)
    assert nm == len(want) > 200
    assert np.array_equal(load("matches.bin", np.int32).reshape(-1, 2), want)

    # stereo row search
    a, b = oracle.kp32_to_kp64(kl), oracle.kp32_to_kp64(kr)
    n2, rp2, dp2, _ = oracle.stereo_match(oracle.stereo_params(386.1, True, scales), a, dl, b, dr)
    assert ns == n2 > 100
    assert load("rp.bin", np.float32).tobytes() == rp2.tobytes() and load("depth.bin", np.float32).tobytes() == dp2.tobytes()

    # vocabulary
    ids, vals, fn, fo, fi = oracle.Vocabulary(VOC).transform(dl, 2)
    assert np.array_equal(load("bow_ids.bin", np.int32), ids) and load("bow_vals.bin", np.float32).tobytes() == vals.tobytes()
    assert np.array_equal(load("fv_nodes.bin", np.int32), fn) and np.array_equal(load("fv_offs.bin", np.int32), fo)
    assert np.array_equal(load("fv_idx.bin", np.int32), fi)

    # undistortion
    un, nrm = oracle.undistort_keypoints(rect, a)
    assert load("und.bin", so.KP64).tobytes() == un.tobytes() and load("nrm.bin", np.float64).tobytes() == nrm.tobytes()
