// snakeorb_saiga_adapter.h — header-only C++ binding of libsnakeorb.so with the reference's class shapes.
//
// Inside the reference tree, define SNAKEORB_WITH_SAIGA before including it: the adapter then takes Saiga's own
// ImageView / KeyPoint / DescriptorORB types, and `Snake/Preprocess/FeatureDetector.cpp:31-41` only has to construct
// `Saiga::ORBExtractorB200` instead of `Saiga::ORBExtractor` / `Saiga::ORBExtractorGPU`.
// Without the macro it ships layout-identical stand-ins so it compiles (and is compile-tested) on its own.
//
// Error convention: the reference aborts through SAIGA_ASSERT / SAIGA_EXIT_ERROR (saiga/core/util/assert.h;
// FeatureDetector.cpp:35); the library returns codes — the adapter is where a code becomes an abort.
#pragma once
#include "../../include/snakeorb.h"

#include <algorithm>
#include <array>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <utility>
#include <vector>

#ifdef SNAKEORB_WITH_SAIGA
#    include "saiga/core/image/imageView.h"
#    include "saiga/vision/features/Features.h"
#    include "saiga/vision/util/ScalePyramid.h"
namespace snkadapter
{
template <typename T>
using ImageView = Saiga::ImageView<T>;
using KeyPointF = Saiga::KeyPoint<float>;
using KeyPointD = Saiga::KeyPoint<double>;
using DescriptorORB = Saiga::DescriptorORB;
}  // namespace snkadapter
#else
namespace snkadapter
{
// saiga/core/image/imageView.h:79-160 — non-owning {height, width, pitchBytes, data}
template <typename T>
struct ImageView
{
    int rows = 0, cols = 0;
    size_t pitchBytes = 0;
    T* data           = nullptr;
    bool empty() const { return rows <= 0 || cols <= 0 || !data; }
};
struct KeyPointF  // saiga/vision/features/Features.h:21-67, T = float
{
    float x, y, size, angle, response;
    int octave;
};
struct alignas(16) KeyPointD  // T = double (Snake::KeyPoint)
{
    double x, y, size, angle, response;
    int octave;
};
using DescriptorORB = std::array<uint64_t, 4>;  // Features.h:70
}  // namespace snkadapter
#endif

namespace snkadapter
{
static_assert(sizeof(KeyPointF) == sizeof(snk_keypoint_f32), "KeyPoint<float> must be 24 bytes");
static_assert(sizeof(KeyPointD) == sizeof(snk_keypoint_f64), "KeyPoint<double> must be 48 bytes");
static_assert(sizeof(DescriptorORB) == 32, "DescriptorORB must be 32 bytes");

inline void check(int rc, const char* what)
{
    if (rc != SNK_OK)
    {
        std::fprintf(stderr, "snakeorb: %s failed (%d): %s\n", what, rc, snk_last_error());
        std::abort();  // == SAIGA_ASSERT behaviour
    }
}

// Drop-in for Saiga::ORBExtractor (ORBExtractor.h:44-90) and Saiga::ORBExtractorGPU (ORBExtractorGPU.h:38-49).
// Same constructor arguments; `num_threads` is accepted and ignored. Not thread-safe, exactly like the originals.
class ORBExtractorB200
{
   public:
    ORBExtractorB200(int nfeatures, float scaleFactor, int num_levels, int th_fast, int th_fast_min, int /*num_threads*/ = 1,
                     int device = 0)
    {
        snk_orb_params p{nfeatures, scaleFactor, num_levels, th_fast, th_fast_min, /*max_batch=*/2, device};
        check(snk_orb_create(&p, &h_), "snk_orb_create");
        num_levels_ = num_levels;
    }
    ~ORBExtractorB200() { snk_orb_destroy(h_); }
    ORBExtractorB200(const ORBExtractorB200&)            = delete;
    ORBExtractorB200& operator=(const ORBExtractorB200&) = delete;

    // void ORBExtractor::Detect(ImageView<uchar>, std::vector<KeyPoint<float>>&, std::vector<DescriptorORB>&)
    // (ORBExtractor.cpp:178). Returns the number of keypoints like ORBExtractorGPU::Detect does.
    int Detect(ImageView<unsigned char> image, std::vector<KeyPointF>& keypoints, std::vector<DescriptorORB>& descriptors)
    {
        descriptors.clear();
        if (image.empty()) return 0;  // ORBExtractor.cpp:182
        const int cap = snk_orb_max_keypoints(h_);
        keypoints.resize(cap);
        descriptors.resize(cap);
        int n = 0;
#ifdef SNAKEORB_WITH_SAIGA
        const unsigned char* ptr = image.data8;
        const int rows = image.rows, cols = image.cols;
#else
        const unsigned char* ptr = image.data;
        const int rows = image.rows, cols = image.cols;
#endif
        check(snk_orb_detect(h_, ptr, rows, cols, image.pitchBytes, /*img_is_device=*/0,
                             reinterpret_cast<snk_keypoint_f32*>(keypoints.data()),
                             reinterpret_cast<uint64_t*>(descriptors.data()), cap, &n),
              "snk_orb_detect");
        keypoints.resize(n);
        descriptors.resize(n);
        return n;
    }

    // Left and right image of a stereo frame (FeatureDetector.cpp:124 + :154 back to back) in ONE batched launch sequence
    // (the handle is created with max_batch = 2): both images are queued as a batch of two, results fetched once.
    void DetectStereo(ImageView<unsigned char> left, ImageView<unsigned char> right, std::vector<KeyPointF>& kl,
                      std::vector<DescriptorORB>& dl, std::vector<KeyPointF>& kr, std::vector<DescriptorORB>& dr)
    {
        if (left.empty() || right.empty() || left.rows != right.rows || left.cols != right.cols || left.pitchBytes != right.pitchBytes)
        {
            Detect(left, kl, dl);  // shapes differ (or an image is missing): two independent calls, as the reference does
            Detect(right, kr, dr);
            return;
        }
#ifdef SNAKEORB_WITH_SAIGA
        const unsigned char *pl = left.data8, *pr = right.data8;
#else
        const unsigned char *pl = left.data, *pr = right.data;
#endif
        const int cap = snk_orb_max_keypoints(h_);
        const size_t img_bytes = left.pitchBytes * (size_t)left.rows;
        std::vector<KeyPointF> k2(2 * (size_t)cap);
        std::vector<DescriptorORB> d2(2 * (size_t)cap);
        int n[2] = {0, 0};
        // the batch entry point takes one base pointer and an image stride: stage the pair contiguously unless it already is
        const unsigned char* base = pl;
        size_t stride             = (size_t)(pr - pl);
        if (pr < pl || stride < img_bytes || stride % 16 != 0)
        {
            stage_.resize(2 * img_bytes);
            std::copy(pl, pl + img_bytes, stage_.begin());
            std::copy(pr, pr + img_bytes, stage_.begin() + img_bytes);
            base = stage_.data(), stride = img_bytes;
        }
        check(snk_orb_detect_batch(h_, base, 2, left.rows, left.cols, left.pitchBytes, stride, /*imgs_are_device=*/0,
                                   reinterpret_cast<snk_keypoint_f32*>(k2.data()), reinterpret_cast<uint64_t*>(d2.data()), cap, n),
              "snk_orb_detect_batch");
        kl.assign(k2.begin(), k2.begin() + n[0]);
        dl.assign(d2.begin(), d2.begin() + n[0]);
        kr.assign(k2.begin() + cap, k2.begin() + cap + n[1]);
        dr.assign(d2.begin() + cap, d2.begin() + cap + n[1]);
    }

    // ScalePyramid accessors used by the matchers (ScalePyramid.h:78-84)
    std::vector<double> Scales() const
    {
        std::vector<double> s(num_levels_);
        check(snk_orb_pyramid(h_, s.data(), nullptr), "snk_orb_pyramid");
        return s;
    }
    std::vector<int32_t> Features() const
    {
        std::vector<int32_t> f(num_levels_);
        check(snk_orb_pyramid(h_, nullptr, f.data()), "snk_orb_pyramid");
        return f;
    }
    snk_orb* handle() { return h_; }

   private:
    snk_orb* h_     = nullptr;
    int num_levels_ = 0;
    std::vector<unsigned char> stage_;  // DetectStereo: the pair side by side when the caller's images are not
};

// Drop-in for Saiga::BruteForceMatcher<DescriptorORB> (Features.h:299-423): same public members.
struct BruteForceMatcherB200
{
    explicit BruteForceMatcherB200(int device = 0) { check(snk_ctx_create(device, &c_), "snk_ctx_create"); }
    ~BruteForceMatcherB200() { snk_ctx_destroy(c_); }
    BruteForceMatcherB200(const BruteForceMatcherB200&)            = delete;
    BruteForceMatcherB200& operator=(const BruteForceMatcherB200&) = delete;

    void matchKnn2(const std::vector<DescriptorORB>& d1, const std::vector<DescriptorORB>& d2)
    {
        knn2.resize(d1.size());
        check(snk_match_knn2(c_, reinterpret_cast<const uint64_t*>(d1.data()), (int)d1.size(),
                             reinterpret_cast<const uint64_t*>(d2.data()), (int)d2.size(), knn2.data()),
              "snk_match_knn2");
    }
    void matchKnn2_omp(const std::vector<DescriptorORB>& d1, const std::vector<DescriptorORB>& d2, int /*threads*/)
    {
        matchKnn2(d1, d2);
    }
    int filterMatches(int threshold, float ratioThreshold)
    {
        std::vector<int32_t> ij(2 * knn2.size() + 2);
        int n = 0;
        check(snk_filter_matches(knn2.data(), (int)knn2.size(), threshold, ratioThreshold, ij.data(), &n), "snk_filter_matches");
        matches.clear();
        for (int i = 0; i < n; ++i) matches.push_back({ij[2 * i], ij[2 * i + 1]});
        return n;
    }
    std::vector<snk_knn2> knn2;                   // {dist0, idx0, dist1, idx1} == knn2(i,0/1).{first,second}
    std::vector<std::pair<int, int>> matches;     // == BruteForceMatcher::matches
    snk_ctx* context() { return c_; }

   private:
    snk_ctx* c_ = nullptr;
};

// Preprocess::StereoMatching (Preprocess.cpp:122-242) body after the two Rectification::Forward loops.
// left/right: rectified keypoints; right_points/depth: Frame::right_points / Frame::depth (resized by the caller's
// Frame::allocateTmp); returns the number of matches.
inline int StereoMatchingB200(snk_ctx* c, double bf, bool relaxed, const std::vector<double>& scales,
                              const std::vector<KeyPointD>& left, const std::vector<DescriptorORB>& dl,
                              const std::vector<KeyPointD>& right, const std::vector<DescriptorORB>& dr,
                              std::vector<float>& right_points, std::vector<float>& depth)
{
    snk_stereo_params p{};
    p.bf      = bf;
    p.relaxed = relaxed;
    p.nlevels = (int)scales.size();
    for (size_t i = 0; i < scales.size() && i < 16; ++i) p.scale[i] = scales[i];
    right_points.resize(left.size());
    depth.resize(left.size());
    int n = 0;
    check(snk_stereo_match(c, &p, reinterpret_cast<const snk_keypoint_f64*>(left.data()),
                           reinterpret_cast<const uint64_t*>(dl.data()), (int)left.size(),
                           reinterpret_cast<const snk_keypoint_f64*>(right.data()),
                           reinterpret_cast<const uint64_t*>(dr.data()), (int)right.size(), right_points.data(),
                           depth.data(), &n),
          "snk_stereo_match");
    return n;
}

// FeatureDetector::Detect (Snake/Preprocess/FeatureDetector.cpp:86-175) for one frame: per image the `.features`
// cache is read when settings.fd_bufferToFile is on and the file exists (:94-111), otherwise the image is extracted,
// the keypoints are widened to double (:129-132) and the cache file is written (:134-139). File names as at :91-92.
// Returns how many of the two images were actually extracted (0 when both came from the cache).
struct FrameFeaturesB200
{
    std::vector<KeyPointD> keypoints, keypoints_right;
    std::vector<DescriptorORB> descriptors, descriptors_right;
};
inline bool file_exists_b200(const std::string& path)
{
    if (FILE* f = std::fopen(path.c_str(), "rb"))
    {
        std::fclose(f);
        return true;
    }
    return false;
}
inline bool read_features_b200(const std::string& path, std::vector<KeyPointD>& kps, std::vector<DescriptorORB>& desc)
{
    int n = 0;
    // first call with capacity 0 reports the count (SNK_ERR_CAPACITY), like sizing a vector before `bf >> vector`
    snk_features_read(path.c_str(), nullptr, nullptr, 0, &n);
    kps.resize(n);
    desc.resize(n);
    return snk_features_read(path.c_str(), reinterpret_cast<snk_keypoint_f64*>(kps.data()),
                             reinterpret_cast<uint64_t*>(desc.data()), n, &n) == SNK_OK;
}
inline int DetectFrameB200(ORBExtractorB200& extractor, bool fd_bufferToFile, const std::string& tmpDir, int file_id,
                           ImageView<unsigned char> image, const ImageView<unsigned char>* right_image,
                           FrameFeaturesB200& frame)
{
    bool compute_left = true, compute_right = right_image != nullptr;
    const std::string feature_file       = tmpDir + "/" + std::to_string(file_id) + ".features";
    const std::string feature_file_right = tmpDir + "/" + std::to_string(file_id) + "_right.features";
    if (fd_bufferToFile)
    {
        if (compute_left && file_exists_b200(feature_file))
            compute_left = !read_features_b200(feature_file, frame.keypoints, frame.descriptors);
        if (compute_right && file_exists_b200(feature_file_right))
            compute_right = !read_features_b200(feature_file_right, frame.keypoints_right, frame.descriptors_right);
    }
    auto run = [&](ImageView<unsigned char> img, std::vector<KeyPointD>& kps, std::vector<DescriptorORB>& desc,
                   const std::string& file) {
        std::vector<KeyPointF> mvKeys;
        extractor.Detect(img, mvKeys, desc);
        kps.resize(mvKeys.size());
        check(snk_keypoints_widen(reinterpret_cast<const snk_keypoint_f32*>(mvKeys.data()), (int)mvKeys.size(),
                                  reinterpret_cast<snk_keypoint_f64*>(kps.data())),
              "snk_keypoints_widen");
        if (fd_bufferToFile)
            check(snk_features_write(file.c_str(), reinterpret_cast<const snk_keypoint_f64*>(kps.data()),
                                     reinterpret_cast<const uint64_t*>(desc.data()), (int)kps.size()),
                  "snk_features_write");
    };
    if (compute_left) run(image, frame.keypoints, frame.descriptors, feature_file);
    if (compute_right) run(*right_image, frame.keypoints_right, frame.descriptors_right, feature_file_right);
    return (int)compute_left + (int)compute_right;
}

// MiniBow2::TemplatedVocabulary<DescriptorORB> (saiga/vision/slam/MiniBow2.h:173-460): the members Snake calls.
// BowVector / FeatureVector are the plain containers MiniBow2 derives its classes from (:43, :98).
using BowVectorB200     = std::vector<std::pair<int, float>>;
using FeatureVectorB200 = std::vector<std::pair<int, std::vector<int>>>;
class VocabularyB200
{
   public:
    // TemplatedVocabulary(const std::string& filename) -> loadRaw (:190, :1020)
    VocabularyB200(const std::string& filename, snk_ctx* ctx, int device = 0) : ctx_(ctx)
    {
        check(snk_voc_load_raw(device, filename.c_str(), &voc_), "snk_voc_load_raw");
        check(snk_voc_info(voc_, &k_, &L_, nullptr, &words_), "snk_voc_info");
    }
    ~VocabularyB200() { snk_voc_destroy(voc_); }
    VocabularyB200(const VocabularyB200&)            = delete;
    VocabularyB200& operator=(const VocabularyB200&) = delete;
    unsigned int size() const { return (unsigned int)words_; }
    bool empty() const { return words_ == 0; }
    int getBranchingFactor() const { return k_; }
    int getDepthLevels() const { return L_; }
    // transform(features, v, fv, levelsup, num_threads) (:883-931); num_threads is meaningless here
    void transform(const std::vector<DescriptorORB>& features, BowVectorB200& v, FeatureVectorB200& fv, int levelsup,
                   int /*num_threads*/ = 1) const
    {
        const int n = (int)features.size();
        std::vector<int32_t> ids(n + 1), nodes(n + 1), offs(n + 2), idx(n + 1);
        std::vector<float> vals(n + 1);
        int nb = 0, nf = 0;
        check(snk_bow_transform(ctx_, voc_, reinterpret_cast<const uint64_t*>(features.data()), n, levelsup, ids.data(),
                                vals.data(), &nb, nodes.data(), offs.data(), idx.data(), &nf),
              "snk_bow_transform");
        v.clear();
        fv.clear();
        v.reserve(nb);
        fv.reserve(nf);
        for (int i = 0; i < nb; ++i) v.emplace_back(ids[i], vals[i]);
        for (int i = 0; i < nf; ++i) fv.emplace_back(nodes[i], std::vector<int>(idx.begin() + offs[i], idx.begin() + offs[i + 1]));
    }
    const snk_voc* handle() const { return voc_; }

   private:
    snk_ctx* ctx_  = nullptr;
    snk_voc* voc_  = nullptr;
    int k_ = 0, L_ = 0, words_ = 0;
};

// Preprocess::undistortKeypoints (Snake/Preprocess/Preprocess.cpp:55-77). rect: K_src fx fy cx cy | D_src.Coeffs() |
// R.coeffs() (x y z w) | K_dst fx fy cx cy s — the fields of Saiga::Rectification, see snk_rectify.
inline void undistortKeypointsB200(snk_ctx* c, const snk_rectify& rect, const std::vector<KeyPointD>& keypoints,
                                   std::vector<KeyPointD>& undistorted_keypoints, std::vector<double>& normalized_points_xy)
{
    undistorted_keypoints.resize(keypoints.size());
    normalized_points_xy.resize(2 * keypoints.size());
    check(snk_undistort_keypoints(c, &rect, reinterpret_cast<const snk_keypoint_f64*>(keypoints.data()), (int)keypoints.size(),
                                  reinterpret_cast<snk_keypoint_f64*>(undistorted_keypoints.data()), normalized_points_xy.data()),
          "snk_undistort_keypoints");
}
}  // namespace snkadapter

#ifdef SNAKEORB_WITH_SAIGA
namespace Saiga
{
using ORBExtractorB200 = snkadapter::ORBExtractorB200;
}
#endif
