// TEST INFRASTRUCTURE ONLY. Thin extern "C" shim over the reference's own classes, compiled together with the
// reference's unmodified translation units (see Makefile: REF_TUS) into oracle/_ref/libsaiga_orb_ref.so.
// Used to pin the restatement in orb_ref.cpp / matchers_ref.cpp and to generate tests/golden.
#include "saiga/vision/features/FeatureDistribution.h"
#include "saiga/vision/features/Features.h"
#include "saiga/vision/features/OrbDescriptors.h"
#include "saiga/vision/features/OrbPattern.h"
#include "saiga/vision/util/FeatureGrid2.h"
#include "saiga/vision/util/ScalePyramid.h"

#include <cstring>
#include <vector>

using namespace Saiga;
#define REF_API extern "C" __attribute__((visibility("default")))

REF_API void ref_scale_pyramid(int levels, float scale_factor, int nfeatures, double* scale, int* features)
{
    ScalePyramid p(levels, scale_factor, nfeatures);
    for (int i = 0; i < levels; ++i)
    {
        scale[i]    = p.Scale(i);
        features[i] = p.Features(i);
    }
}
REF_API void ref_scale_pyramid_fn(int levels, float scale_factor, double ref_dist, int ref_level, double dist,
                                  double* out4)
{
    ScalePyramid p(levels, scale_factor, 1000);
    auto mm = p.EstimateMinMaxDistance(ref_dist, ref_level);
    out4[0] = mm.first;
    out4[1] = mm.second;
    out4[2] = p.PredictScaleLevel(ref_dist, ref_level, dist);
    out4[3] = p.ScaleForContiniousLevel(out4[2]);
}
REF_API void ref_umax(int* out16)
{
    auto u = ORBPattern::AngleUmax();
    for (int i = 0; i < 16; ++i) out16[i] = u[i];
}
REF_API void ref_pattern(int* out1024)
{
    auto p = ORBPattern::DescriptorPattern();
    for (int i = 0; i < 512; ++i)
    {
        out1024[2 * i]     = p[i].x();
        out1024[2 * i + 1] = p[i].y();
    }
}
REF_API void ref_angles(const unsigned char* img, int rows, int cols, size_t pitch, const float* xy, int n, float* out)
{
    ORB orb;
    ImageView<unsigned char> view(rows, cols, pitch, (void*)img);
    for (int i = 0; i < n; ++i) out[i] = orb.ComputeAngle(view, vec2(xy[2 * i], xy[2 * i + 1]));
}
REF_API void ref_descriptors(const unsigned char* img, int rows, int cols, size_t pitch, const float* xy,
                             const float* angle, int n, uint64_t* out)
{
    ORB orb;
    ImageView<unsigned char> view(rows, cols, pitch, (void*)img);
    for (int i = 0; i < n; ++i)
    {
        DescriptorORB d = orb.ComputeDescriptor(view, vec2(xy[2 * i], xy[2 * i + 1]), angle[i]);
        std::memcpy(out + 4 * (size_t)i, d.data(), 32);
    }
}
static_assert(sizeof(KeyPoint<float>) == 24, "KeyPoint<float> layout");
static_assert(sizeof(KeyPoint<double>) == 48, "KeyPoint<double> layout");
REF_API int ref_distribute(KeyPoint<float>* kps, int n, float minx, float miny, float maxx, float maxy, int target,
                           KeyPoint<float>* out, int cap)
{
    QuadtreeFeatureDistributor d;
    auto r = d.Distribute(ArrayView<KeyPoint<float>>(kps, n), vec2(minx, miny), vec2(maxx, maxy), target);
    for (int i = 0; i < (int)r.size() && i < cap; ++i) out[i] = r[i];
    return (int)r.size();
}
REF_API int ref_distance(const uint64_t* a, const uint64_t* b)
{
    DescriptorORB x, y;
    std::memcpy(x.data(), a, 32);
    std::memcpy(y.data(), b, 32);
    return distance(x, y);
}
REF_API void ref_knn2(const uint64_t* d1, int n, const uint64_t* d2, int m, int* out4)
{
    std::vector<DescriptorORB> a(n), b(m);
    std::memcpy(a.data(), d1, 32 * (size_t)n);
    std::memcpy(b.data(), d2, 32 * (size_t)m);
    BruteForceMatcher<DescriptorORB> bf;
    bf.matchKnn2(a, b);
    for (int i = 0; i < n; ++i)
    {
        out4[4 * i]     = bf.knn2(i, 0).first;
        out4[4 * i + 1] = bf.knn2(i, 0).second;
        out4[4 * i + 2] = bf.knn2(i, 1).first;
        out4[4 * i + 3] = bf.knn2(i, 1).second;
    }
}
REF_API int ref_knn2_filter(const uint64_t* d1, int n, const uint64_t* d2, int m, int th, float ratio, int* ij)
{
    std::vector<DescriptorORB> a(n), b(m);
    std::memcpy(a.data(), d1, 32 * (size_t)n);
    std::memcpy(b.data(), d2, 32 * (size_t)m);
    BruteForceMatcher<DescriptorORB> bf;
    bf.matchKnn2(a, b);
    int c = bf.filterMatches(th, ratio);
    for (int i = 0; i < c; ++i)
    {
        ij[2 * i]     = bf.matches[i].first;
        ij[2 * i + 1] = bf.matches[i].second;
    }
    return c;
}
// FeatureGrid2::create with explicit bounds; cells out as (begin,end) per row-major cell; returns permutation
REF_API void ref_feature_grid(double bminx, double bminy, double bmaxx, double bmaxy, int rows, int cols,
                              const KeyPoint<double>* kps, int n, int* permutation, int* cells)
{
    FeatureGridBounds2<double, 20> b;
    b.bmin        = Vec2(bminx, bminy);
    b.bmax        = Vec2(bmaxx, bmaxy);
    b.cellSize    = Vec2(20, 20);
    b.cellSizeInv = b.cellSize.array().inverse();
    b.Rows        = rows;
    b.Cols        = cols;
    std::vector<KeyPoint<double>> v(kps, kps + n);
    FeatureGrid2 g;
    auto perm = g.create(b, v);
    for (int i = 0; i < n; ++i) permutation[i] = perm[i];
    for (int r = 0; r < rows; ++r)
        for (int c = 0; c < cols; ++c)
        {
            cells[2 * (r * cols + c)]     = g.grid(r, c).first;
            cells[2 * (r * cols + c) + 1] = g.grid(r, c).second;
        }
}

// ---- Preprocess::undistortKeypoints's per-keypoint math with the reference's own camera-model headers --------
#include "saiga/vision/cameraModel/Rectify.h"
// rect: K_src fx fy cx cy s | D_src k1..k6 p1 p2 | R x y z w | K_dst fx fy cx cy s   (22 doubles)
REF_API void ref_undistort_points(const double* rect, const double* xy, int n, double* normalized, double* undistorted)
{
    Rectification r;
    r.K_src = IntrinsicsPinholed(rect[0], rect[1], rect[2], rect[3], rect[4]);
    Eigen::Matrix<double, 8, 1> c;
    for (int i = 0; i < 8; ++i) c(i) = rect[5 + i];
    r.D_src = Distortion(c);
    r.R     = Quat(rect[16], rect[13], rect[14], rect[15]);
    r.K_dst = IntrinsicsPinholed(rect[17], rect[18], rect[19], rect[20], rect[21]);
    for (int i = 0; i < n; ++i)
    {
        // the statements of Snake/Preprocess/Preprocess.cpp:64-75
        Vec2 p            = Vec2(xy[2 * i], xy[2 * i + 1]);
        p                 = r.K_src.unproject2(p);
        Vec2 p_normalized = undistortPointGN(p, p, r.D_src);
        Vec3 p_r          = r.R * Vec3(p_normalized(0), p_normalized(1), 1);
        p                 = Vec2(p_r(0) / p_r(2), p_r(1) / p_r(2));
        normalized[2 * i]     = p(0);
        normalized[2 * i + 1] = p(1);
        p                      = r.K_dst.normalizedToImage(p);
        undistorted[2 * i]     = p(0);
        undistorted[2 * i + 1] = p(1);
    }
}

// ---- MiniBow2 vocabulary: the reference's own tree builder, file format and transform -------------------------
#include "saiga/vision/slam/MiniBow2.h"
// trains a vocabulary with the reference's create() (k-means++ seeded by rand()) and writes it with saveRaw()
REF_API int ref_voc_create(const uint64_t* desc, const int* counts, int n_images, int k, int L, unsigned seed,
                           const char* out_path)
{
    std::vector<std::vector<MiniBow2::Descriptor>> training(n_images);
    size_t at = 0;
    for (int i = 0; i < n_images; ++i)
    {
        training[i].resize(counts[i]);
        std::memcpy(training[i].data(), desc + 4 * at, 32 * (size_t)counts[i]);
        at += counts[i];
    }
    srand(seed);
    MiniBow2::TemplatedVocabulary<MiniBow2::Descriptor> voc(k, L);
    voc.create(training);
    voc.saveRaw(out_path);
    return (int)voc.size();
}
// loadRaw() + transform(); BowVector as (id, value) arrays, FeatureVector as CSR in list order
REF_API int ref_voc_transform(const char* voc_path, const uint64_t* desc, int n, int levelsup, int* bow_ids,
                              float* bow_vals, int* n_bow, int* fv_nodes, int* fv_offsets, int* fv_indices, int* n_fv)
{
    static std::string loaded;
    static MiniBow2::TemplatedVocabulary<MiniBow2::Descriptor>* voc = nullptr;
    if (loaded != voc_path)
    {
        delete voc;
        voc    = new MiniBow2::TemplatedVocabulary<MiniBow2::Descriptor>(std::string(voc_path));
        loaded = voc_path;
    }
    std::vector<MiniBow2::Descriptor> f(n);
    std::memcpy(f.data(), desc, 32 * (size_t)n);
    MiniBow2::BowVector bv;
    MiniBow2::FeatureVector fv;
    voc->transform(f, bv, fv, levelsup, 1);
    *n_bow = (int)bv.size();
    for (size_t i = 0; i < bv.size(); ++i) bow_ids[i] = bv[i].first, bow_vals[i] = bv[i].second;
    *n_fv  = (int)fv.size();
    int at = 0;
    for (size_t i = 0; i < fv.size(); ++i)
    {
        fv_nodes[i]   = fv[i].first;
        fv_offsets[i] = at;
        for (int idx : fv[i].second) fv_indices[at++] = idx;
    }
    fv_offsets[fv.size()] = at;
    return (int)voc->size();
}
// FeatureVector::score (MiniBow2.h:134-170) on two BowVectors
REF_API float ref_bow_score(const int* ids1, const float* v1, int n1, const int* ids2, const float* v2, int n2)
{
    MiniBow2::BowVector a, b;
    for (int i = 0; i < n1; ++i) a.push_back({ids1[i], v1[i]});
    for (int i = 0; i < n2; ++i) b.push_back({ids2[i], v2[i]});
    return MiniBow2::FeatureVector::score(a, b);
}


// ---- Saiga::EpipolarDistanceSquared from the reference's own translation unit (Epipolar.cpp, see Makefile) -------
#include "saiga/vision/reconstruction/Epipolar.h"
REF_API double ref_epipolar_distance_squared(const double* p1, const double* p2, const double* F_rowmajor)
{
    Mat3 F;
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) F(r, c) = F_rowmajor[3 * r + c];
    return EpipolarDistanceSquared(Vec2(p1[0], p1[1]), Vec2(p2[0], p2[1]), F);
}

// ---- candidate enumeration of Features::GetFeaturesInArea (Snake/Map/Features.cpp:13-30) on the reference's own grid
// classes: FeatureGridBounds2::minMaxCellWithRadius / cellClamped and FeatureGrid2::create / cellIt are the
// reference's; the three nested loops and the distance test follow Features.cpp (that file needs the whole Snake map
// to compile). kps must already be grid-ordered (ref_feature_grid's permutation applied). Returns the count.
REF_API int ref_features_in_area(double bminx, double bminy, double bmaxx, double bmaxy, int rows, int cols,
                                 const KeyPoint<double>* kps, int n, double px, double py, float r, int* out)
{
    FeatureGridBounds2<double, 20> b;
    b.bmin        = Vec2(bminx, bminy);
    b.bmax        = Vec2(bmaxx, bmaxy);
    b.cellSize    = Vec2(20, 20);
    b.cellSizeInv = b.cellSize.array().inverse();
    b.Rows        = rows;
    b.Cols        = cols;
    std::vector<KeyPoint<double>> v(kps, kps + n);
    FeatureGrid2 g;
    auto perm = g.create(b, v);
    for (int i = 0; i < n; ++i)
        if (perm[i] != i) return -1;  // the caller did not pass grid-ordered keypoints
    const Vec2 position(px, py);
    auto [cellMin, cellMax] = b.minMaxCellWithRadius(position, r);
    auto r2                 = r * r;
    int c                   = 0;
    for (auto cx : Range(cellMin.first, cellMax.first + 1))
        for (auto cy : Range(cellMin.second, cellMax.second + 1))
            for (auto pid : g.cellIt({cx, cy}))
                if ((v[pid].point - position).squaredNorm() < r2) out[c++] = pid;
    return c;
}

// ---- FeatureGridBounds2<double,20>::computeFromIntrinsicsDist (FeatureGrid2.h:63-104), the reference's own ------
// K: fx fy cx cy s, dist: k1..k6 p1 p2; out: bmin.x bmin.y bmax.x bmax.y cellSizeInv.x cellSizeInv.y; rows_cols[2]
REF_API void ref_grid_bounds(int w, int h, const double* K, const double* dist, int resolution, double* out6, int* rows_cols)
{
    FeatureGridBounds2<double, 20> b;
    Eigen::Matrix<double, 8, 1> c;
    for (int i = 0; i < 8; ++i) c(i) = dist[i];
    b.computeFromIntrinsicsDist(w, h, IntrinsicsPinholed(K[0], K[1], K[2], K[3], K[4]), Distortion(c), resolution);
    out6[0] = b.bmin(0), out6[1] = b.bmin(1), out6[2] = b.bmax(0), out6[3] = b.bmax(1);
    out6[4] = b.cellSizeInv(0), out6[5] = b.cellSizeInv(1);
    rows_cols[0] = b.Rows, rows_cols[1] = b.Cols;
}
