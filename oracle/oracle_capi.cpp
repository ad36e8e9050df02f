// TEST INFRASTRUCTURE ONLY — CPU oracle, never on the product path.
// extern "C" surface of liborb_oracle.so, loaded with ctypes by tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py. Nothing in snake-slam_b200/ links or loads this.
#include "cvlite.h"
#include "matchers_ref.h"
#include "orb_ref.h"

#include <algorithm>
#include <cstring>
#include <memory>
#include <vector>

using namespace orbref;

#include "bow_ref.h"
extern "C" {
#define ORC_API __attribute__((visibility("default")))

ORC_API void orc_resize(const uint8_t* src, int sh, int sw, size_t sp, uint8_t* dst, int dh, int dw, size_t dp)
{
    cvlite::resize_linear_u8(src, sh, sw, sp, dst, dh, dw, dp);
}
ORC_API void orc_rgba_to_gray(const uint8_t* src, int h, int w, size_t sp, uint8_t* dst, size_t dp)
{
    cvlite::rgba_to_gray_u8(src, h, w, sp, dst, dp);
}
ORC_API void orc_blur(const uint8_t* src, int h, int w, size_t sp, uint8_t* dst, size_t dp)
{
    cvlite::gaussian7_u8(src, h, w, sp, dst, dp);
}
// out: triples (x, y, score); returns count (may exceed cap; only cap written)
ORC_API int orc_fast(const uint8_t* roi, int h, int w, size_t pitch, int threshold, int32_t* out, int cap)
{
    std::vector<cvlite::FastKp> k;
    cvlite::fast9_16_nms(roi, h, w, pitch, threshold, k);
    for (int i = 0; i < (int)k.size() && i < cap; ++i)
    {
        out[3 * i] = k[i].x, out[3 * i + 1] = k[i].y, out[3 * i + 2] = k[i].score;
    }
    return (int)k.size();
}

ORC_API void orc_scale_pyramid(int levels, float scale_factor, int nfeatures, double* scale, int32_t* features)
{
    ScalePyramid p(levels, scale_factor, nfeatures);
    for (int i = 0; i < levels; ++i)
    {
        if (scale) scale[i] = p.scale[i];
        if (features) features[i] = p.num_features[i];
    }
}
ORC_API void orc_umax(int32_t* out16)
{
    auto u = AngleUmax();
    for (int i = 0; i < 16; ++i) out16[i] = u[i];
}
ORC_API void orc_cell_grid(int rows, int cols, int32_t* out7)
{
    CellGrid g = MakeCellGrid(rows, cols);
    out7[0] = g.nCols, out7[1] = g.nRows, out7[2] = g.wCell, out7[3] = g.hCell, out7[4] = g.minB, out7[5] = g.maxBX,
    out7[6] = g.maxBY;
}
ORC_API void orc_angles(const uint8_t* img, size_t pitch, const float* xy, int n, float* out)
{
    auto u = AngleUmax();
    for (int i = 0; i < n; ++i) out[i] = ComputeAngle(img, pitch, xy[2 * i], xy[2 * i + 1], u.data());
}
ORC_API void orc_descriptors(const uint8_t* img, size_t pitch, const float* xy, const float* angle, int n,
                             uint64_t* out)
{
    for (int i = 0; i < n; ++i)
    {
        Descriptor d = ComputeDescriptor(img, pitch, xy[2 * i], xy[2 * i + 1], angle[i]);
        std::memcpy(out + 4 * (size_t)i, d.data(), 32);
    }
}
// kps: n records of 24 bytes, permuted in place; out: result records; returns count
ORC_API int orc_distribute(KeyPointF* kps, int n, float minx, float miny, float maxx, float maxy, int target,
                           KeyPointF* out, int cap)
{
    std::vector<KeyPointF> v(kps, kps + n);
    auto r = Distribute(v, minx, miny, maxx, maxy, target);
    std::copy(v.begin(), v.end(), kps);
    for (int i = 0; i < (int)r.size() && i < cap; ++i) out[i] = r[i];
    return (int)r.size();
}

struct OrcExtractor
{
    std::unique_ptr<OrbExtractor> ex;
};
ORC_API void* orc_extractor_create(int nfeatures, float scale, int nlevels, int th, int thmin, int threads)
{
    auto* e = new OrcExtractor;
    e->ex.reset(new OrbExtractor(nfeatures, scale, nlevels, th, thmin, threads));
    return e;
}
ORC_API void orc_extractor_destroy(void* h) { delete (OrcExtractor*)h; }
ORC_API int orc_extractor_detect(void* h, const uint8_t* img, int rows, int cols, size_t pitch, KeyPointF* kps,
                                 uint64_t* desc, int cap)
{
    auto* e = (OrcExtractor*)h;
    std::vector<KeyPointF> k;
    std::vector<Descriptor> d;
    e->ex->Detect(img, rows, cols, pitch, k, d);
    int n = (int)k.size();
    for (int i = 0; i < n && i < cap; ++i)
    {
        kps[i] = k[i];
        std::memcpy(desc + 4 * (size_t)i, d[i].data(), 32);
    }
    return n;
}
ORC_API int orc_extractor_level_dims(void* h, int level, int32_t* rows, int32_t* cols)
{
    auto& L = ((OrcExtractor*)h)->ex->levels[level];
    *rows = L.rows, *cols = L.cols;
    return 0;
}
ORC_API void orc_extractor_level(void* h, int level, int blurred, uint8_t* dst)
{
    auto& L = ((OrcExtractor*)h)->ex->levels[level];
    auto& s = blurred ? L.gauss : L.image;
    std::memcpy(dst, s.data(), s.size());
}
// which: 0 raw FAST candidates (relative to (16,16)), 1 kept after distribution (+16, angle set, unscaled)
ORC_API int orc_extractor_level_kps(void* h, int level, int which, KeyPointF* out, int cap)
{
    auto& L = ((OrcExtractor*)h)->ex->levels[level];
    auto& v = which ? L.kept : L.raw;
    for (int i = 0; i < (int)v.size() && i < cap; ++i) out[i] = v[i];
    return (int)v.size();
}

ORC_API int orc_hamming(const uint64_t* a, const uint64_t* b) { return hamming(a, b); }
ORC_API void orc_match_knn2(const uint64_t* d1, int n, const uint64_t* d2, int m, snk_knn2* out, int threads)
{
    match_knn2(d1, n, d2, m, out, threads);
}
ORC_API int orc_filter_matches(const snk_knn2* k, int n, int th, float ratio, int32_t* ij)
{
    return filter_matches(k, n, th, ratio, ij);
}
ORC_API int orc_stereo_match(const snk_stereo_params* p, const snk_keypoint_f64* l, const uint64_t* dl, int nl,
                             const snk_keypoint_f64* r, const uint64_t* dr, int nr, float* rp, float* depth,
                             long long* ops)
{
    return stereo_match(p, l, dl, nl, r, dr, nr, rp, depth, ops);
}
ORC_API int orc_stereo_from_rgbd(const snk_rgbd_params* p, const snk_keypoint_f64* kp, int n, const float* img, int rows,
                                 int cols, size_t pitch, float* rp, float* depth)
{
    return stereo_from_rgbd(p, kp, n, img, rows, cols, pitch, rp, depth);
}
ORC_API void orc_undistort_keypoints(const snk_rectify* r, const snk_keypoint_f64* kps, int n, snk_keypoint_f64* kp_un,
                                     double* normalized)
{
    undistort_keypoints(r, kps, n, kp_un, normalized);
}
ORC_API void orc_feature_grid(const snk_grid_view* b, const snk_keypoint_f64* kps, int n, int32_t* perm, int32_t* cells)
{
    feature_grid(b, kps, n, perm, cells);
}
ORC_API int orc_search_projection_fine(const snk_frame_view* f, snk_point_fine* pts, int n, const snk_camera* K,
                                       const snk_scale_pyramid* sp, float th, float ratio, int32_t* k2p, int32_t* bi,
                                       int32_t* bd, uint8_t* vis, long long* ops)
{
    return search_projection_fine(f, pts, n, K, sp, th, ratio, k2p, bi, bd, vis, ops);
}
ORC_API int orc_search_projection_coarse(const snk_frame_view* f, const snk_point_coarse* pts, int n,
                                         const snk_camera* K, const snk_scale_pyramid* sp, float th, int fe,
                                         int motion, int32_t* k2p, long long* ops)
{
    return search_projection_coarse(f, pts, n, K, sp, th, fe, motion, k2p, ops);
}
ORC_API int orc_search_projection_frame_to_kf(const snk_frame_view* f, const snk_point_kf* pts, int n,
                                              const snk_camera* K, float th, int fe, int32_t* k2p, long long* ops)
{
    return search_projection_frame_to_kf(f, pts, n, K, th, fe, k2p, ops);
}
ORC_API int orc_match_bow(const uint64_t* d1, int n1, const uint64_t* d2, int n2, const snk_featvec_view* fv1,
                          const snk_featvec_view* fv2, const uint8_t* v1, const uint8_t* v2, int th, float ratio,
                          int32_t* m12, long long* ops)
{
    return match_bow(d1, n1, d2, n2, fv1, fv2, v1, v2, th, ratio, m12, ops);
}

// ---- MiniBow2 ----
ORC_API void* orc_voc_load_raw(const char* path)
{
    auto* v = new bowref::Vocabulary;
    if (!v->load_raw(path))
    {
        delete v;
        return nullptr;
    }
    return v;
}
ORC_API void* orc_voc_from_arrays(int k, int L, int n, const int32_t* parent, const double* weight, const int32_t* word_id,
                                  const uint64_t* desc)
{
    auto* v = new bowref::Vocabulary;
    if (!v->from_arrays(k, L, n, parent, weight, word_id, desc))
    {
        delete v;
        return nullptr;
    }
    return v;
}
ORC_API int orc_voc_save_raw(void* v, const char* path) { return ((bowref::Vocabulary*)v)->save_raw(path) ? 0 : 1; }
ORC_API void orc_voc_destroy(void* v) { delete (bowref::Vocabulary*)v; }
ORC_API void orc_voc_info(void* v, int* out4)
{
    auto* p = (bowref::Vocabulary*)v;
    out4[0] = p->k, out4[1] = p->L, out4[2] = (int)p->parent.size(), out4[3] = p->n_words;
}
ORC_API void orc_bow_transform(void* v, const uint64_t* desc, int n, int levelsup, int32_t* bow_ids, float* bow_vals,
                               int* n_bow, int32_t* fv_nodes, int32_t* fv_offsets, int32_t* fv_indices, int* n_fv)
{
    std::vector<int32_t> bi, fn, fo, fi;
    std::vector<float> bv;
    bowref::transform(*(bowref::Vocabulary*)v, desc, n, levelsup, bi, bv, fn, fo, fi);
    *n_bow = (int)bi.size(), *n_fv = (int)fn.size();
    std::copy(bi.begin(), bi.end(), bow_ids), std::copy(bv.begin(), bv.end(), bow_vals);
    std::copy(fn.begin(), fn.end(), fv_nodes), std::copy(fo.begin(), fo.end(), fv_offsets);
    std::copy(fi.begin(), fi.end(), fv_indices);
}
ORC_API float orc_bow_score(const int32_t* ids1, const float* v1, int n1, const int32_t* ids2, const float* v2, int n2)
{
    return bowref::score(ids1, v1, n1, ids2, v2, n2);
}
// ---- local mapping / loop closing searches ----
ORC_API double orc_epipolar_distance_squared(const double* p1, const double* p2, const double* F)
{
    return epipolar_distance_squared(p1, p2, F);
}
ORC_API int orc_fuse_search(const snk_frame_view* kf, const uint8_t* mask, const snk_point_fusion* pts, int n,
                            const snk_camera* K, const snk_scale_pyramid* sp, float th, float obs_factor, int feature_th,
                            int32_t* cand_kp, int32_t* cand_id, long long* ops)
{
    return fuse_search(kf, mask, pts, n, K, sp, th, obs_factor, feature_th, cand_kp, cand_id, ops);
}
ORC_API void orc_fuse_search_loop(const snk_frame_view* kf, const uint8_t* skip, const snk_point_fusion* pts, int n,
                                  const snk_camera* K, float th, int32_t* best_idx, int32_t* best_dist)
{
    fuse_search_loop(kf, skip, pts, n, K, th, best_idx, best_dist);
}
ORC_API int orc_search_triangulation_bow(const snk_frame_view* kf1, const double* np1, const snk_featvec_view* fv1,
                                         const snk_frame_view* kf2, const double* np2, const snk_featvec_view* fv2,
                                         const double* E, const snk_camera* K, float epi, int fd, int32_t* pairs,
                                         long long* ops)
{
    return search_triangulation_bow(kf1, np1, fv1, kf2, np2, fv2, E, K, epi, fd, pairs, ops);
}
ORC_API int orc_search_triangulation_project(const snk_frame_view* kf1, const double* np1, const snk_frame_view* kf2,
                                             const double* np2, const double* grid, int gr, int gc, const double* E,
                                             const snk_camera* K, float epi, int fd, int32_t* pairs, long long* ops)
{
    return search_triangulation_project(kf1, np1, kf2, np2, grid, gr, gc, E, K, epi, fd, pairs, ops);
}
ORC_API int orc_match_prediction(const uint64_t* dl, int n, double* pos, const snk_frame_view* right, float radius, int fe,
                                 int32_t* matches)
{
    return match_prediction(dl, n, pos, right, radius, fe, matches);
}
ORC_API int orc_features_in_area(const snk_frame_view* f, double px, double py, float r, int32_t* out)
{
    return features_in_area(f, px, py, r, out);
}
ORC_API void orc_grid_bounds(int w, int h, const double* K, const double* dist, int resolution, snk_grid_view* out)
{
    grid_bounds(w, h, K, dist, resolution, out);
}
}  // extern "C"
