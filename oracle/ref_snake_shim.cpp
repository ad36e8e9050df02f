// TEST INFRASTRUCTURE ONLY. extern "C" shim over the reference's own Snake matchers, compiled together with the
// reference's UNMODIFIED translation units where they lie under /root/reference (see Makefile: SNAKE_TUS) into
// oracle/_ref/libsnake_ref.so:
//   Snake/Preprocess/Preprocess.cpp            StereoMatching, ComputeStereoFromRGBD, undistortKeypoints
//   Snake/Tracking/SnakeORBMatcher.cpp         SearchByProjection2 / FrameFrame2 / FrameToKeyframe
//   Snake/LoopClosing/LoopORBMatcher.cpp       MatchBoW, Fuse
//   Snake/LocalMapping/MappingORBMatcher.cpp   Fuse, SearchForTriangulation2, SearchForTriangulationProject
//   Snake/Tracking/Initialization/MonoInitializer.cpp   ComputeMatchesPrediction
//   Snake/Map/{Features,Frame,Keyframe,KeyframeGraph,MapPoint}.cpp, saiga ScalePyramid / Epipolar / Rectify
// The Snake and Saiga HEADERS are the reference's own too (Map/Frame.h, Map/Keyframe.h, Map/MapPoint.h, Map/LocalMap.h,
// System/SnakeGlobal.h ...); only three absent third-party / generated headers have stand-ins in oracle/ref_stubs_snake
// (glog/logging.h, opencv2/imgproc.hpp — Preprocess.cpp includes it but the functions used here call nothing of it —
// and the cmake-generated buildconfig.h). Reference symbols outside the matcher path that these files mention (viewer,
// optimiser, IMU, RANSAC ...) are resolved by abort()-ing stubs generated at build time (gen_link_stubs.py).
//
// Every entry point takes the plain structs of include/snakeorb.h (the same arguments as the restatement's orc_*
// functions in oracle_capi.cpp), builds real Snake::Frame / Keyframe / MapPoint / LocalMap objects and the globals the
// functions read (K, stereo_cam, featureGridBounds, scalePyramid, settings, rect_left/right, rgbd_intrinsics), calls
// the reference function, and copies its effects back. Private members are reached with -fno-access-control.
// Not thread-safe (the reference's globals).
#include "LocalMapping/MappingORBMatcher.h"
#include "LoopClosing/LoopORBMatcher.h"
#include "Map/Map.h"
#include "Preprocess/Preprocess.h"
#include "Tracking/Initialization/MonoInitializer.h"
#include "Tracking/SnakeORBMatcher.h"

#include "../include/snakeorb.h"

// Built a second time with -DSNK_USE_ADAPTER (oracle/Makefile: _ref/libsnake_adapter.so): the same object builders, but
// the calls go to the drop-in classes of snake-slam_b200/host/snakeorb_snake_adapter.h (libsnakeorb.so, the GPU) instead
// of the reference's classes — the adapter is tested with real Snake objects against the reference's frozen outputs.
#ifdef SNK_USE_ADAPTER
#    define SNAKEORB_WITH_SAIGA
#    include "../snake-slam_b200/host/snakeorb_snake_adapter.h"
using ProjectionMatcherT = snkadapter::SnakeORBMatcherB200;
#else
using ProjectionMatcherT = Snake::SnakeORBMatcher;
#endif

#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <new>

using namespace Snake;
#define REF_API extern "C" __attribute__((visibility("default")))

// The global `inline Map map` (Map/Map.h:293) is constructed when the library is loaded; the reference's constructor
// (Map/Map.cpp:24) pre-builds 10 000 keyframes and 10 000 000 map points. Map.cpp is not linked (viewer / optimiser
// dependencies); the matchers never touch the map container, so an empty one stands in.
Snake::Map::Map(int, int) : allKeyframes(1), allPoints(1)
{
    pointValid.resize(1 << 21, true);  // MapPoint::Replace ends in map.EraseMapPoint(*this), which flips this flag
}

namespace
{
Saiga::Distortion make_distortion(const double* d)
{
    Eigen::Matrix<double, 8, 1> c;
    for (int i = 0; i < 8; ++i) c(i) = d[i];
    return Saiga::Distortion(c);
}

Saiga::Rectification make_rect(const snk_rectify* r, double bf)
{
    Saiga::Rectification o;
    o.K_src = IntrinsicsPinholed(r->fx, r->fy, r->cx, r->cy, 0);
    o.D_src = make_distortion(r->dist);
    o.R     = Quat(r->quat[3], r->quat[0], r->quat[1], r->quat[2]);  // Eigen ctor order: w, x, y, z
    o.K_dst = IntrinsicsPinholed(r->dfx, r->dfy, r->dcx, r->dcy, r->ds);
    o.bf    = bf;
    return o;
}

Snake::KeyPoint to_kp(const snk_keypoint_f64& k)
{
    Snake::KeyPoint o;
    o.point    = Vec2(k.x, k.y);
    o.size     = k.size;
    o.angle    = k.angle;
    o.response = k.response;
    o.octave   = k.octave;
    return o;
}

FeatureDescriptor to_desc(const uint64_t* d)
{
    FeatureDescriptor o;
    for (int i = 0; i < 4; ++i) o[i] = d[4 * 0 + i];
    return o;
}

void set_pyramid(int levels, double scale_factor) { scalePyramid = ScalePyramid(levels, scale_factor, 1000); }

void set_camera(const snk_camera* c)
{
    K          = IntrinsicsPinholed(c->fx, c->fy, c->cx, c->cy, c->s);
    stereo_cam = StereoCamera4(K, c->bf);
}

void set_bounds(const snk_grid_view& g)
{
    featureGridBounds.bmin        = Vec2(g.bmin_x, g.bmin_y);
    featureGridBounds.bmax        = Vec2(g.bmax_x, g.bmax_y);
    featureGridBounds.cellSizeInv = Vec2(g.cell_size_inv_x, g.cell_size_inv_y);
    featureGridBounds.cellSize    = Vec2(1.0 / g.cell_size_inv_x, 1.0 / g.cell_size_inv_y);
    featureGridBounds.Rows        = g.rows;
    featureGridBounds.Cols        = g.cols;
}

// MapPoint objects standing for the integer ids of the plain structs
struct PointPool
{
    std::map<int, std::unique_ptr<MapPoint>> by_id;
    MapPoint* get(int id)
    {
        if (id < 0) return nullptr;
        auto& p = by_id[id];
        if (!p) p = std::make_unique<MapPoint>(id);
        return p.get();
    }
};

// A Snake::Frame from a frame view: the members the searches read (Features.h:18-41, Frame.h)
std::shared_ptr<Frame> make_frame(const snk_frame_view* f, PointPool& pool, const double* normalized = nullptr)
{
    auto fr = std::make_shared<Frame>(0);
    fr->N   = f->n;
    fr->undistorted_keypoints.resize(f->n);
    fr->descriptors.resize(f->n);
    fr->right_points.assign(f->n, -1.f);
    fr->depth.assign(f->n, -1.f);
    fr->mvpMapPoints.assign(f->n, nullptr);
    fr->mvbOutlier.assign(f->n, 0);
    fr->normalized_points.resize(f->n);
    for (int i = 0; i < f->n; ++i)
    {
        fr->undistorted_keypoints[i] = to_kp(f->kp_un[i]);
        fr->descriptors[i]           = to_desc(f->desc + 4 * (size_t)i);
        if (f->right_points) fr->right_points[i] = f->right_points[i];
        if (f->depth) fr->depth[i] = f->depth[i];
        if (f->kp_to_point) fr->mvpMapPoints[i] = pool.get(f->kp_to_point[i]);
        if (normalized) fr->normalized_points[i] = Vec2(normalized[2 * i], normalized[2 * i + 1]);
    }
    fr->keypoints = fr->undistorted_keypoints;
    fr->tmpPose   = SE3(Quat(f->pose_q[3], f->pose_q[0], f->pose_q[1], f->pose_q[2]),
                        Vec3(f->pose_t[0], f->pose_t[1], f->pose_t[2]));
    fr->grid.Rows = f->grid.rows;
    fr->grid.Cols = f->grid.cols;
    fr->grid.grid.resize(f->grid.rows, f->grid.cols);
    for (int r = 0; r < f->grid.rows; ++r)
        for (int c = 0; c < f->grid.cols; ++c)
        {
            const int32_t* cell = f->grid.cells + 2 * ((size_t)r * f->grid.cols + c);
            fr->grid.grid(r, c) = {cell[0], cell[1]};
        }
    return fr;
}

void copy_assignment(const Frame& fr, int32_t* kp_to_point_out)
{
    for (int i = 0; i < fr.N; ++i) kp_to_point_out[i] = fr.mvpMapPoints[i] ? fr.mvpMapPoints[i]->id() : -1;
}

void fill_featvec(FeatureVector& fv, const snk_featvec_view* v)
{
    fv.clear();
    for (int k = 0; k < v->n_nodes; ++k)
    {
        std::vector<int> idx(v->indices + v->offsets[k], v->indices + v->offsets[k + 1]);
        fv.push_back({v->node_ids[k], idx});
    }
}
}  // namespace

// ---- Preprocess::StereoMatching (Preprocess.cpp:122-242) on RAW keypoints, with its own Rectification::Forward loops ----
REF_API int snkref_stereo_matching(const snk_rectify* rl, const snk_rectify* rr, double bf, int relaxed, int levels,
                                   double scale_factor, const snk_keypoint_f64* left, const uint64_t* dl, int nl,
                                   const snk_keypoint_f64* right, const uint64_t* dr, int nr, float* right_points,
                                   float* depth)
{
    settings.inputType         = InputType::Stereo;
    settings.fd_relaxed_stereo = relaxed != 0;
    rect_left                  = make_rect(rl, bf);
    rect_right                 = make_rect(rr, bf);
    set_pyramid(levels, scale_factor);
    Frame fr(0);
    fr.N = nl;
    for (int i = 0; i < nl; ++i) fr.keypoints.push_back(to_kp(left[i])), fr.descriptors.push_back(to_desc(dl + 4 * (size_t)i));
    for (int i = 0; i < nr; ++i)
        fr.keypoints_right.push_back(to_kp(right[i])), fr.descriptors_right.push_back(to_desc(dr + 4 * (size_t)i));
    fr.allocateTmp();  // right_points / depth = -1000 (Frame.cpp:21-29)
#ifdef SNK_USE_ADAPTER
    snkadapter::PreprocessB200 pp;
    int m = pp.StereoMatching(fr);
#else
    alignas(Preprocess) static unsigned char raw[sizeof(Preprocess)];  // the method reads no member; the ctor would start a thread
    int m = reinterpret_cast<Preprocess*>(raw)->StereoMatching(fr);
#endif
    for (int i = 0; i < nl; ++i) right_points[i] = fr.right_points[i], depth[i] = fr.depth[i];
    return m;
}

// ---- Preprocess::ComputeStereoFromRGBD (Preprocess.cpp:79-120) ----
REF_API int snkref_stereo_from_rgbd(const snk_rgbd_params* p, const snk_keypoint_f64* kp, int n, const float* img, int rows,
                                    int cols, size_t pitch, float* right_points, float* depth)
{
    settings.inputType            = InputType::RGBD;
    K                             = IntrinsicsPinholed(p->fx, p->fy, p->cx, p->cy, 0);
    rgbd_intrinsics.depthModel.K  = IntrinsicsPinholed(p->dfx, p->dfy, p->dcx, p->dcy, p->ds);
    rgbd_intrinsics.depthModel.dis = make_distortion(p->dist);
    rgbd_intrinsics.bf            = p->bf;
    Frame fr(0);
    fr.N = n;
    for (int i = 0; i < n; ++i) fr.undistorted_keypoints.push_back(to_kp(kp[i]));
    fr.right_points.assign(n, -1000.f);
    fr.depth.assign(n, -1000.f);
    fr.depth_image.create(rows, cols);
    for (int y = 0; y < rows; ++y)
        std::memcpy(&fr.depth_image(y, 0), reinterpret_cast<const char*>(img) + (size_t)y * pitch, (size_t)cols * 4);
#ifdef SNK_USE_ADAPTER
    snkadapter::PreprocessB200 pp;
    int m = pp.ComputeStereoFromRGBD(fr);
#else
    alignas(Preprocess) static unsigned char raw[sizeof(Preprocess)];
    int m = reinterpret_cast<Preprocess*>(raw)->ComputeStereoFromRGBD(fr);
#endif
    for (int i = 0; i < n; ++i) right_points[i] = fr.right_points[i], depth[i] = fr.depth[i];
    return m;
}

// ---- Preprocess::undistortKeypoints (Preprocess.cpp:55-77) ----
REF_API void snkref_undistort_keypoints(const snk_rectify* r, const snk_keypoint_f64* kps, int n, int img_rows, int img_cols,
                                        snk_keypoint_f64* kp_un, double* normalized)
{
    rect_left = make_rect(r, 1.0);
    Frame fr(0);
    fr.N = n;
    for (int i = 0; i < n; ++i) fr.keypoints.push_back(to_kp(kps[i]));
    fr.image.create(img_rows, img_cols);  // only its size is read (an assert)
    fr.allocateTmp();
#ifdef SNK_USE_ADAPTER
    snkadapter::PreprocessB200 pp;
    pp.undistortKeypoints(fr);
#else
    alignas(Preprocess) static unsigned char raw[sizeof(Preprocess)];
    reinterpret_cast<Preprocess*>(raw)->undistortKeypoints(fr);
#endif
    for (int i = 0; i < n; ++i)
    {
        kp_un[i]   = kps[i];
        kp_un[i].x = fr.undistorted_keypoints[i].point.x();
        kp_un[i].y = fr.undistorted_keypoints[i].point.y();
        normalized[2 * i] = fr.normalized_points[i].x(), normalized[2 * i + 1] = fr.normalized_points[i].y();
    }
}

// ---- SnakeORBMatcher::SearchByProjection2 (SnakeORBMatcher.cpp:365-526) ----
REF_API int snkref_search_projection_fine(const snk_frame_view* f, snk_point_fine* pts, int n, const snk_camera* cam,
                                          const snk_scale_pyramid* sp, float th, float ratio, int32_t* kp_to_point_out,
                                          int32_t* best_idx, int32_t* best_dist, uint8_t* visible, int threads)
{
    set_camera(cam);
    set_bounds(f->grid);
    set_pyramid(sp->nlevels, sp->scale_factor);
    PointPool pool;
    auto fr = make_frame(f, pool);
    PointPool lm_pool;  // the local map's points are distinct objects from the frame's earlier matches
    LocalMap<FineTrackingPoint> lm(16);
    std::vector<std::unique_ptr<MapPoint>> mps;
    for (int i = 0; i < n; ++i)
    {
        mps.push_back(std::make_unique<MapPoint>(pts[i].id));
        MapPoint* mp              = mps.back().get();
        mp->descriptor            = to_desc(pts[i].desc);
        mp->position              = Vec3(pts[i].pos[0], pts[i].pos[1], pts[i].pos[2]);
        mp->normal                = Vec3(pts[i].normal[0], pts[i].normal[1], pts[i].normal[2]);
        mp->reference_depth       = pts[i].ref_depth;
        mp->reference_scale_level = pts[i].ref_level;
        mp->mnVisible             = 0;
        FineTrackingPoint ftp(mp);  // copies descriptor, position, normal, depth, level (LocalMap.h:45-54)
        ftp.valid = pts[i].valid != 0;
        lm.points.push_back(ftp);
    }
    ProjectionMatcherT matcher;
    int m = matcher.SearchByProjection2(*fr, lm, th, ratio, threads);
    for (int i = 0; i < n; ++i)
    {
        pts[i].valid = lm.points[i].valid ? 1 : 0;
        visible[i]   = (uint8_t)mps[i]->mnVisible;
        best_idx[i]  = matcher.tmp_matches_world_image[i].first;
        best_dist[i] = best_idx[i] >= 0 ? matcher.tmp_matches_world_image[i].second : -1;
    }
    copy_assignment(*fr, kp_to_point_out);
    return m;
}

// ---- SnakeORBMatcher::SearchByProjectionFrameFrame2 (SnakeORBMatcher.cpp:191-354) ----
REF_API int snkref_search_projection_coarse(const snk_frame_view* f, const snk_point_coarse* pts, int n,
                                            const snk_camera* cam, const snk_scale_pyramid* sp, float th,
                                            int feature_error, int motion, int32_t* kp_to_point_out, int threads)
{
    set_camera(cam);
    set_bounds(f->grid);
    set_pyramid(sp->nlevels, sp->scale_factor);
    PointPool pool;
    auto fr = make_frame(f, pool);
    // bForward / bBackward are derived from the frame's velocity and the stereo baseline (:209-212)
    const double base = stereo_cam.baseLine();
    settings.inputType = motion == 0 ? InputType::Mono : InputType::Stereo;
    fr->local_velocity = SE3(Quat::Identity(), Vec3(0, 0, motion == 1 ? -(std::abs(base) * 2 + 1) : (std::abs(base) * 2 + 1)));
    LocalMap<CoarseTrackingPoint> lm(16);
    std::vector<std::unique_ptr<MapPoint>> mps;
    for (int i = 0; i < n; ++i)
    {
        mps.push_back(std::make_unique<MapPoint>(pts[i].id));
        MapPoint* mp = mps.back().get();
        mp->position = Vec3(pts[i].pos[0], pts[i].pos[1], pts[i].pos[2]);
        mp->normal   = Vec3(pts[i].normal[0], pts[i].normal[1], pts[i].normal[2]);
        CoarseTrackingPoint ctp(mp);
        ctp.descriptor = to_desc(pts[i].desc);
        ctp.octave     = pts[i].octave;
        ctp.angle      = pts[i].angle;
        lm.points.push_back(ctp);
    }
    ProjectionMatcherT matcher;
    int m = matcher.SearchByProjectionFrameFrame2(*fr, lm, th, feature_error, threads);
    copy_assignment(*fr, kp_to_point_out);
    return m;
}

// ---- SnakeORBMatcher::SearchByProjectionFrameToKeyframe (SnakeORBMatcher.cpp:71-188) ----
REF_API int snkref_search_projection_frame_to_kf(const snk_frame_view* f, const snk_point_kf* pts, int n,
                                                 const snk_camera* cam, float th, int feature_error,
                                                 int32_t* kp_to_point_out)
{
    set_camera(cam);
    set_bounds(f->grid);
    PointPool pool;  // shared: a keyframe point the frame already holds must be the same object (sAlreadyFound)
    auto fr = make_frame(f, pool);
    Keyframe kf(0);
    kf.frame = std::make_shared<Frame>(1);
    auto& obs = kf.MapPointRef();
    obs.assign(n, nullptr);
    for (int i = 0; i < n; ++i)
    {
        MapPoint* mp = pool.get(pts[i].id);
        if (!mp) continue;
        mp->descriptor = to_desc(pts[i].desc);
        mp->position   = Vec3(pts[i].pos[0], pts[i].pos[1], pts[i].pos[2]);
        obs[i]         = mp;
    }
    ProjectionMatcherT matcher;
    int m = matcher.SearchByProjectionFrameToKeyframe(*fr, kf, th, feature_error);
    copy_assignment(*fr, kp_to_point_out);
    return m;
}

// ---- LoopORBmatcher::MatchBoW (LoopORBMatcher.cpp:121-215) ----
REF_API int snkref_match_bow(const uint64_t* d1, int n1, const uint64_t* d2, int n2, const snk_featvec_view* fv1,
                             const snk_featvec_view* fv2, const uint8_t* valid1, const uint8_t* valid2, int th,
                             float ratio, int32_t* match12)
{
    // valid: 1 = good map point; 0 = none; 2 = a map point that is bad (isBad()) — both 0 and 2 are skipped by the reference
    Keyframe kf1(0), kf2(1);
    kf1.frame = std::make_shared<Frame>(0), kf2.frame = std::make_shared<Frame>(1);
    std::vector<std::unique_ptr<MapPoint>> own;
    auto fill = [&](Keyframe& kf, const uint64_t* d, int n, const snk_featvec_view* fv, const uint8_t* valid) {
        kf.frame->N = n;
        kf.frame->descriptors.resize(n);
        auto& obs = kf.MapPointRef();
        obs.assign(n, nullptr);
        for (int i = 0; i < n; ++i)
        {
            kf.frame->descriptors[i] = to_desc(d + 4 * (size_t)i);
            if (valid[i])
            {
                own.push_back(std::make_unique<MapPoint>(i));
                own.back()->valid = valid[i] == 1;
                obs[i]            = own.back().get();
            }
        }
        fill_featvec(kf.frame->bow_feature_vec, fv);
    };
    fill(kf1, d1, n1, fv1, valid1);
    fill(kf2, d2, n2, fv2, valid2);
    std::vector<MapPoint*> matches12;
#ifdef SNK_USE_ADAPTER
    snkadapter::LoopORBmatcherB200 lm;
    int m = lm.MatchBoW(&kf1, &kf2, matches12, th, ratio);
#else
    int m = LoopORBmatcher::MatchBoW(&kf1, &kf2, matches12, th, ratio);
#endif
    for (int i = 0; i < n1; ++i) match12[i] = matches12[i] ? matches12[i]->id() : -1;
    return m;
}

// ---- the candidate loop around MatchBoW (LoopDetector.cpp:384-391 -> :225), every candidate evaluated ----
REF_API int snkref_match_bow_batch(const uint64_t* d1, int n1, const snk_featvec_view* fv1, const uint8_t* valid1,
                                   const snk_bow_candidate* cands, int n_cand, int th, float ratio, int32_t* match12,
                                   int32_t* n_matches)
{
    std::vector<std::unique_ptr<MapPoint>> own;
    auto fill = [&](Keyframe& kf, int frame_id, const uint64_t* d, int n, const snk_featvec_view* fv, const uint8_t* valid) {
        kf.frame    = std::make_shared<Frame>(frame_id);
        kf.frame->N = n;
        kf.frame->descriptors.resize(n);
        auto& obs = kf.MapPointRef();
        obs.assign(n, nullptr);
        for (int i = 0; i < n; ++i)
        {
            kf.frame->descriptors[i] = to_desc(d + 4 * (size_t)i);
            if (valid[i])
            {
                own.push_back(std::make_unique<MapPoint>(i));
                own.back()->valid = valid[i] == 1;
                obs[i]            = own.back().get();
            }
        }
        fill_featvec(kf.frame->bow_feature_vec, fv);
    };
    Keyframe kf1(0);
    fill(kf1, 0, d1, n1, fv1, valid1);
    std::vector<std::unique_ptr<Keyframe>> kfs;
    std::vector<Keyframe*> ptrs;
    for (int k = 0; k < n_cand; ++k)
    {
        kfs.push_back(std::make_unique<Keyframe>(k + 1));
        fill(*kfs.back(), k + 1, cands[k].desc, cands[k].n, cands[k].fv, cands[k].valid);
        ptrs.push_back(kfs.back().get());
    }
    std::vector<std::vector<MapPoint*>> all(n_cand);
#ifdef SNK_USE_ADAPTER
    snkadapter::LoopORBmatcherB200 lm;
    const std::vector<int> counts = lm.MatchBoWBatch(&kf1, ptrs, all, th, ratio);
    for (int k = 0; k < n_cand; ++k) n_matches[k] = counts[k];
#else
    for (int k = 0; k < n_cand; ++k) n_matches[k] = LoopORBmatcher::MatchBoW(&kf1, ptrs[k], all[k], th, ratio);
#endif
    for (int k = 0; k < n_cand; ++k)
        for (int i = 0; i < n1; ++i) match12[(size_t)k * n1 + i] = all[k][i] ? all[k][i]->id() : -1;
    return 0;
}

// ---- MappingORBMatcher::Fuse(kf, pose, point_mask, LocalMap<FusionPoint>, ...) (MappingORBMatcher.cpp:359-482) ----
REF_API int snkref_fuse_search(const snk_frame_view* f, const uint8_t* mask, const snk_point_fusion* pts, int n,
                               const snk_camera* cam, const snk_scale_pyramid* sp, float th, float obs_factor,
                               int feature_th, int32_t* cand_kp, int32_t* cand_id)
{
    set_camera(cam);
    set_bounds(f->grid);
    set_pyramid(sp->nlevels, sp->scale_factor);
    PointPool pool;
    Keyframe kf(0);
    kf.frame = make_frame(f, pool);
    LocalMap<FusionPoint> lm(16);
    MapPoint proto(0);
    for (int i = 0; i < n; ++i)
    {
        FusionPoint fp(&proto);
        fp.descriptor            = to_desc(pts[i].desc);
        fp.position              = Vec3(pts[i].pos[0], pts[i].pos[1], pts[i].pos[2]);
        fp.normal                = Vec3(pts[i].normal[0], pts[i].normal[1], pts[i].normal[2]);
        fp.reference_depth       = pts[i].ref_depth;
        fp.reference_scale_level = pts[i].ref_level;
        fp.observations          = pts[i].observations;
        fp.id                    = pts[i].id;
        lm.points.push_back(fp);
    }
    std::vector<bool> vmask;
    if (mask) vmask.assign(mask, mask + n);
    std::vector<std::pair<int, int>> cands;
    MappingORBMatcher matcher;
    int m = matcher.Fuse(&kf, kf.frame->Pose(), mask ? &vmask : nullptr, lm, cands, th, obs_factor, feature_th);
    for (size_t j = 0; j < cands.size(); ++j) cand_kp[j] = cands[j].first, cand_id[j] = cands[j].second;
    return m;
}

// ---- MappingORBMatcher::SearchForTriangulation2 (MappingORBMatcher.cpp:14-101) ----
static void make_kf(Keyframe& kf, const snk_frame_view* f, const double* np, const snk_featvec_view* fv, PointPool& pool)
{
    kf.frame = make_frame(f, pool, np);
    kf.SetPose(kf.frame->Pose());
    auto& obs = kf.MapPointRef();
    obs.assign(f->n, nullptr);
    for (int i = 0; i < f->n; ++i) obs[i] = kf.frame->mvpMapPoints[i];
    if (fv) fill_featvec(kf.frame->bow_feature_vec, fv);
}

REF_API int snkref_search_triangulation_bow(const snk_frame_view* f1, const double* np1, const snk_featvec_view* fv1,
                                            const snk_frame_view* f2, const double* np2, const snk_featvec_view* fv2,
                                            const double* E, const snk_camera* cam, float epi, int fd, int32_t* pairs)
{
    set_camera(cam);
    PointPool p1, p2;
    Keyframe kf1(0), kf2(1);
    snk_frame_view a = *f1, b = *f2;  // only n, desc, kp_to_point are read: the views may come without grid / keypoints
    std::vector<snk_keypoint_f64> z1(f1->n), z2(f2->n);
    if (!a.kp_un) a.kp_un = z1.data();
    if (!b.kp_un) b.kp_un = z2.data();
    if (!a.grid.cells) a.grid.rows = a.grid.cols = 0;
    if (!b.grid.cells) b.grid.rows = b.grid.cols = 0;
    make_kf(kf1, &a, np1, fv1, p1);
    make_kf(kf2, &b, np2, fv2, p2);
    Mat3 Em;
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) Em(r, c) = E[3 * r + c];
    std::vector<std::pair<int, int>> out;
    MappingORBMatcher matcher;
    int m = matcher.SearchForTriangulation2(&kf1, &kf2, Em, out, epi, fd);
    for (size_t j = 0; j < out.size(); ++j) pairs[2 * j] = out[j].first, pairs[2 * j + 1] = out[j].second;
    return m;
}

// ---- MappingORBMatcher::SearchForTriangulationProject (MappingORBMatcher.cpp:167-247) ----
REF_API int snkref_search_triangulation_project(const snk_frame_view* f1, const double* np1, const snk_frame_view* f2,
                                                const double* np2, const double* grid, int gr, int gc, const double* E,
                                                const snk_camera* cam, float epi, int fd, int32_t* pairs)
{
    set_camera(cam);
    set_bounds(f1->grid);
    PointPool p1, p2;
    Keyframe kf1(0), kf2(1);
    make_kf(kf1, f1, np1, nullptr, p1);
    make_kf(kf2, f2, np2, nullptr, p2);
    Eigen::Matrix<double, -1, -1, Eigen::RowMajor> g(gr, gc);
    for (int r = 0; r < gr; ++r)
        for (int c = 0; c < gc; ++c) g(r, c) = grid[(size_t)r * gc + c];
    Mat3 Em;
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) Em(r, c) = E[3 * r + c];
    std::vector<std::pair<int, int>> out;
    MappingORBMatcher matcher;
    int m = matcher.SearchForTriangulationProject(g, kf1.Pose(), kf2.Pose(), &kf1, &kf2, Em, out, epi, fd);
    for (size_t j = 0; j < out.size(); ++j) pairs[2 * j] = out[j].first, pairs[2 * j + 1] = out[j].second;
    return m;
}

// ---- LoopORBmatcher::Fuse (LoopORBMatcher.cpp:13-96), search AND map edits ----
// skip[i]: 1 = the point is bad, 2 = the point is already in the keyframe (observed at feature `in_kf_idx` = the first
// free feature), 0 = neither. Reports what the reference did with every point: state[i] = feature index the point
// ended up observing in the keyframe (-1 none); the returned value is nFused.
REF_API int snkref_fuse_loop(const snk_frame_view* f, const uint8_t* skip, const snk_point_fusion* pts, int n,
                             const snk_camera* cam, float th, int32_t* state, uint8_t* kf_point_bad_out)
{
    set_camera(cam);
    set_bounds(f->grid);
    PointPool pool;
    Keyframe kf(0);
    make_kf(kf, f, nullptr, nullptr, pool);
    // the keyframe's own points observe it (needed by Replace / IsInKeyFrame)
    for (int i = 0; i < f->n; ++i)
        if (kf.GetMapPoint(i)) kf.GetMapPoint(i)->AddObservation(&kf, i);
    std::vector<std::unique_ptr<MapPoint>> own;
    std::vector<MapPoint*> vp;
    for (int i = 0; i < n; ++i)
    {
        own.push_back(std::make_unique<MapPoint>(1000000 + i));
        MapPoint* mp   = own.back().get();
        mp->descriptor = to_desc(pts[i].desc);
        mp->position   = Vec3(pts[i].pos[0], pts[i].pos[1], pts[i].pos[2]);
        mp->normal     = Vec3(pts[i].normal[0], pts[i].normal[1], pts[i].normal[2]);
        if (skip && skip[i] == 1) mp->valid = false;
        vp.push_back(mp);
    }
    LoopORBmatcher matcher;
    int m = matcher.Fuse(&kf, vp, th);
    for (int i = 0; i < n; ++i) state[i] = vp[i]->GetIndexInKeyFrame(&kf);
    for (int i = 0; i < f->n; ++i)
    {
        MapPoint* orig      = kf.frame->mvpMapPoints[i];
        kf_point_bad_out[i] = orig ? (orig->isBad() ? 1 : 0) : 2;
    }
    return m;
}

// ---- MonoInitializer::ComputeMatchesPrediction (MonoInitializer.cpp:619-697) ----
REF_API int snkref_match_prediction(const uint64_t* desc_left, int n_left, double* last_positions,
                                    const snk_frame_view* right, float radius, int feature_error, int threads,
                                    int32_t* matches)
{
    // radius (50) and featureError (50) are constants inside the reference function; the caller passes the same values
    if (radius != 50.f || feature_error != 50) return -2;
    set_bounds(right->grid);
    settings.num_tracking_threads = threads;
    PointPool pool;
    auto fr   = make_frame(right, pool);
    auto left = std::make_shared<Frame>(1);
    left->N   = n_left;
    left->descriptors.resize(n_left);
    left->undistorted_keypoints.resize(n_left);
    left->normalized_points.resize(n_left);
    for (int i = 0; i < n_left; ++i) left->descriptors[i] = to_desc(desc_left + 4 * (size_t)i);
    // the constructor sets up RANSAC solvers that are not linked; the method only uses plain containers, which are
    // valid when zero-filled (libstdc++ vectors)
    void* raw = std::calloc(1, sizeof(MonoInitializer));
    auto* mi  = reinterpret_cast<MonoInitializer*>(raw);
    new (&mi->last_frame_feature_position) decltype(mi->last_frame_feature_position)();
    mi->last_frame_feature_position.resize(n_left);
    for (int i = 0; i < n_left; ++i) mi->last_frame_feature_position[i] = Vec2(last_positions[2 * i], last_positions[2 * i + 1]);
    int m = mi->ComputeMatchesPrediction(left, fr);
    for (int i = 0; i < n_left; ++i)
    {
        matches[i]                = mi->matches[i];
        last_positions[2 * i]     = mi->last_frame_feature_position[i].x();
        last_positions[2 * i + 1] = mi->last_frame_feature_position[i].y();
    }
    return m;  // the object is leaked on purpose: its destructor would touch never-constructed solver members
}

// ---- Features::GetFeaturesInArea (Snake/Map/Features.cpp:13-30), the reference's own translation unit ----
REF_API int snkref_features_in_area(const snk_frame_view* f, double px, double py, float r, int32_t* out)
{
    set_bounds(f->grid);
    PointPool pool;
    auto fr = make_frame(f, pool);
    std::vector<int> idx;
    fr->GetFeaturesInArea(idx, Vec2(px, py), r);
    for (size_t i = 0; i < idx.size(); ++i) out[i] = idx[i];
    return (int)idx.size();
}
