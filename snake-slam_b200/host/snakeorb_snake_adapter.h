// snakeorb_snake_adapter.h — the Snake-SLAM matcher classes on top of libsnakeorb.so, taking the reference's own structs.
//
// To be compiled INSIDE the reference tree (it includes Snake's and Saiga's headers): the classes below have the
// member functions of Snake::SnakeORBMatcher (Snake/Tracking/SnakeORBMatcher.h:15-53), LoopORBmatcher::MatchBoW
// (Snake/LoopClosing/LoopORBMatcher.h:21-23) and the matching half of Snake::Preprocess (Snake/Preprocess/Preprocess.h:
// 31-34) with the same signatures, read the same globals (K, stereo_cam, featureGridBounds, scalePyramid, settings,
// rect_left / rect_right, rgbd_intrinsics) and leave the same effects in Frame / LocalMap / MapPoint. The only
// translation at the boundary is MapPoint* <-> integer id (the C ABI carries no pointers into the caller's map).
// A call site changes one type name: `SnakeORBMatcher matcher;` -> `snkadapter::SnakeORBMatcherB200 matcher;`.
//
// Tested by tests/test_gpu_adapter_snake.py: oracle/ref_snake_shim.cpp built a second time with SNK_USE_ADAPTER drives
// these classes with real Snake objects on the GPU and must reproduce the outputs frozen from the reference's own
// compiled matchers (tests/golden/golden_snake.npz).
#pragma once
#include "snakeorb_saiga_adapter.h"

#include "LoopClosing/LoopORBMatcher.h"
#include "Map/Frame.h"
#include "Map/LocalMap.h"
#include "Map/Map.h"
#include "Preprocess/Preprocess.h"
#include "Tracking/SnakeORBMatcher.h"

#include <unordered_map>

namespace snkadapter
{
static_assert(sizeof(Snake::KeyPoint) == sizeof(snk_keypoint_f64), "Snake::KeyPoint must be the 48-byte record of the C ABI");
static_assert(sizeof(std::pair<int, int>) == 8, "FeatureGrid2 cells must be (begin, end) int pairs");

// MapPoint* <-> id for one call
struct PointTable
{
    std::vector<Snake::MapPoint*> points;
    std::unordered_map<Snake::MapPoint*, int> index;
    int id(Snake::MapPoint* mp)
    {
        if (!mp) return -1;
        auto it = index.find(mp);
        if (it != index.end()) return it->second;
        index.emplace(mp, (int)points.size());
        points.push_back(mp);
        return (int)points.size() - 1;
    }
    Snake::MapPoint* ptr(int id) const { return id < 0 ? nullptr : points[id]; }
};

inline snk_camera camera_b200()
{
    return snk_camera{Snake::K.fx, Snake::K.fy, Snake::K.cx, Snake::K.cy, Snake::K.s, Snake::stereo_cam.bf};
}
inline snk_scale_pyramid pyramid_b200()
{
    snk_scale_pyramid p{};
    p.nlevels      = Snake::scalePyramid.num_levels;
    p.scale_factor = Snake::scalePyramid.scale_factor;
    for (int i = 0; i < p.nlevels && i < 16; ++i) p.scale[i] = Snake::scalePyramid.Scale(i);
    return p;
}
inline snk_rectify rectify_b200(const Saiga::Rectification& r)
{
    snk_rectify o{};
    o.fx = r.K_src.fx, o.fy = r.K_src.fy, o.cx = r.K_src.cx, o.cy = r.K_src.cy;
    const auto d = r.D_src.Coeffs();
    for (int i = 0; i < 8; ++i) o.dist[i] = d(i);
    o.quat[0] = r.R.x(), o.quat[1] = r.R.y(), o.quat[2] = r.R.z(), o.quat[3] = r.R.w();
    o.dfx = r.K_dst.fx, o.dfy = r.K_dst.fy, o.dcx = r.K_dst.cx, o.dcy = r.K_dst.cy, o.ds = r.K_dst.s;
    return o;
}

// The parts of a Frame the searches read, as the C ABI's view (no copies except the MapPoint* -> id column)
struct FrameViewB200
{
    snk_frame_view v{};
    std::vector<int32_t> k2p;
    FrameViewB200(Snake::Frame& f, const Saiga::SE3& pose, const std::vector<Snake::MapPoint*>* points, PointTable& table)
    {
        v.n            = f.N;
        v.kp_un        = reinterpret_cast<const snk_keypoint_f64*>(f.undistorted_keypoints.data());
        v.desc         = reinterpret_cast<const uint64_t*>(f.descriptors.data());
        v.right_points = f.right_points.empty() ? nullptr : f.right_points.data();
        v.depth        = f.depth.empty() ? nullptr : f.depth.data();
        k2p.assign(std::max(f.N, 1), -1);
        if (points)
            for (int i = 0; i < f.N && i < (int)points->size(); ++i) k2p[i] = table.id((*points)[i]);
        v.kp_to_point = k2p.data();
        const auto q  = pose.unit_quaternion();
        v.pose_q[0] = q.x(), v.pose_q[1] = q.y(), v.pose_q[2] = q.z(), v.pose_q[3] = q.w();
        for (int i = 0; i < 3; ++i) v.pose_t[i] = pose.translation()(i);
        auto& b                = Snake::featureGridBounds;
        v.grid.bmin_x          = b.bmin(0), v.grid.bmin_y = b.bmin(1), v.grid.bmax_x = b.bmax(0), v.grid.bmax_y = b.bmax(1);
        v.grid.cell_size_inv_x = b.cellSizeInv(0), v.grid.cell_size_inv_y = b.cellSizeInv(1);
        v.grid.rows = b.Rows, v.grid.cols = b.Cols;
        v.grid.cells = reinterpret_cast<const int32_t*>(f.grid.grid.data());  // row-major (begin, end) pairs
    }
};

template <typename P>
inline void copy_point_common(P& o, const Snake::FeatureDescriptor& d, const Snake::Vec3& pos)
{
    for (int k = 0; k < 4; ++k) o.desc[k] = d[k];
    for (int k = 0; k < 3; ++k) o.pos[k] = pos(k);
}

// Snake::SnakeORBMatcher (Snake/Tracking/SnakeORBMatcher.h:15-53, .cpp:71-526)
class SnakeORBMatcherB200
{
   public:
    explicit SnakeORBMatcherB200(int device = 0) { check(snk_ctx_create(device, &c_), "snk_ctx_create"); }
    ~SnakeORBMatcherB200() { snk_ctx_destroy(c_); }
    SnakeORBMatcherB200(const SnakeORBMatcherB200&)            = delete;
    SnakeORBMatcherB200& operator=(const SnakeORBMatcherB200&) = delete;

    // SnakeORBMatcher.cpp:71-188
    int SearchByProjectionFrameToKeyframe(Snake::Frame& CurrentFrame, const Snake::Keyframe& kf, float th,
                                          Snake::FeatureDistance featureError)
    {
        PointTable table;
        FrameViewB200 fv(CurrentFrame, CurrentFrame.Pose(), &CurrentFrame.mvpMapPoints, table);
        auto& kfp = kf.GetMapPointMatches();
        std::vector<snk_point_kf> pts(kfp.size());
        for (size_t i = 0; i < kfp.size(); ++i)
        {
            pts[i]    = snk_point_kf{};
            pts[i].id = table.id(kfp[i]);
            if (kfp[i]) copy_point_common(pts[i], kfp[i]->GetDescriptor(), kfp[i]->getPosition());
        }
        const snk_camera cam = camera_b200();
        std::vector<int32_t> out(std::max(CurrentFrame.N, 1));
        int matches = 0;
        check(snk_search_projection_frame_to_kf(c_, &fv.v, pts.data(), (int)pts.size(), &cam, th, featureError, out.data(), &matches),
              "snk_search_projection_frame_to_kf");
        for (int i = 0; i < CurrentFrame.N; ++i) CurrentFrame.mvpMapPoints[i] = table.ptr(out[i]);
        return matches;
    }

    // SnakeORBMatcher.cpp:365-526; num_threads is accepted and ignored
    int SearchByProjection2(Snake::Frame& CurrentFrame, Snake::LocalMap<Snake::FineTrackingPoint>& lm, const float th, float ratio,
                            int /*num_threads*/)
    {
        PointTable table;
        FrameViewB200 fv(CurrentFrame, CurrentFrame.Pose(), &CurrentFrame.mvpMapPoints, table);
        const int N = (int)lm.points.size();
        std::vector<snk_point_fine> pts(N);
        for (int i = 0; i < N; ++i)
        {
            auto& lmp = lm.points[i];
            pts[i]    = snk_point_fine{};
            copy_point_common(pts[i], lmp.descriptor, lmp.position);
            for (int k = 0; k < 3; ++k) pts[i].normal[k] = lmp.normal(k);
            pts[i].ref_depth = lmp.reference_depth;
            pts[i].ref_level = lmp.reference_scale_level;
            pts[i].valid     = lmp.valid ? 1 : 0;
            pts[i].id        = table.id(lmp.mp);
        }
        const snk_camera cam        = camera_b200();
        const snk_scale_pyramid pyr = pyramid_b200();
        std::vector<int32_t> out(std::max(CurrentFrame.N, 1)), bi(std::max(N, 1)), bd(std::max(N, 1));
        std::vector<uint8_t> vis(std::max(N, 1));
        int matches = 0;
        check(snk_search_projection_fine(c_, &fv.v, pts.data(), N, &cam, &pyr, th, ratio, out.data(), bi.data(), bd.data(), vis.data(),
                                         &matches),
              "snk_search_projection_fine");
        tmp_matches_world_image.resize(N);
        for (int i = 0; i < N; ++i)
        {
            lm.points[i].valid = pts[i].valid != 0;
            if (vis[i]) lm.points[i].mp->IncreaseVisible();
            tmp_matches_world_image[i] = {bi[i], bd[i]};
        }
        for (int i = 0; i < CurrentFrame.N; ++i) CurrentFrame.mvpMapPoints[i] = table.ptr(out[i]);
        return matches;
    }

    // SnakeORBMatcher.cpp:191-354
    int SearchByProjectionFrameFrame2(Snake::Frame& CurrentFrame, const Snake::LocalMap<Snake::CoarseTrackingPoint>& lm, const float th,
                                      Snake::FeatureDistance featureError, int /*num_threads*/)
    {
        PointTable table;
        FrameViewB200 fv(CurrentFrame, CurrentFrame.Pose(), &CurrentFrame.mvpMapPoints, table);
        const int N = (int)lm.points.size();
        std::vector<snk_point_coarse> pts(N);
        for (int i = 0; i < N; ++i)
        {
            auto& lmp = lm.points[i];
            pts[i]    = snk_point_coarse{};
            copy_point_common(pts[i], lmp.descriptor, lmp.position);
            for (int k = 0; k < 3; ++k) pts[i].normal[k] = lmp.normal(k);
            pts[i].octave = lmp.octave;
            pts[i].angle  = lmp.angle;
            pts[i].id     = table.id(lmp.mp);
        }
        // :209-212 — negative because the motion model lives in world -> camera space
        auto z_diff          = -CurrentFrame.local_velocity.translation().z();
        const bool bForward  = Snake::settings.inputType != Snake::InputType::Mono && z_diff > Snake::stereo_cam.baseLine();
        const bool bBackward = Snake::settings.inputType != Snake::InputType::Mono && z_diff < Snake::stereo_cam.baseLine();
        const snk_camera cam        = camera_b200();
        const snk_scale_pyramid pyr = pyramid_b200();
        std::vector<int32_t> out(std::max(CurrentFrame.N, 1));
        int matches = 0;
        check(snk_search_projection_coarse(c_, &fv.v, pts.data(), N, &cam, &pyr, th, featureError, bForward ? 1 : (bBackward ? 2 : 0),
                                           out.data(), &matches),
              "snk_search_projection_coarse");
        for (int i = 0; i < CurrentFrame.N; ++i) CurrentFrame.mvpMapPoints[i] = table.ptr(out[i]);
        return matches;
    }

    std::vector<std::pair<int, int>> tmp_matches_world_image;  // (best keypoint, distance) per local-map point, as in the reference
    snk_ctx* context() { return c_; }

   private:
    snk_ctx* c_ = nullptr;
};

// LoopORBmatcher::MatchBoW (Snake/LoopClosing/LoopORBMatcher.cpp:121-215)
class LoopORBmatcherB200
{
   public:
    explicit LoopORBmatcherB200(int device = 0) { check(snk_ctx_create(device, &c_), "snk_ctx_create"); }
    ~LoopORBmatcherB200() { snk_ctx_destroy(c_); }
    LoopORBmatcherB200(const LoopORBmatcherB200&)            = delete;
    LoopORBmatcherB200& operator=(const LoopORBmatcherB200&) = delete;

    int MatchBoW(Snake::KeyFrame* pKF1, Snake::KeyFrame* pKF2, std::vector<Snake::MapPoint*>& vpMatches12, int threshold, float ratio)
    {
        auto& mp1 = pKF1->GetMapPointMatches();
        auto& mp2 = pKF2->GetMapPointMatches();
        Csr fv1(pKF1->frame->bow_feature_vec), fv2(pKF2->frame->bow_feature_vec);
        const auto v1 = valid(mp1), v2 = valid(mp2);
        std::vector<int32_t> m12(std::max<size_t>(mp1.size(), 1), -1);
        int n = 0;
        check(snk_match_bow(c_, reinterpret_cast<const uint64_t*>(pKF1->frame->descriptors.data()), (int)mp1.size(),
                            reinterpret_cast<const uint64_t*>(pKF2->frame->descriptors.data()), (int)mp2.size(), &fv1.view, &fv2.view,
                            v1.data(), v2.data(), threshold, ratio, m12.data(), &n),
              "snk_match_bow");
        vpMatches12.assign(mp1.size(), nullptr);
        for (size_t i = 0; i < mp1.size(); ++i)
            if (m12[i] >= 0) vpMatches12[i] = mp2[m12[i]];
        return n;
    }

    // MatchBoW(pKF1, candidates[k], vpMatches12[k], ...) for every k in one device call: what the candidate loop of
    // LoopDetector::ComputeSim3(new_kf) (LoopDetector.cpp:384-391) asks for, ahead of its early exit. Returns the match counts.
    std::vector<int> MatchBoWBatch(Snake::KeyFrame* pKF1, const std::vector<Snake::KeyFrame*>& candidates,
                                   std::vector<std::vector<Snake::MapPoint*>>& vpMatches12, int threshold, float ratio)
    {
        auto& mp1 = pKF1->GetMapPointMatches();
        Csr fv1(pKF1->frame->bow_feature_vec);
        const auto v1 = valid(mp1);
        const size_t K = candidates.size(), n1 = mp1.size();
        std::vector<Csr> fvs;
        std::vector<std::vector<uint8_t>> vs;
        fvs.reserve(K), vs.reserve(K);  // the views point into the elements: no reallocation after this
        std::vector<snk_bow_candidate> cs(K);
        for (size_t k = 0; k < K; ++k)
        {
            auto& mp2 = candidates[k]->GetMapPointMatches();
            fvs.emplace_back(candidates[k]->frame->bow_feature_vec);
            fvs.back().rebind();
            vs.push_back(valid(mp2));
            cs[k] = snk_bow_candidate{reinterpret_cast<const uint64_t*>(candidates[k]->frame->descriptors.data()), (int32_t)mp2.size(),
                                      &fvs[k].view, vs[k].data()};
        }
        std::vector<int32_t> m12(std::max<size_t>(K * n1, 1), -1);
        std::vector<int> counts(std::max<size_t>(K, 1), 0);
        check(snk_match_bow_batch(c_, reinterpret_cast<const uint64_t*>(pKF1->frame->descriptors.data()), (int)n1, &fv1.view, v1.data(),
                                  cs.data(), (int)K, threshold, ratio, m12.data(), counts.data()),
              "snk_match_bow_batch");
        vpMatches12.assign(K, std::vector<Snake::MapPoint*>(n1, nullptr));
        for (size_t k = 0; k < K; ++k)
        {
            auto& mp2 = candidates[k]->GetMapPointMatches();
            for (size_t i = 0; i < n1; ++i)
                if (m12[k * n1 + i] >= 0) vpMatches12[k][i] = mp2[m12[k * n1 + i]];
        }
        counts.resize(K);
        return counts;
    }

   private:
    struct Csr  // FeatureVector (MiniBow2.h:98-132) as the CSR arrays of snk_featvec_view
    {
        std::vector<int32_t> ids, off, idx;
        snk_featvec_view view{};
        explicit Csr(const Snake::FeatureVector& fv)
        {
            off.push_back(0);
            for (auto& node : fv)
            {
                ids.push_back((int32_t)node.first);
                for (int i : node.second) idx.push_back(i);
                off.push_back((int32_t)idx.size());
            }
            rebind();
        }
        void rebind() { view = snk_featvec_view{(int32_t)ids.size(), ids.data(), off.data(), idx.data()}; }
    };
    static std::vector<uint8_t> valid(const std::vector<Snake::MapPoint*>& mps)
    {
        std::vector<uint8_t> v(std::max<size_t>(mps.size(), 1), 0);
        for (size_t i = 0; i < mps.size(); ++i) v[i] = (mps[i] && !mps[i]->isBad()) ? 1 : 0;
        return v;
    }
   private:
    snk_ctx* c_ = nullptr;
};

// The matching half of Snake::Preprocess (Snake/Preprocess/Preprocess.cpp:55-242)
class PreprocessB200
{
   public:
    explicit PreprocessB200(int device = 0) { check(snk_ctx_create(device, &c_), "snk_ctx_create"); }
    ~PreprocessB200() { snk_ctx_destroy(c_); }
    PreprocessB200(const PreprocessB200&)            = delete;
    PreprocessB200& operator=(const PreprocessB200&) = delete;

    // Preprocess.cpp:55-77
    void undistortKeypoints(Snake::Frame& frame)
    {
        const snk_rectify r = rectify_b200(Snake::rect_left);
        frame.undistorted_keypoints.resize(frame.N);
        frame.normalized_points.resize(frame.N);
        std::vector<double> nrm(2 * std::max(frame.N, 1));
        check(snk_undistort_keypoints(c_, &r, reinterpret_cast<const snk_keypoint_f64*>(frame.keypoints.data()), frame.N,
                                      reinterpret_cast<snk_keypoint_f64*>(frame.undistorted_keypoints.data()), nrm.data()),
              "snk_undistort_keypoints");
        for (int i = 0; i < frame.N; ++i) frame.normalized_points[i] = Snake::Vec2(nrm[2 * i], nrm[2 * i + 1]);
    }

    // Preprocess.cpp:122-242 (the Rectification::Forward loops included)
    int StereoMatching(Snake::Frame& frame)
    {
        const snk_rectify rl = rectify_b200(Snake::rect_left), rr = rectify_b200(Snake::rect_right);
        snk_stereo_params p{};
        p.bf      = Snake::rect_left.bf;
        p.relaxed = Snake::settings.fd_relaxed_stereo ? 1 : 0;
        p.nlevels = Snake::scalePyramid.num_levels;
        for (int i = 0; i < p.nlevels && i < 16; ++i) p.scale[i] = Snake::scalePyramid.Scale(i);
        const int nl = (int)frame.keypoints.size(), nr = (int)frame.keypoints_right.size();
        frame.right_points.resize(nl);
        frame.depth.resize(nl);
        int n = 0;
        check(snk_stereo_matching(c_, &rl, &rr, &p, reinterpret_cast<const snk_keypoint_f64*>(frame.keypoints.data()),
                                  reinterpret_cast<const uint64_t*>(frame.descriptors.data()), nl,
                                  reinterpret_cast<const snk_keypoint_f64*>(frame.keypoints_right.data()),
                                  reinterpret_cast<const uint64_t*>(frame.descriptors_right.data()), nr, frame.right_points.data(),
                                  frame.depth.data(), &n),
              "snk_stereo_matching");
        return n;
    }

    // Preprocess.cpp:79-120
    int ComputeStereoFromRGBD(Snake::Frame& frame)
    {
        auto& ri = Snake::rgbd_intrinsics;
        snk_rgbd_params p{};
        p.fx = Snake::K.fx, p.fy = Snake::K.fy, p.cx = Snake::K.cx, p.cy = Snake::K.cy;
        p.dfx = ri.depthModel.K.fx, p.dfy = ri.depthModel.K.fy, p.dcx = ri.depthModel.K.cx, p.dcy = ri.depthModel.K.cy, p.ds = ri.depthModel.K.s;
        const auto d = ri.depthModel.dis.Coeffs();
        for (int i = 0; i < 8; ++i) p.dist[i] = d(i);
        p.bf = ri.bf;
        frame.right_points.resize(frame.N);
        frame.depth.resize(frame.N);
        int n = 0;
        check(snk_stereo_from_rgbd(c_, &p, reinterpret_cast<const snk_keypoint_f64*>(frame.undistorted_keypoints.data()), frame.N,
                                   frame.depth_image.data(), frame.depth_image.rows, frame.depth_image.cols, frame.depth_image.pitchBytes,
                                   frame.right_points.data(), frame.depth.data(), &n),
              "snk_stereo_from_rgbd");
        return n;
    }

   private:
    snk_ctx* c_ = nullptr;
};
}  // namespace snkadapter
