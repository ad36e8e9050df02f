// Local-mapping and loop-closing searches: MappingORBMatcher::Fuse / SearchForTriangulation2 /
// SearchForTriangulationProject (Snake/LocalMapping/MappingORBMatcher.cpp), the search of LoopORBmatcher::Fuse
// (Snake/LoopClosing/LoopORBMatcher.cpp:13-72) and MonoInitializer::ComputeMatchesPrediction
// (Snake/Tracking/Initialization/MonoInitializer.cpp:619-697).
//
// In the reference every point / feature is processed independently (the "already matched" flags of the
// triangulation searches are never set), so one thread owns one point and walks its candidates in the reference's
// order; results are written per point and listed in order by the host wrapper. Geometry as in project_kernels.cuh:
// IEEE double, unfused, in the reference's operation order.
#pragma once
#include "project_kernels.cuh"

namespace snk
{
struct Mat3Dev
{
    double m[9];  // row-major
};

// Saiga::EpipolarDistanceSquared (saiga/vision/reconstruction/Epipolar.cpp:47-56). Eigen's fixed 3x3 * 3x1 product
// (SSE2 packets): rows 0 and 1 accumulate left to right, the scalar tail row sums as t0 + (t1 + t2).
__device__ __forceinline__ double epi_dist2(const double* __restrict__ p1, const double* __restrict__ p2, const Mat3Dev& F)
{
    const double x1 = p1[0], y1 = p1[1], x2 = p2[0], y2 = p2[1];
    const double l0 = (F.m[0] * x1 + F.m[1] * y1) + F.m[2] * 1.0;
    const double l1 = (F.m[3] * x1 + F.m[4] * y1) + F.m[5] * 1.0;
    const double l2 = F.m[6] * x1 + (F.m[7] * y1 + F.m[8] * 1.0);
    const double d       = (x2 * l0 + y2 * l1) + 1.0 * l2;
    const double lengSqr = l0 * l0 + l1 * l1;
    return d * d / lengSqr;
}

struct Projected
{
    D3 cam;  // point in the camera frame
    double u, v;
    bool ok;
};
// pose * wp, z > 0, K.project, featureGridBounds.inImage
__device__ __forceinline__ Projected project_point(const FrameDev& f, const snk_camera& K, const D3& wp)
{
    Projected p;
    const D3 r0 = qrot(f.q, wp);
    p.cam       = D3{r0.x + f.t[0], r0.y + f.t[1], r0.z + f.t[2]};
    p.ok        = false;
    p.u = p.v = 0;
    if (p.cam.z <= 0) return p;
    const double invz = 1.0 / p.cam.z;
    const double xn = p.cam.x * invz, yn = p.cam.y * invz;
    p.u  = K.fx * xn + K.s * yn + K.cx;
    p.v  = K.fy * yn + K.cy;
    p.ok = in_image(f, p.u, p.v);
    return p;
}

// MappingORBMatcher::Fuse, LocalMap<FusionPoint> overload (MappingORBMatcher.cpp:359-482)
__global__ void k_fuse(FrameDev f, const float* __restrict__ depth, const uint8_t* __restrict__ mask,
                       const snk_point_fusion* __restrict__ pts, int n, snk_camera K, snk_scale_pyramid sp, float th,
                       float obs_factor, int feature_th, double log_sf, int* __restrict__ best_idx)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    best_idx[i] = -1;
    if (mask && !mask[i]) return;
    const snk_point_fusion& lmp = pts[i];
    const D3 wp{lmp.pos[0], lmp.pos[1], lmp.pos[2]};
    const Projected pr = project_point(f, K, wp);
    if (!pr.ok) return;
    const D3 cam = cam_position(f);
    const D3 PO{cam.x - wp.x, cam.y - wp.y, cam.z - wp.z};
    const double dist = sqrt(PO.x * PO.x + PO.y * PO.y + PO.z * PO.z);
    double max_d = (double)lmp.ref_depth * sp.scale[lmp.ref_level];
    double min_d = max_d / sp.scale[sp.nlevels - 1];
    min_d /= sp.scale_factor;
    max_d *= sp.scale_factor;
    if (dist < min_d || dist > max_d) return;
    if (PO.x * lmp.normal[0] + PO.y * lmp.normal[1] + PO.z * lmp.normal[2] < 0.5 * dist) return;
    const float observationFactor = (lmp.observations <= 2) ? obs_factor : 1.0f;
    const float radius            = __fmul_rn(observationFactor, th);
    const double gate             = (double)__fmul_rn(__fmul_rn(th, th), observationFactor);
    const double pmax       = (double)lmp.ref_depth * sp.scale[lmp.ref_level];
    const double prediction = log(pmax / dist) / log_sf;
    const double r = (double)radius, r2 = r * r;
    // stereo_cam.projectStereo(np) (Intrinsics4.h:221-230)
    const double invz = 1.0 / pr.cam.z;
    const double sx = K.fx * (pr.cam.x * invz) + K.cx, sy = K.fy * (pr.cam.y * invz) + K.cy;
    const double sr = sx - K.bf * invz;
    const Desc q = load_desc(lmp.desc);
    int bestDist = 256, bestIdx = -1;
    for_cells(f, pr.u, pr.v, radius, [&](int pid) {
        const snk_keypoint_f64& kp = f.kp_un[pid];
        const double pe = prediction - kp.octave;
        if (!(pe * pe < (1.2 * 1.2))) return;
        const double dx = kp.x - pr.u, dy = kp.y - pr.v;
        if (!((dx * dx + dy * dy) < r2)) return;
        double e2;
        if (depth[pid] > 0.f)
        {
            const double d0 = sx - kp.x, d1 = sy - kp.y, d2 = sr - (double)f.right_points[pid];
            e2 = d0 * d0 + d1 * d1 + d2 * d2;
        }
        else
        {
            const double d0 = pr.u - kp.x, d1 = pr.v - kp.y;
            e2 = d0 * d0 + d1 * d1;
        }
        if (e2 > gate) return;
        const int d = hamming(q, load_desc(f.desc + 4 * (size_t)pid));
        if (d < bestDist) bestDist = d, bestIdx = pid;
    });
    if (bestIdx >= 0 && bestDist <= feature_th) best_idx[i] = bestIdx;
}

// LoopORBmatcher::Fuse, the search (LoopORBMatcher.cpp:13-72)
__global__ void k_fuse_loop(FrameDev f, const uint8_t* __restrict__ skip, const snk_point_fusion* __restrict__ pts,
                            int n, snk_camera K, float th, int* __restrict__ best_idx, int* __restrict__ best_dist)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    best_idx[i]  = -1;
    best_dist[i] = 256;
    if (skip && skip[i]) return;
    const snk_point_fusion& mp = pts[i];
    const D3 wp{mp.pos[0], mp.pos[1], mp.pos[2]};
    const Projected pr = project_point(f, K, wp);
    if (!pr.ok) return;
    const D3 cam = cam_position(f);
    const D3 PO{cam.x - wp.x, cam.y - wp.y, cam.z - wp.z};
    const double dist3D = sqrt(PO.x * PO.x + PO.y * PO.y + PO.z * PO.z);
    if (PO.x * mp.normal[0] + PO.y * mp.normal[1] + PO.z * mp.normal[2] < 0.5 * dist3D) return;
    const float radius = th;
    const double r2    = (double)__fmul_rn(radius, radius);  // GetFeaturesInArea keeps r*r in float (Features.cpp:17)
    const Desc q = load_desc(mp.desc);
    int bestDist = 256, bestIdx = -1;
    for_cells(f, pr.u, pr.v, radius, [&](int pid) {
        const snk_keypoint_f64& kp = f.kp_un[pid];
        const double dx = kp.x - pr.u, dy = kp.y - pr.v;
        if (!((dx * dx + dy * dy) < r2)) return;
        const int d = hamming(q, load_desc(f.desc + 4 * (size_t)pid));
        if (d < bestDist) bestDist = d, bestIdx = pid;
    });
    best_idx[i]  = bestIdx;
    best_dist[i] = bestDist;
}

// MappingORBMatcher::SearchForTriangulation2 (MappingORBMatcher.cpp:14-101). One thread per CSR slot of kf1's
// FeatureVector (the reference's visiting order of idx1); the merge over node ids is a binary search for the node in
// kf2's vector. best2[slot] = idx2 or -1.
__global__ void k_tri_bow(FeatVecDev fv1, FeatVecDev fv2, int n_slots, const int* __restrict__ k2p1,
                          const int* __restrict__ k2p2, const uint64_t* __restrict__ desc1,
                          const uint64_t* __restrict__ desc2, const double* __restrict__ np1,
                          const double* __restrict__ np2, Mat3Dev E, double th_chi2, int feature_distance,
                          int* __restrict__ best2)
{
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= n_slots) return;
    best2[slot] = -1;
    // node of this slot: last a with offsets[a] <= slot
    int lo = 0, hi = fv1.n_nodes;
    while (hi - lo > 1)
    {
        const int mid = (lo + hi) >> 1;
        if (fv1.offsets[mid] <= slot)
            lo = mid;
        else
            hi = mid;
    }
    const int node = fv1.node_ids[lo];
    int b0 = 0, b1 = fv2.n_nodes;
    while (b0 < b1)
    {
        const int mid = (b0 + b1) >> 1;
        if (fv2.node_ids[mid] < node)
            b0 = mid + 1;
        else
            b1 = mid;
    }
    if (b0 >= fv2.n_nodes || fv2.node_ids[b0] != node) return;
    const int idx1 = fv1.indices[slot];
    if (k2p1[idx1] >= 0) return;
    const Desc q = load_desc(desc1 + 4 * (size_t)idx1);
    int bestDist = 50, bestIdx2 = -1;  // TH_LOW; "dist > bestDist -> continue": the last minimum wins
    for (int ib = fv2.offsets[b0]; ib < fv2.offsets[b0 + 1]; ++ib)
    {
        const int idx2 = fv2.indices[ib];
        if (k2p2[idx2] >= 0) continue;
        const int dist = hamming(q, load_desc(desc2 + 4 * (size_t)idx2));
        if (dist > feature_distance || dist > bestDist) continue;
        if (epi_dist2(np1 + 2 * idx1, np2 + 2 * idx2, E) < th_chi2) bestIdx2 = idx2, bestDist = dist;
    }
    best2[slot] = bestIdx2;
}

// MappingORBMatcher::SearchForTriangulationProject (MappingORBMatcher.cpp:167-247). f1: kf1 (keypoints, pose, bounds),
// f2: kf2 (grid, keypoints, descriptors, pose). best2[idx1] = idx2 or -1.
__global__ void k_tri_project(FrameDev f1, FrameDev f2, const double* __restrict__ np1, const double* __restrict__ np2,
                              const double* __restrict__ grid, int grid_cols, Mat3Dev E, snk_camera K, double th_chi2,
                              int feature_distance, int* __restrict__ best2)
{
    const int idx1 = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx1 >= f1.n) return;
    best2[idx1] = -1;
    if (f1.kp_to_point[idx1] >= 0) return;
    const snk_keypoint_f64& kp1 = f1.kp_un[idx1];
    const int2 cell = cell_clamped(f1, kp1.x, kp1.y);
    const double z  = grid[(size_t)(cell.y / 4) * grid_cols + cell.x / 4];
    // K.unproject(ip, z) (Intrinsics4.h:70-74), then PoseInv() * p: Sophus inverse = (q^-1, q^-1 * (t * -1))
    const D3 p{((kp1.x - K.cx) / K.fx) * z, ((kp1.y - K.cy) / K.fy) * z, 1.0 * z};
    const double qi[4] = {-f1.q[0], -f1.q[1], -f1.q[2], f1.q[3]};
    const D3 ti = cam_position(f1);
    const D3 r0 = qrot(qi, p);
    const D3 wp{r0.x + ti.x, r0.y + ti.y, r0.z + ti.z};
    // K.project(pKF2->Pose() * wp) has no z test; inImage uses the global bounds
    const D3 r1 = qrot(f2.q, wp);
    const D3 c2{r1.x + f2.t[0], r1.y + f2.t[1], r1.z + f2.t[2]};
    const double invz = 1.0 / c2.z;
    const double xn = c2.x * invz, yn = c2.y * invz;
    const double u = K.fx * xn + K.s * yn + K.cx, v = K.fy * yn + K.cy;
    if (!in_image(f1, u, v)) return;
    const Desc q = load_desc(f1.desc + 4 * (size_t)idx1);
    int bestDist = 50, bestIdx2 = -1;
    for_cells(f2, u, v, 20.f, [&](int idx2) {
        const snk_keypoint_f64& kp = f2.kp_un[idx2];
        const double dx = kp.x - u, dy = kp.y - v;
        if (!((dx * dx + dy * dy) < 400.0)) return;
        if (f2.kp_to_point[idx2] >= 0) return;
        if (epi_dist2(np1 + 2 * idx1, np2 + 2 * idx2, E) > th_chi2) return;
        const int dist = hamming(q, load_desc(f2.desc + 4 * (size_t)idx2));
        if (dist > feature_distance || dist > bestDist) return;
        bestIdx2 = idx2, bestDist = dist;
    });
    best2[idx1] = bestIdx2;
}

// MonoInitializer::ComputeMatchesPrediction (MonoInitializer.cpp:619-697)
__global__ void k_match_prediction(FrameDev right, const uint64_t* __restrict__ desc_left, int n_left,
                                   double* __restrict__ positions, float radius, int feature_error,
                                   int* __restrict__ matches)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_left) return;
    const double px = positions[2 * i], py = positions[2 * i + 1];
    const double r2 = (double)__fmul_rn(radius, radius);
    const Desc q = load_desc(desc_left + 4 * (size_t)i);
    int bestDist = 256, bestIdx2 = -1;
    for_cells(right, px, py, radius, [&](int i2) {
        const snk_keypoint_f64& kp = right.kp_un[i2];
        const double dx = kp.x - px, dy = kp.y - py;
        if (!((dx * dx + dy * dy) < r2)) return;
        const int d = hamming(q, load_desc(right.desc + 4 * (size_t)i2));
        if (d < bestDist) bestDist = d, bestIdx2 = i2;
    });
    if (bestDist < feature_error)
    {
        positions[2 * i]     = right.kp_un[bestIdx2].x;
        positions[2 * i + 1] = right.kp_un[bestIdx2].y;
        matches[i]           = bestIdx2;
    }
    else
        matches[i] = -1;
}
}  // namespace snk
