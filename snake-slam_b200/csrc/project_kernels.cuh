// Projection searches (SnakeORBMatcher, Snake/Tracking/SnakeORBMatcher.cpp) and BoW-candidate matching
// (LoopORBmatcher::MatchBoW, Snake/LoopClosing/LoopORBMatcher.cpp:121-215).
//
// Geometry is double precision and unfused (the library is compiled with -fmad=false), in the reference's operation
// order: Sophus SE3 * point through Eigen's quaternion rotation, pinhole projection with one reciprocal
// (saiga/vision/cameraModel/Intrinsics4.h:53-59), grid cells by iFloor((p - bmin) * cellSizeInv) clamped
// (saiga/vision/util/FeatureGrid2.h:24-60). Candidates are visited cx-outer, cy-inner, index-ascending
// (Snake/Map/Features.cpp:13-76), one map point per thread, so strict '<' keeps the reference's first minimum.
// The serial "first point wins the keypoint" pass (:321-332, :516-524) becomes an atomicMin on the point index.
#pragma once
#include "match_kernels.cuh"

namespace snk
{
struct FrameDev
{
    int n;
    const snk_keypoint_f64* kp_un;
    const uint64_t* desc;
    const float* right_points;
    const int* kp_to_point;
    double q[4], t[3];
    double bmin_x, bmin_y, bmax_x, bmax_y, inv_x, inv_y;
    int rows, cols;
    const int* cells;
};

struct D3
{
    double x, y, z;
};
__device__ __forceinline__ D3 cross3(const D3& a, const D3& b)
{
    return D3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
// Eigen::QuaternionBase::_transformVector: uv = 2 * (q.vec x v);  v + w*uv + q.vec x uv
__device__ __forceinline__ D3 qrot(const double* q, const D3& v)
{
    const D3 qv{q[0], q[1], q[2]};
    D3 uv = cross3(qv, v);
    uv    = D3{uv.x + uv.x, uv.y + uv.y, uv.z + uv.z};
    const D3 c = cross3(qv, uv);
    return D3{v.x + q[3] * uv.x + c.x, v.y + q[3] * uv.y + c.y, v.z + q[3] * uv.z + c.z};
}
__device__ __forceinline__ D3 cam_position(const FrameDev& f)
{
    const double qi[4] = {-f.q[0], -f.q[1], -f.q[2], f.q[3]};
    return qrot(qi, D3{f.t[0] * -1, f.t[1] * -1, f.t[2] * -1});
}
__device__ __forceinline__ int ifloor_d(double v) { return ((int)v) - (((int)v) > v); }
__device__ __forceinline__ int2 cell_clamped(const FrameDev& f, double px, double py)
{
    int cx = ifloor_d((px - f.bmin_x) * f.inv_x), cy = ifloor_d((py - f.bmin_y) * f.inv_y);
    return make_int2(min(max(cx, 0), f.cols - 1), min(max(cy, 0), f.rows - 1));
}
__device__ __forceinline__ bool in_image(const FrameDev& f, double x, double y)
{
    return x >= f.bmin_x && y >= f.bmin_y && x < f.bmax_x && y < f.bmax_y;
}

// F(pid) is called for every keypoint of the visited cells, in the reference's order
template <typename F>
__device__ __forceinline__ void for_cells(const FrameDev& f, double u, double v, float r_cells, F&& fn)
{
    const int2 lo = cell_clamped(f, u - r_cells, v - r_cells), hi = cell_clamped(f, u + r_cells, v + r_cells);
    for (int cx = lo.x; cx <= hi.x; ++cx)
        for (int cy = lo.y; cy <= hi.y; ++cy)
        {
            const int2 be = reinterpret_cast<const int2*>(f.cells)[cy * f.cols + cx];
            for (int pid = be.x; pid < be.y; ++pid) fn(pid);
        }
}

// SnakeORBMatcher::SearchByProjection2, parallel part (:380-514)
__global__ void k_proj_fine(FrameDev f, snk_point_fine* __restrict__ pts, int n, snk_camera K, snk_scale_pyramid sp,
                            float th, float ratio, double log_sf, int* __restrict__ best_idx,
                            int* __restrict__ best_dist, uint8_t* __restrict__ visible, int* __restrict__ winner)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    best_idx[i]  = -1;
    best_dist[i] = 0;
    visible[i]   = 0;
    snk_point_fine& lmp = pts[i];
    if (!lmp.valid) return;
    const D3 P{lmp.pos[0], lmp.pos[1], lmp.pos[2]};
    const D3 r0 = qrot(f.q, P);
    const D3 Pc{r0.x + f.t[0], r0.y + f.t[1], r0.z + f.t[2]};
    const double z = Pc.z;
    if (z < 0)
    {
        lmp.valid = 0;
        return;
    }
    const double invz = 1.0 / z;
    const double xn = Pc.x * invz, yn = Pc.y * invz;
    const double u = K.fx * xn + K.s * yn + K.cx, v = K.fy * yn + K.cy;
    if (!in_image(f, u, v))
    {
        lmp.valid = 0;
        return;
    }
    const D3 cam = cam_position(f);
    const D3 PO{cam.x - P.x, cam.y - P.y, cam.z - P.z};
    const double dist = sqrt(PO.x * PO.x + PO.y * PO.y + PO.z * PO.z);
    double max_d = (double)lmp.ref_depth * sp.scale[lmp.ref_level];
    double min_d = max_d / sp.scale[sp.nlevels - 1];
    min_d /= sp.scale_factor;
    max_d *= sp.scale_factor;
    if (dist < min_d || dist > max_d)
    {
        lmp.valid = 0;
        return;
    }
    const double viewCos = (PO.x * lmp.normal[0] + PO.y * lmp.normal[1] + PO.z * lmp.normal[2]) / dist;
    if (viewCos < 0.5)
    {
        lmp.valid = 0;
        return;
    }
    visible[i] = 1;
    float r    = ((double)(float)viewCos > 0.998) ? 2.5f : 4.0f;
    if (th != 1.0f) r = __fmul_rn(r, th);
    const double pmax       = (double)lmp.ref_depth * sp.scale[lmp.ref_level];
    const double prediction = log(pmax / dist) / log_sf;
    r                       = (float)((double)r * pow(sp.scale_factor, prediction));
    const double rd = (double)r, r2 = rd * rd;
    const double right_pred = u - K.bf / z;
    const Desc q            = load_desc(lmp.desc);
    int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
    for_cells(f, u, v, r, [&](int pid) {
        const snk_keypoint_f64& kp = f.kp_un[pid];
        const double e = prediction - kp.octave;
        if (!(e * e < (1.2 * 1.2))) return;
        const double dx = kp.x - u, dy = kp.y - v;
        if (!((dx * dx + dy * dy) < r2)) return;
        if (f.kp_to_point[pid] >= 0) return;
        const float rpt = f.right_points[pid];
        if (rpt > 0)
        {
            const double er = fabs(right_pred - (double)rpt);
            if (er > (double)r * 0.5) return;
        }
        const int d = hamming(q, load_desc(f.desc + 4 * (size_t)pid));
        if (d < bestDist)
        {
            bestDist2  = bestDist;
            bestDist   = d;
            bestLevel2 = bestLevel;
            bestLevel  = kp.octave;
            bestIdx    = pid;
        }
        else if (d < bestDist2)
        {
            bestLevel2 = kp.octave;
            bestDist2  = d;
        }
    });
    if (bestIdx >= 0 && bestDist <= 100)
    {
        if (bestLevel == bestLevel2 && (float)bestDist > __fmul_rn(ratio, (float)bestDist2)) return;
        best_idx[i]  = bestIdx;
        best_dist[i] = bestDist;
        atomicMin(&winner[bestIdx], i);
    }
}

// SnakeORBMatcher::SearchByProjectionFrameFrame2, parallel part (:220-318)
__global__ void k_proj_coarse(FrameDev f, const snk_point_coarse* __restrict__ pts, int n, snk_camera K,
                              snk_scale_pyramid sp, float th, int feature_error, int motion,
                              int* __restrict__ best_idx, int* __restrict__ best_bin, int* __restrict__ winner)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    best_idx[i] = -1;
    best_bin[i] = 0;
    const snk_point_coarse& lmp = pts[i];
    const D3 wp{lmp.pos[0], lmp.pos[1], lmp.pos[2]};
    const D3 r0 = qrot(f.q, wp);
    const D3 Pc{r0.x + f.t[0], r0.y + f.t[1], r0.z + f.t[2]};
    const double z = Pc.z;
    if (z <= 0) return;
    const double invz = 1.0 / z;
    const double xn = Pc.x * invz, yn = Pc.y * invz;
    const double u = K.fx * xn + K.s * yn + K.cx, v = K.fy * yn + K.cy;
    if (!in_image(f, u, v)) return;
    const D3 cam = cam_position(f);
    const D3 PO{cam.x - wp.x, cam.y - wp.y, cam.z - wp.z};
    const double dist    = sqrt(PO.x * PO.x + PO.y * PO.y + PO.z * PO.z);
    const double viewCos = (PO.x * lmp.normal[0] + PO.y * lmp.normal[1] + PO.z * lmp.normal[2]) / dist;
    if (viewCos < 0.5) return;
    const int lvl = lmp.octave;
    const float r = (float)((double)th * sp.scale[lvl]);
    int lo, hi;
    if (motion == 1)
        lo = lvl - 1, hi = 100;
    else if (motion == 2)
        lo = 0, hi = lvl;
    else
        lo = lvl - 1, hi = lvl + 1;
    const double rd = (double)r, r2 = rd * rd;
    const double right_pred = u - K.bf / z;
    const Desc q            = load_desc(lmp.desc);
    int bestDist = 256, bestIdx = -1;
    for_cells(f, u, v, r, [&](int pid) {
        const snk_keypoint_f64& kp = f.kp_un[pid];
        if (kp.octave < lo || kp.octave > hi) return;
        const double dx = kp.x - u, dy = kp.y - v;
        if (!((dx * dx + dy * dy) < r2)) return;
        if (f.kp_to_point[pid] >= 0) return;
        const float rpt = f.right_points[pid];
        if (rpt > 0)
        {
            const double er = fabs(right_pred - (double)rpt);
            if (er > (double)r * 0.5) return;
        }
        const int d = hamming(q, load_desc(f.desc + 4 * (size_t)pid));
        if (d < bestDist)
        {
            bestDist = d;
            bestIdx  = pid;
        }
    });
    if (bestIdx >= 0 && bestDist <= feature_error)
    {
        float rot = (float)((double)lmp.angle - f.kp_un[bestIdx].angle);
        if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
        int bin = (int)roundf(__fmul_rn(rot, 1.0f / 30));
        if (bin == 30) bin = 0;
        best_idx[i] = bestIdx;
        best_bin[i] = bin;
        atomicMin(&winner[bestIdx], i);
    }
}

// first-come-first-served assignment as a parallel pass: keypoint k goes to the lowest point index that chose it
template <typename PT>
__global__ void k_assign(const int* __restrict__ kp_to_point_in, const int* __restrict__ winner, const PT* __restrict__ pts,
                         const int* __restrict__ best_bin, int n_kp, int* __restrict__ kp_to_point_out,
                         int* __restrict__ hist /*30 bins or null*/, int* __restrict__ n_matches)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    int got     = 0;
    if (k < n_kp)
    {
        int v = kp_to_point_in[k];
        const int w = winner[k];
        if (v < 0 && w != 0x7fffffff)
        {
            v   = pts[w].id;
            got = 1;
            if (hist) atomicAdd(&hist[best_bin[w]], 1);
        }
        kp_to_point_out[k] = v;
    }
    got = warp_sum(got);
    if ((threadIdx.x & 31) == 0 && got) atomicAdd(n_matches, got);
}

// rotation-histogram cull of the coarse search (ComputeThreeMaxima :27-68 and :334-351)
__global__ void k_rot_cull(const int* __restrict__ kp_to_point_in, const int* __restrict__ winner,
                           const int* __restrict__ best_bin, int n_kp, const int* __restrict__ hist,
                           int* __restrict__ kp_to_point_out, int* __restrict__ n_matches)
{
    __shared__ int keep[3];
    if (threadIdx.x == 0)
    {
        int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < 30; ++i)
        {
            const int s = hist[i];
            if (s > max1)
            {
                max3 = max2, max2 = max1, max1 = s;
                ind3 = ind2, ind2 = ind1, ind1 = i;
            }
            else if (s > max2)
            {
                max3 = max2, max2 = s;
                ind3 = ind2, ind2 = i;
            }
            else if (s > max3)
            {
                max3 = s, ind3 = i;
            }
        }
        if ((float)max2 < __fmul_rn(0.1f, (float)max1))
            ind2 = -1, ind3 = -1;
        else if ((float)max3 < __fmul_rn(0.1f, (float)max1))
            ind3 = -1;
        keep[0] = ind1, keep[1] = ind2, keep[2] = ind3;
    }
    __syncthreads();
    int removed = 0;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n_kp; k += gridDim.x * blockDim.x)
    {
        const int w = winner[k];
        if (kp_to_point_in[k] < 0 && w != 0x7fffffff)
        {
            const int b = best_bin[w];
            if (b != keep[0] && b != keep[1] && b != keep[2])
            {
                kp_to_point_out[k] = -1;
                ++removed;
            }
        }
    }
    if (removed) atomicSub(n_matches, removed);
}

// SnakeORBMatcher::SearchByProjectionFrameToKeyframe (:71-188): every accepted match removes a keypoint for all later
// points, so the points are walked in order by one warp; the lanes share the candidate evaluation of the current point.
__global__ void k_proj_frame_to_kf(FrameDev f, const snk_point_kf* __restrict__ pts, int n, snk_camera K, float th,
                                   int feature_error, int* __restrict__ kp_to_point_out, int* __restrict__ n_matches)
{
    const int lane = threadIdx.x;
    int matches    = 0;
    for (int i = 0; i < n; ++i)
    {
        const int id = pts[i].id;
        if (id < 0) continue;
        // sAlreadyFound: ids matched before the call (kp_to_point_in), not those added by this loop
        int found = 0;
        for (int k = lane; k < f.n; k += 32) found |= (f.kp_to_point[k] == id);
        if (__any_sync(0xffffffffu, found)) continue;
        const D3 wp{pts[i].pos[0], pts[i].pos[1], pts[i].pos[2]};
        const D3 r0 = qrot(f.q, wp);
        const D3 Pc{r0.x + f.t[0], r0.y + f.t[1], r0.z + f.t[2]};
        const double z = Pc.z;
        const double invz = 1.0 / z;
        const double xn = Pc.x * invz, yn = Pc.y * invz;
        const double u = K.fx * xn + K.s * yn + K.cx, v = K.fy * yn + K.cy;
        if (z <= 0) continue;
        if (!in_image(f, u, v)) continue;
        const float r  = th;
        const double r2 = (double)__fmul_rn(r, r);  // GetFeaturesInArea keeps r*r in float (Features.cpp:17)
        const double right_pred = u - K.bf / z;
        const Desc q = load_desc(pts[i].desc);
        // lanes take candidates round-robin in visiting order; key = dist << 20 | visiting rank keeps the first minimum
        uint32_t best = 0xffffffffu;
        int bestPid   = -1;
        int rank      = 0;
        for_cells(f, u, v, r, [&](int pid) {
            const int my = rank++;
            if ((my & 31) != lane) return;
            const snk_keypoint_f64& kp = f.kp_un[pid];
            const double dx = kp.x - u, dy = kp.y - v;
            if (!((dx * dx + dy * dy) < r2)) return;
            if (kp_to_point_out[pid] >= 0) return;
            const float rpt = f.right_points[pid];
            if (rpt > 0)
            {
                const float er = (float)fabs(right_pred - (double)rpt);
                if ((double)er > (double)r * 0.5) return;
            }
            const uint32_t key = ((uint32_t)hamming(q, load_desc(f.desc + 4 * (size_t)pid)) << 20) | (uint32_t)my;
            if (key < best)
            {
                best    = key;
                bestPid = pid;
            }
        });
#pragma unroll
        for (int s = 16; s > 0; s >>= 1)
        {
            const uint32_t ob = __shfl_xor_sync(0xffffffffu, best, s);
            const int op      = __shfl_xor_sync(0xffffffffu, bestPid, s);
            if (ob < best) best = ob, bestPid = op;
        }
        if (bestPid >= 0 && (int)(best >> 20) <= feature_error)
        {
            if (lane == 0) kp_to_point_out[bestPid] = id;
            ++matches;
        }
        __syncwarp();
    }
    if (lane == 0) *n_matches = matches;
}

// FeatureGrid2::create (saiga/vision/util/FeatureGrid2.h:130-192) for one frame: the stable counting sort is done as
// a bitonic sort of (cell << 13 | index) keys in shared memory; cell ranges come from binary searches on the sorted keys
// (an empty cell gets begin == end == number of keypoints in earlier cells, exactly the reference's running prefix).
__global__ void __launch_bounds__(256) k_feature_grid(FrameDev f, int n, int np2, int* __restrict__ permutation,
                                                      int* __restrict__ cells_out)
{
    extern __shared__ uint32_t gkeys[];
    const int tid = threadIdx.x, ncell = f.rows * f.cols;
    for (int i = tid; i < np2; i += 256)
    {
        uint32_t k = 0xffffffffu;
        if (i < n)
        {
            const double x = f.kp_un[i].x, y = f.kp_un[i].y;
            int cell = ncell;  // outside the bounds: sorted behind every cell, in input order
            if (in_image(f, x, y))
            {
                // LogicalCell: iFloor((p - bmin) * cellSizeInv); inside the bounds it is a valid cell
                const int cx = ifloor_d((x - f.bmin_x) * f.inv_x), cy = ifloor_d((y - f.bmin_y) * f.inv_y);
                cell = min(max(cy, 0), f.rows - 1) * f.cols + min(max(cx, 0), f.cols - 1);
            }
            k = ((uint32_t)cell << 13) | (uint32_t)i;
        }
        gkeys[i] = k;
    }
    __syncthreads();
    for (int k = 2; k <= np2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1)
        {
            for (int i = tid; i < np2; i += 256)
            {
                const int ixj = i ^ j;
                if (ixj > i)
                {
                    const uint32_t a = gkeys[i], b = gkeys[ixj];
                    if ((a > b) == ((i & k) == 0))
                    {
                        gkeys[i]   = b;
                        gkeys[ixj] = a;
                    }
                }
            }
            __syncthreads();
        }
    for (int p = tid; p < n; p += 256) permutation[gkeys[p] & 8191u] = p;
    for (int c = tid; c < ncell; c += 256)
    {
        int be[2];
#pragma unroll
        for (int e = 0; e < 2; ++e)
        {
            const uint32_t target = (uint32_t)(c + e) << 13;
            int lo = 0, hi = n;
            while (lo < hi)
            {
                const int mid = (lo + hi) >> 1;
                if (gkeys[mid] < target)
                    lo = mid + 1;
                else
                    hi = mid;
            }
            be[e] = lo;
        }
        cells_out[2 * c]     = be[0];
        cells_out[2 * c + 1] = be[1];
    }
}

// LoopORBmatcher::MatchBoW: one warp per vocabulary node of keyframe 1. Features live in exactly one node, so nodes
// are independent; inside a node the idx1 loop is sequential (greedy vbMatched2), the idx2 scan is split over lanes.
struct FeatVecDev
{
    int n_nodes;
    const int* node_ids;
    const int* offsets;
    const int* indices;
};
__device__ __forceinline__ void match_bow_node(const uint64_t* __restrict__ d1, const uint64_t* __restrict__ d2, const FeatVecDev& fv1,
                                               const FeatVecDev& fv2, const uint8_t* __restrict__ valid1,
                                               const uint8_t* __restrict__ valid2, int th, float ratio, int* __restrict__ match12,
                                               uint8_t* __restrict__ matched2, int* __restrict__ n_matches, int a, int lane)
{
    const int id = fv1.node_ids[a];
    int lo = 0, hi = fv2.n_nodes;
    while (lo < hi)
    {
        const int mid = (lo + hi) >> 1;
        if (fv2.node_ids[mid] < id)
            lo = mid + 1;
        else
            hi = mid;
    }
    if (lo >= fv2.n_nodes || fv2.node_ids[lo] != id) return;
    const int b0 = fv2.offsets[lo], b1 = fv2.offsets[lo + 1];
    int matches = 0;
    for (int i1 = fv1.offsets[a]; i1 < fv1.offsets[a + 1]; ++i1)
    {
        const int idx1 = fv1.indices[i1];
        if (!valid1[idx1]) continue;
        const Desc q = load_desc(d1 + 4 * (size_t)idx1);
        uint32_t k1 = (256u << 20) | 0xfffffu, k2 = 256u;  // best (dist << 20 | position), second-best distance
        for (int i2 = b0 + lane; i2 < b1; i2 += 32)
        {
            const int idx2 = fv2.indices[i2];
            if (matched2[idx2] || !valid2[idx2]) continue;
            const uint32_t d   = (uint32_t)hamming(q, load_desc(d2 + 4 * (size_t)idx2));
            const uint32_t key = (d << 20) | (uint32_t)(i2 - b0);
            if (d < (k1 >> 20))
            {
                k2 = k1 >> 20;
                k1 = key;
            }
            else if (d < k2)
                k2 = d;
        }
#pragma unroll
        for (int s = 16; s > 0; s >>= 1)
        {
            const uint32_t o1 = __shfl_xor_sync(0xffffffffu, k1, s), o2 = __shfl_xor_sync(0xffffffffu, k2, s);
            // two smallest distances of the union (multiset), first position on ties
            const uint32_t lo1 = min(k1, o1), hi1 = max(k1, o1);
            k2 = min(min(k2, o2), hi1 >> 20);
            k1 = lo1;
        }
        const int best1 = (int)(k1 >> 20), best2 = (int)k2;
        if (best1 < th && (float)best1 < __fmul_rn(ratio, (float)best2))
        {
            const int idx2 = fv2.indices[b0 + (int)(k1 & 0xfffffu)];
            if (lane == 0)
            {
                matched2[idx2] = 1;
                match12[idx1]  = idx2;
            }
            ++matches;
        }
        __syncwarp();
    }
    if (lane == 0 && matches) atomicAdd(n_matches, matches);
}
__global__ void k_match_bow(const uint64_t* __restrict__ d1, const uint64_t* __restrict__ d2, FeatVecDev fv1, FeatVecDev fv2,
                            const uint8_t* __restrict__ valid1, const uint8_t* __restrict__ valid2, int th, float ratio,
                            int* __restrict__ match12, uint8_t* __restrict__ matched2, int* __restrict__ n_matches)
{
    const int a = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (a >= fv1.n_nodes) return;
    match_bow_node(d1, d2, fv1, fv2, valid1, valid2, th, ratio, match12, matched2, n_matches, a, threadIdx.x & 31);
}

// One keyframe against many (LoopDetector::ComputeSim3(new_kf) walks candidateKFs, LoopDetector.cpp:384-391; MatchBoW reads
// nothing the previous candidate's pass wrote): blockIdx.y = candidate, all candidates' arrays in one uploaded blob.
struct BowCandDev
{
    int desc_off, valid_off, ids_off, offs_off, inds_off;  // byte offsets into the blob
    int n_nodes, n2, m2_off;                               // m2_off: offset of the candidate's vbMatched2 flags
};
__global__ void k_match_bow_batch(const unsigned char* __restrict__ blob, const BowCandDev* __restrict__ cands, const uint64_t* __restrict__ d1,
                                  FeatVecDev fv1, const uint8_t* __restrict__ valid1, int n1, int th, float ratio,
                                  int* __restrict__ match12, uint8_t* __restrict__ matched2, int* __restrict__ n_matches)
{
    const int a = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (a >= fv1.n_nodes) return;
    const BowCandDev cd = cands[blockIdx.y];
    if (cd.n_nodes == 0 || cd.n2 == 0) return;
    const FeatVecDev fv2{cd.n_nodes, reinterpret_cast<const int*>(blob + cd.ids_off), reinterpret_cast<const int*>(blob + cd.offs_off),
                         reinterpret_cast<const int*>(blob + cd.inds_off)};
    match_bow_node(d1, reinterpret_cast<const uint64_t*>(blob + cd.desc_off), fv1, fv2, valid1, blob + cd.valid_off, th, ratio,
                   match12 + (size_t)blockIdx.y * n1, matched2 + cd.m2_off, n_matches + blockIdx.y, a, threadIdx.x & 31);
}
}  // namespace snk
