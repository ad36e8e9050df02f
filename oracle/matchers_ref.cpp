// TEST INFRASTRUCTURE ONLY — CPU oracle, never on the product path.
//
// Restatements of the reference's Hamming matchers on the plain structs of include/snakeorb.h. PINNED: the originals
// (which work on Snake::Frame / Map / globals) are compiled unmodified into oracle/_ref/libsnake_ref.so
// (ref_snake_shim.cpp) and tests/test_oracle_snake_live.py requires identical results on 1000 seeded cases per function;
// their outputs are frozen in tests/golden/golden_snake.npz. Paths relative to /root/reference/.
// Built with -ffp-contract=off: the double geometry is evaluated unfused, in the reference's operation order.
#include "matchers_ref.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <set>
#include <vector>

namespace orbref
{
// External/saiga/src/saiga/vision/features/Features.h:124-138
int hamming(const uint64_t* a, const uint64_t* b)
{
    int d = 0;
    for (int i = 0; i < 4; ++i) d += __builtin_popcountll(a[i] ^ b[i]);
    return d;
}

// Features.h:323-354
void match_knn2(const uint64_t* d1, int n, const uint64_t* d2, int m, snk_knn2* out, int threads)
{
#pragma omp parallel for num_threads(threads)
    for (int i = 0; i < n; ++i)
    {
        snk_knn2 r{1000, -1, 1000, -1};
        for (int j = 0; j < m; ++j)
        {
            int dis = hamming(d1 + 4 * (size_t)i, d2 + 4 * (size_t)j);
            if (dis < r.dist0)
            {
                r.dist1 = r.dist0;
                r.idx1  = r.idx0;
                r.dist0 = dis;
                r.idx0  = j;
            }
            else if (dis < r.dist1)
            {
                r.dist1 = dis;
                r.idx1  = j;
            }
        }
        out[i] = r;
    }
}

// Features.h:394-414
int filter_matches(const snk_knn2* knn2, int n, int threshold, float ratio, int32_t* matches_ij)
{
    int c = 0;
    for (int i = 0; i < n; ++i)
    {
        if (knn2[i].dist0 > threshold) continue;
        if (float(knn2[i].dist0) > float(knn2[i].dist1) * ratio) continue;
        matches_ij[2 * c]     = i;
        matches_ij[2 * c + 1] = knn2[i].idx0;
        ++c;
    }
    return c;
}

static inline int iRoundF(float v) { return (int)(v + (v >= 0 ? 0.5f : -0.5f)); }

// Snake/Preprocess/Preprocess.cpp:122-242 (after the Rectification::Forward loops)
int stereo_match(const snk_stereo_params* p, const snk_keypoint_f64* left, const uint64_t* dl, int nl,
                 const snk_keypoint_f64* right, const uint64_t* dr, int nr, float* right_points, float* depth,
                 long long* pair_ops)
{
    for (int i = 0; i < nl; ++i) right_points[i] = depth[i] = -1000.f;  // Frame::allocateTmp
    long long ops = 0;
    if (nl == 0 || nr == 0)
    {
        if (pair_ops) *pair_ops = 0;
        return 0;
    }
    int min_y = 1023123, max_y = -19284;
    const float min_disp = 0;
    const float max_disp = (float)(p->bf * 0.5);
    for (int i = 0; i < nr; ++i)
    {
        int r = iRoundF((float)right[i].y);  // iRound(float) receives a double
        min_y = std::min(min_y, r);
        max_y = std::max(max_y, r);
    }
    std::vector<std::vector<int>> row_map(max_y - min_y + 1);
    for (int i = 0; i < nr; ++i) row_map[iRoundF((float)right[i].y) - min_y].push_back(i);

    int num_matches = 0;
    for (int i = 0; i < nl; ++i)
    {
        const snk_keypoint_f64& kp = left[i];
        int y                      = iRoundF((float)kp.y);
        int y_center               = y - min_y;
        float r                    = (float)std::ceil(2.0f * p->scale[kp.octave]);
        int best_id = -1, best_dist = 250, second = 250;
        for (int row = (int)(y_center - r); row <= y_center + r; ++row)
        {
            if (row < 0 || row >= (int)row_map.size()) continue;
            for (int o : row_map[row])
            {
                double disparity = kp.x - right[o].x;
                if (disparity < min_disp || disparity > max_disp) continue;
                if (std::abs(kp.octave - right[o].octave) > 1) continue;
                int dist = hamming(dl + 4 * (size_t)i, dr + 4 * (size_t)o);
                ++ops;
                if (dist < best_dist)
                {
                    second    = best_dist;
                    best_dist = dist;
                    best_id   = o;
                }
                else if (dist < second)
                {
                    second = dist;
                }
            }
        }
        if (best_dist > (p->relaxed ? 75 : 40)) continue;
        if (best_dist > (p->relaxed ? 0.9 : 0.7) * second) continue;
        double a1 = kp.angle, a2 = right[best_id].angle;
        float rot = (float)std::min(std::abs(a1 - a2), std::min(std::abs((a1 + 365) - a2), std::abs(a1 - (a2 + 365))));
        if (rot > (p->relaxed ? 25 : 5)) continue;
        double right_point = right[best_id].x;
        double disparity   = kp.x - right_point;
        if (disparity <= 0.001)
        {
            disparity   = 0.001;
            right_point = kp.x - disparity;
        }
        right_points[i] = (float)right_point;
        depth[i]        = (float)(p->bf / disparity);
        ++num_matches;
    }
    if (pair_ops) *pair_ops = ops;
    return num_matches;
}

// Snake/Preprocess/Preprocess.cpp:79-120 with distortNormalizedPoint of
// External/saiga/src/saiga/vision/cameraModel/Distortion.h:128-170
int stereo_from_rgbd(const snk_rgbd_params* p, const snk_keypoint_f64* kp, int n, const float* depth_img, int rows,
                     int cols, size_t pitch, float* right_points, float* depth)
{
    int num = 0;
    for (int i = 0; i < n; ++i)
    {
        const double x = (kp[i].x - p->cx) / p->fx, y = (kp[i].y - p->cy) / p->fy;
        const double k1 = p->dist[0], k2 = p->dist[1], k3 = p->dist[2], k4 = p->dist[3], k5 = p->dist[4], k6 = p->dist[5];
        const double p1 = p->dist[6], p2 = p->dist[7];
        const double x2 = x * x, y2 = y * y;
        const double r2 = x2 + y2, _2xy = 2.0 * x * y;
        const double r4 = r2 * r2;
        const double r6 = r4 * r2;
        const double radial_u  = 1.0 + k1 * r2 + k2 * r4 + k3 * r6;
        const double radial_v  = 1.0 + k4 * r2 + k5 * r4 + k6 * r6;
        const double radial_iv = 1.0 / radial_v;
        const double radial    = radial_u * radial_iv;
        const double tx = p1 * _2xy + p2 * (r2 + 2.0 * x2);
        const double ty = p1 * (r2 + 2.0 * y2) + p2 * _2xy;
        double xd = x * radial + tx, yd = y * radial + ty;
        if (r2 > 100000.0 * 100000.0) xd = yd = 100000;
        const double u = p->dfx * xd + p->ds * yd + p->dcx, v = p->dfy * yd + p->dcy;
        const int px = (int)(u + 0.5), py = (int)(v + 0.5);
        float d = 0.f;  // the reference asserts the lookup is inside the image; outside counts as no measurement here
        if (px >= 0 && px < cols && py >= 0 && py < rows) d = *(const float*)((const char*)depth_img + (size_t)py * pitch + (size_t)px * 4);
        if (d > 0)
        {
            depth[i]        = d;
            right_points[i] = (float)(kp[i].x - p->bf / d);
            ++num;
        }
        else
        {
            depth[i]        = -1;
            right_points[i] = -1;
        }
    }
    return num;
}

// ---------------------------------------------------------------------------------------------
// Preprocess::undistortKeypoints (Snake/Preprocess/Preprocess.cpp:55-77)
// ---------------------------------------------------------------------------------------------
namespace
{
// distortNormalizedPoint with its point Jacobian (saiga/vision/cameraModel/Distortion.h:128-212)
inline void distort_with_jacobian(const double* D, double x, double y, double& xd, double& yd, double* J /*2x2 row-major or null*/)
{
    const double k1 = D[0], k2 = D[1], k3 = D[2], k4 = D[3], k5 = D[4], k6 = D[5], p1 = D[6], p2 = D[7];
    const double x2 = x * x, y2 = y * y;
    const double r2 = x2 + y2, _2xy = 2.0 * x * y;
    const double r4 = r2 * r2;
    const double r6 = r4 * r2;
    const double radial_u  = 1.0 + k1 * r2 + k2 * r4 + k3 * r6;
    const double radial_v  = 1.0 + k4 * r2 + k5 * r4 + k6 * r6;
    const double radial_iv = 1.0 / radial_v;
    const double radial    = radial_u * radial_iv;
    const double tx = p1 * _2xy + p2 * (r2 + 2.0 * x2);
    const double ty = p1 * (r2 + 2.0 * y2) + p2 * _2xy;
    xd = x * radial + tx;
    yd = y * radial + ty;
    if (r2 > 100000.0 * 100000.0) xd = yd = 100000;
    if (!J) return;
    double u0 = k1 * (2 * x), u1 = k1 * (2 * y);
    u0 += k2 * (4 * x * x * x + 4 * x * y * y);
    u1 += k2 * (4 * y * y * y + 4 * y * x * x);
    u0 += k3 * (6 * x2 * x2 * x + 12 * x2 * x * y2 + 6 * x * y2 * y2);
    u1 += k3 * (6 * y2 * y2 * y + 12 * y2 * y * x2 + 6 * y * x2 * x2);
    double v0 = k4 * (2 * x), v1 = k4 * (2 * y);
    v0 += k5 * (4 * x * x * x + 4 * x * y * y);
    v1 += k5 * (4 * y * y * y + 4 * y * x * x);
    v0 += k6 * (6 * x2 * x2 * x + 12 * x2 * x * y2 + 6 * x * y2 * y2);
    v1 += k6 * (6 * y2 * y2 * y + 12 * y2 * y * x2 + 6 * y * x2 * x2);
    // J_rad: both rows equal (Distortion.h:181-196)
    const double jr0 = (u0 * radial_v - v0 * radial_u) * radial_iv * radial_iv;
    const double jr1 = (u1 * radial_v - v1 * radial_u) * radial_iv * radial_iv;
    J[0] = (x * jr0 + radial) + (2 * p1 * y + 6 * p2 * x);
    J[1] = (x * jr1) + (2 * p1 * x + 2 * p2 * y);
    J[2] = (y * jr0) + (2 * p2 * y + 2 * p1 * x);
    J[3] = (y * jr1 + radial) + (2 * p2 * x + 6 * p1 * y);
}

// Eigen::LDLT<Matrix2d>::solve as the vendored Eigen executes it for a 2x2 (Eigen/src/Cholesky/LDLT.h:300-400
// unblocked factorisation with diagonal pivoting, :560-600 solve with the 1/min() pivot cut-off)
inline void ldlt2_solve(double a, double b, double c, double r0, double r1, double& s0, double& s1)
{
    const bool swap = !(std::abs(a) >= std::abs(c));  // maxCoeff keeps the first of equal maxima; NaN never wins
    double d0 = swap ? c : a, rest = swap ? a : c;
    double l = b;
    if (!(std::abs(d0) > 0))
    {  // whole diagonal zero: transpositions reset, D = 0 => the solve zeroes both rows
        s0 = s1 = 0;
        return;
    }
    l /= d0;
    const double d1 = rest - l * (d0 * l);
    double y0 = swap ? r1 : r0, y1 = swap ? r0 : r1;
    y1 -= l * y0;
    const double tol = std::numeric_limits<double>::min();
    y0 = (std::abs(d0) > tol) ? y0 / d0 : 0.0;
    y1 = (std::abs(d1) > tol) ? y1 / d1 : 0.0;
    y0 -= l * y1;
    s0 = swap ? y1 : y0;
    s1 = swap ? y0 : y1;
}

// undistortPointGN (Distortion.h:241-285), point == guess as Preprocess.cpp:68 calls it
inline void undistort_gn(const double* D, double px, double py, double& ox, double& oy)
{
    double x = px, y = py, lx = px, ly = py;
    double last_chi2 = std::numeric_limits<double>::infinity();
    for (int it = 0; it < 5; ++it)
    {
        double J[4], xd, yd;
        distort_with_jacobian(D, x, y, xd, yd, J);
        const double r0 = xd - px, r1 = yd - py;
        const double chi2 = r0 * r0 + r1 * r1;
        if (chi2 > last_chi2)
        {
            x = lx;
            y = ly;
            continue;
        }
        const double a = J[0] * J[0] + J[2] * J[2], b = J[1] * J[0] + J[3] * J[2], c = J[1] * J[1] + J[3] * J[3];
        const double g0 = (-J[0]) * r0 + (-J[2]) * r1, g1 = (-J[1]) * r0 + (-J[3]) * r1;
        lx        = x;
        ly        = y;
        last_chi2 = chi2;
        double d0, d1;
        ldlt2_solve(a, b, c, g0, g1, d0, d1);
        x += d0;
        y += d1;
    }
    double xd, yd;
    distort_with_jacobian(D, x, y, xd, yd, nullptr);
    const double r0 = xd - px, r1 = yd - py;
    if (r0 * r0 + r1 * r1 > last_chi2)
    {
        x = lx;
        y = ly;
    }
    ox = x;
    oy = y;
}
}  // namespace

void undistort_keypoints(const snk_rectify* r, const snk_keypoint_f64* kps, int n, snk_keypoint_f64* kp_un,
                         double* normalized)
{
    for (int i = 0; i < n; ++i)
    {
        kp_un[i] = kps[i];
        const double px = (kps[i].x - r->cx) / r->fx, py = (kps[i].y - r->cy) / r->fy;  // unproject2 (no skew)
        double nx, ny;
        undistort_gn(r->dist, px, py, nx, ny);
        // Eigen::Quaternion * Vector3 (Eigen/src/Geometry/Quaternion.h _transformVector): uv = 2 (q.vec x v);
        // v + w uv + q.vec x uv
        const double qx = r->quat[0], qy = r->quat[1], qz = r->quat[2], qw = r->quat[3];
        const double v0 = nx, v1 = ny, v2 = 1.0;
        double u0 = qy * v2 - qz * v1, u1 = qz * v0 - qx * v2, u2 = qx * v1 - qy * v0;
        u0 += u0;
        u1 += u1;
        u2 += u2;
        const double c0 = qy * u2 - qz * u1, c1 = qz * u0 - qx * u2, c2 = qx * u1 - qy * u0;
        const double p0 = (v0 + qw * u0) + c0, p1 = (v1 + qw * u1) + c1, p2 = (v2 + qw * u2) + c2;
        const double mx = p0 / p2, my = p1 / p2;
        normalized[2 * i]     = mx;
        normalized[2 * i + 1] = my;
        kp_un[i].x = r->dfx * mx + r->ds * my + r->dcx;  // K_dst.normalizedToImage (Intrinsics4.h:80)
        kp_un[i].y = r->dfy * my + r->dcy;
    }
}

// ---------------------------------------------------------------------------------------------
// projection geometry helpers
// ---------------------------------------------------------------------------------------------
namespace
{
struct V3
{
    double x, y, z;
};
inline V3 cross(const V3& a, const V3& b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }

// Eigen::QuaternionBase::_transformVector: uv = 2 * (q.vec x v); v + w*uv + q.vec x uv
inline V3 rotate(const double q[4], const V3& v)
{
    V3 qv{q[0], q[1], q[2]};
    V3 uv = cross(qv, v);
    uv    = {uv.x + uv.x, uv.y + uv.y, uv.z + uv.z};
    V3 c  = cross(qv, uv);
    return {v.x + q[3] * uv.x + c.x, v.y + q[3] * uv.y + c.y, v.z + q[3] * uv.z + c.z};
}
// Sophus SE3 * point
inline V3 se3_apply(const double q[4], const double t[3], const V3& p)
{
    V3 r = rotate(q, p);
    return {r.x + t[0], r.y + t[1], r.z + t[2]};
}
// Sophus SE3::inverse().translation(): conj(q) * (t * -1)
inline V3 camera_position(const double q[4], const double t[3])
{
    double qi[4] = {-q[0], -q[1], -q[2], q[3]};
    return rotate(qi, V3{t[0] * -1, t[1] * -1, t[2] * -1});
}
inline int iFloorD(double v) { return ((int)v) - (((int)v) > v); }

struct Cell
{
    int cx, cy;
};
// External/saiga/src/saiga/vision/util/FeatureGrid2.h:24-60
inline Cell cell_clamped(const snk_grid_view& g, double px, double py)
{
    double x = (px - g.bmin_x) * g.cell_size_inv_x;
    double y = (py - g.bmin_y) * g.cell_size_inv_y;
    Cell c{iFloorD(x), iFloorD(y)};
    c.cx = std::min(std::max(c.cx, 0), g.cols - 1);
    c.cy = std::min(std::max(c.cy, 0), g.rows - 1);
    return c;
}
inline bool in_image(const snk_grid_view& g, double x, double y)
{
    return x >= g.bmin_x && y >= g.bmin_y && x < g.bmax_x && y < g.bmax_y;
}

// Visit grid candidates in the reference's order: cx outer, cy inner, index ascending (Snake/Map/Features.cpp:13-76).
// minMaxCellWithRadius takes a *float* radius (FeatureGrid2.h:24) while the distance test uses the caller's r*r.
template <typename F>
inline void for_candidates(const snk_frame_view* f, double px, double py, float r_cells, double r2, F&& fn)
{
    const snk_grid_view& g = f->grid;
    Cell lo                = cell_clamped(g, px - r_cells, py - r_cells);
    Cell hi                = cell_clamped(g, px + r_cells, py + r_cells);
    for (int cx = lo.cx; cx <= hi.cx; ++cx)
        for (int cy = lo.cy; cy <= hi.cy; ++cy)
        {
            int b = g.cells[2 * (cy * g.cols + cx)], e = g.cells[2 * (cy * g.cols + cx) + 1];
            for (int pid = b; pid < e; ++pid)
            {
                double dx = f->kp_un[pid].x - px, dy = f->kp_un[pid].y - py;
                bool inside = (dx * dx + dy * dy) < r2;
                fn(pid, inside);
            }
        }
}

// Intrinsics4.h:53-59
inline void project(const snk_camera* K, const V3& X, double& u, double& v)
{
    double invz = 1.0 / X.z;
    double x = X.x * invz, y = X.y * invz;
    u = K->fx * x + K->s * y + K->cx;
    v = K->fy * y + K->cy;
}
}  // namespace

// External/saiga/src/saiga/vision/util/FeatureGrid2.h:130-192 (FeatureGrid2::create) with the cell lookup of :37-52
void feature_grid(const snk_grid_view* b, const snk_keypoint_f64* kps, int n, int32_t* permutation, int32_t* cells)
{
    const int ncell = b->rows * b->cols;
    std::vector<int> cid(n), first(ncell, 0), second(ncell, 0);
    for (int i = 0; i < n; ++i)
    {
        if (in_image(*b, kps[i].x, kps[i].y))
        {
            const int cx = iFloorD((kps[i].x - b->bmin_x) * b->cell_size_inv_x);
            const int cy = iFloorD((kps[i].y - b->bmin_y) * b->cell_size_inv_y);
            cid[i]       = cy * b->cols + cx;
            first[cid[i]]++;
        }
        else
            cid[i] = -1;
    }
    int total = 0;
    for (int c = 0; c < ncell; ++c)
    {
        int count = first[c];
        first[c] = second[c] = total;
        total += count;
    }
    for (int i = 0; i < n; ++i)
        permutation[i] = (cid[i] == -1) ? total++ : second[cid[i]]++;
    for (int c = 0; c < ncell; ++c) cells[2 * c] = first[c], cells[2 * c + 1] = second[c];
}

// Snake/Tracking/SnakeORBMatcher.cpp:365-526
int search_projection_fine(const snk_frame_view* f, snk_point_fine* pts, int n, const snk_camera* K,
                           const snk_scale_pyramid* sp, float th, float ratio, int32_t* kp_to_point_out,
                           int32_t* best_idx, int32_t* best_dist, uint8_t* visible, long long* pair_ops)
{
    const V3 cam       = camera_position(f->pose_q, f->pose_t);
    const bool bFactor = th != 1.0;
    const double logsf = std::log(sp->scale_factor);
    long long ops      = 0;
    std::memcpy(kp_to_point_out, f->kp_to_point, sizeof(int32_t) * f->n);
    for (int i = 0; i < n; ++i)
    {
        best_idx[i]  = -1;
        best_dist[i] = 0;
        visible[i]   = 0;
        snk_point_fine& lmp = pts[i];
        if (!lmp.valid) continue;
        V3 P{lmp.pos[0], lmp.pos[1], lmp.pos[2]};
        V3 Pc    = se3_apply(f->pose_q, f->pose_t, P);
        double z = Pc.z;
        if (z < 0)
        {
            lmp.valid = 0;
            continue;
        }
        double u, v;
        project(K, Pc, u, v);
        if (!in_image(f->grid, u, v))
        {
            lmp.valid = 0;
            continue;
        }
        V3 PO{cam.x - P.x, cam.y - P.y, cam.z - P.z};
        double dist = std::sqrt(PO.x * PO.x + PO.y * PO.y + PO.z * PO.z);
        // ScalePyramid::EstimateMinMaxDistance (ScalePyramid.h:26-41)
        double max_d = (double)lmp.ref_depth * sp->scale[lmp.ref_level];
        double min_d = max_d / sp->scale[sp->nlevels - 1];
        min_d /= sp->scale_factor;
        max_d *= sp->scale_factor;
        if (dist < min_d || dist > max_d)
        {
            lmp.valid = 0;
            continue;
        }
        double viewCos = (PO.x * lmp.normal[0] + PO.y * lmp.normal[1] + PO.z * lmp.normal[2]) / dist;
        if (viewCos < 0.5)
        {
            lmp.valid = 0;
            continue;
        }
        visible[i] = 1;
        // RadiusByViewingCos(float) (:357-363): the double viewCos is narrowed to float first
        float r = ((float)viewCos > 0.998) ? 2.5f : 4.0f;
        if (bFactor) r *= th;
        // PredictScaleLevel (ScalePyramid.h:44-57), ScaleForContiniousLevel (:78)
        double pmax       = (double)lmp.ref_depth * sp->scale[lmp.ref_level];
        double prediction = std::log(pmax / dist) / logsf;
        r                 = (float)(r * std::pow(sp->scale_factor, prediction));
        const double rd   = r;
        const double r2   = rd * rd;

        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        const double right_pred = u - K->bf / z;  // stereo_cam.LeftPointToRight
        for_candidates(f, u, v, r, r2, [&](int pid, bool inside) {
            double e = prediction - f->kp_un[pid].octave;
            if (!(e * e < (1.2 * 1.2))) return;
            if (!inside) return;
            if (f->kp_to_point[pid] >= 0) return;
            if (f->right_points[pid] > 0)
            {
                double er = std::abs(right_pred - f->right_points[pid]);
                if (er > r * 0.5) return;
            }
            int d = hamming(lmp.desc, f->desc + 4 * (size_t)pid);
            ++ops;
            if (d < bestDist)
            {
                bestDist2  = bestDist;
                bestDist   = d;
                bestLevel2 = bestLevel;
                bestLevel  = f->kp_un[pid].octave;
                bestIdx    = pid;
            }
            else if (d < bestDist2)
            {
                bestLevel2 = f->kp_un[pid].octave;
                bestDist2  = d;
            }
        });
        if (bestIdx >= 0 && bestDist <= 100)  // TH_HIGH
        {
            if (bestLevel == bestLevel2 && bestDist > ratio * bestDist2) continue;
            best_idx[i]  = bestIdx;
            best_dist[i] = bestDist;
        }
    }
    int matches = 0;
    for (int i = 0; i < n; ++i)  // :516-524 first come, first served
    {
        int b = best_idx[i];
        if (b == -1) continue;
        if (kp_to_point_out[b] >= 0) continue;
        kp_to_point_out[b] = pts[i].id;
        ++matches;
    }
    if (pair_ops) *pair_ops = ops;
    return matches;
}

// Snake/Tracking/SnakeORBMatcher.cpp:27-68
static void three_maxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3)
{
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; ++i)
    {
        const int s = (int)histo[i].size();
        if (s > max1)
        {
            max3 = max2;
            max2 = max1;
            max1 = s;
            ind3 = ind2;
            ind2 = ind1;
            ind1 = i;
        }
        else if (s > max2)
        {
            max3 = max2;
            max2 = s;
            ind3 = ind2;
            ind2 = i;
        }
        else if (s > max3)
        {
            max3 = s;
            ind3 = i;
        }
    }
    if (max2 < 0.1f * (float)max1)
    {
        ind2 = -1;
        ind3 = -1;
    }
    else if (max3 < 0.1f * (float)max1)
    {
        ind3 = -1;
    }
}

// Snake/Tracking/SnakeORBMatcher.cpp:191-354
int search_projection_coarse(const snk_frame_view* f, const snk_point_coarse* pts, int n, const snk_camera* K,
                             const snk_scale_pyramid* sp, float th, int feature_error, int motion,
                             int32_t* kp_to_point_out, long long* pair_ops)
{
    const int HISTO    = 30;
    const float factor = 1.0f / HISTO;
    std::vector<int> hist[30];
    const V3 cam  = camera_position(f->pose_q, f->pose_t);
    long long ops = 0;
    std::memcpy(kp_to_point_out, f->kp_to_point, sizeof(int32_t) * f->n);
    std::vector<int> bidx(n, -1), bbin(n, 0);
    for (int i = 0; i < n; ++i)
    {
        const snk_point_coarse& lmp = pts[i];
        V3 wp{lmp.pos[0], lmp.pos[1], lmp.pos[2]};
        V3 Pc = se3_apply(f->pose_q, f->pose_t, wp);
        double u, v;
        project(K, Pc, u, v);
        double z = Pc.z;
        if (z <= 0) continue;
        if (!in_image(f->grid, u, v)) continue;
        V3 PO{cam.x - wp.x, cam.y - wp.y, cam.z - wp.z};
        double dist    = std::sqrt(PO.x * PO.x + PO.y * PO.y + PO.z * PO.z);
        double viewCos = (PO.x * lmp.normal[0] + PO.y * lmp.normal[1] + PO.z * lmp.normal[2]) / dist;
        if (viewCos < 0.5) continue;
        int lvl = lmp.octave;
        float r = th;
        r       = (float)(r * sp->scale[lvl]);
        int lo, hi;
        if (motion == 1)
            lo = lvl - 1, hi = 100;
        else if (motion == 2)
            lo = 0, hi = lvl;
        else
            lo = lvl - 1, hi = lvl + 1;
        const double rd = r, r2 = rd * rd;
        int bestDist = 256, bestIdx = -1;
        const double right_pred = u - K->bf / z;
        for_candidates(f, u, v, r, r2, [&](int pid, bool inside) {
            int oc = f->kp_un[pid].octave;
            if (oc < lo || oc > hi) return;
            if (!inside) return;
            if (f->kp_to_point[pid] >= 0) return;
            if (f->right_points[pid] > 0)
            {
                double er = std::abs(right_pred - f->right_points[pid]);
                if (er > r * 0.5) return;
            }
            int d = hamming(lmp.desc, f->desc + 4 * (size_t)pid);
            ++ops;
            if (d < bestDist)
            {
                bestDist = d;
                bestIdx  = pid;
            }
        });
        if (bestIdx >= 0 && bestDist <= feature_error)
        {
            float rot = (float)(lmp.angle - f->kp_un[bestIdx].angle);
            if (rot < 0.0) rot += 360.0f;
            int bin = (int)std::round(rot * factor);
            if (bin == HISTO) bin = 0;
            bidx[i] = bestIdx;
            bbin[i] = bin;
        }
    }
    int matches = 0;
    for (int i = 0; i < n; ++i)
    {
        int b = bidx[i];
        if (b == -1) continue;
        if (kp_to_point_out[b] >= 0) continue;
        kp_to_point_out[b] = pts[i].id;
        hist[bbin[i]].push_back(b);
        ++matches;
    }
    int i1 = -1, i2 = -1, i3 = -1;
    three_maxima(hist, HISTO, i1, i2, i3);
    for (int i = 0; i < HISTO; ++i)
        if (i != i1 && i != i2 && i != i3)
            for (int b : hist[i])
            {
                kp_to_point_out[b] = -1;
                --matches;
            }
    if (pair_ops) *pair_ops = ops;
    return matches;
}

// Snake/Tracking/SnakeORBMatcher.cpp:71-188 (mbCheckOrientation = false, :87)
int search_projection_frame_to_kf(const snk_frame_view* f, const snk_point_kf* pts, int n, const snk_camera* K,
                                  float th, int feature_error, int32_t* kp_to_point_out, long long* pair_ops)
{
    std::set<int> found;
    long long ops = 0;
    std::memcpy(kp_to_point_out, f->kp_to_point, sizeof(int32_t) * f->n);
    for (int i = 0; i < f->n; ++i)
        if (f->kp_to_point[i] >= 0) found.insert(f->kp_to_point[i]);
    int matches = 0;
    for (int i = 0; i < n; ++i)
    {
        if (pts[i].id < 0) continue;
        if (found.count(pts[i].id)) continue;
        V3 wp{pts[i].pos[0], pts[i].pos[1], pts[i].pos[2]};
        V3 Pc = se3_apply(f->pose_q, f->pose_t, wp);
        double u, v;
        project(K, Pc, u, v);
        double z = Pc.z;
        if (z <= 0) continue;
        if (!in_image(f->grid, u, v)) continue;
        float r          = th;
        const float r2f  = r * r;  // GetFeaturesInArea(…, float r): r2 is float (Features.cpp:17)
        int bestDist = 256, bestIdx = -1;
        const double right_pred = u - K->bf / z;
        for_candidates(f, u, v, r, (double)r2f, [&](int pid, bool inside) {
            if (!inside) return;
            if (kp_to_point_out[pid] >= 0) return;
            if (f->right_points[pid] > 0)
            {
                const float er = (float)std::abs(right_pred - f->right_points[pid]);
                if (er > r * 0.5) return;
            }
            int d = hamming(pts[i].desc, f->desc + 4 * (size_t)pid);
            ++ops;
            if (d < bestDist)
            {
                bestDist = d;
                bestIdx  = pid;
            }
        });
        if (bestIdx >= 0 && bestDist <= feature_error)
        {
            kp_to_point_out[bestIdx] = pts[i].id;
            ++matches;
        }
    }
    if (pair_ops) *pair_ops = ops;
    return matches;
}

// Snake/LoopClosing/LoopORBMatcher.cpp:121-215
int match_bow(const uint64_t* d1, int n1, const uint64_t* d2, int n2, const snk_featvec_view* fv1,
              const snk_featvec_view* fv2, const uint8_t* valid1, const uint8_t* valid2, int th, float ratio,
              int32_t* match12, long long* pair_ops)
{
    for (int i = 0; i < n1; ++i) match12[i] = -1;
    std::vector<char> matched2(n2, 0);
    long long ops = 0;
    int nmatches  = 0;
    int a = 0, b = 0;
    while (a < fv1->n_nodes && b < fv2->n_nodes)
    {
        int ida = fv1->node_ids[a], idb = fv2->node_ids[b];
        if (ida == idb)
        {
            for (int i1 = fv1->offsets[a]; i1 < fv1->offsets[a + 1]; ++i1)
            {
                int idx1 = fv1->indices[i1];
                if (!valid1[idx1]) continue;
                int best1 = 256, bestIdx2 = -1, best2 = 256;
                for (int i2 = fv2->offsets[b]; i2 < fv2->offsets[b + 1]; ++i2)
                {
                    int idx2 = fv2->indices[i2];
                    if (matched2[idx2] || !valid2[idx2]) continue;
                    int dist = hamming(d1 + 4 * (size_t)idx1, d2 + 4 * (size_t)idx2);
                    ++ops;
                    if (dist < best1)
                    {
                        best2    = best1;
                        best1    = dist;
                        bestIdx2 = idx2;
                    }
                    else if (dist < best2)
                    {
                        best2 = dist;
                    }
                }
                if (best1 < th)
                {
                    if ((float)best1 < ratio * (float)best2)
                    {
                        matched2[bestIdx2] = 1;
                        match12[idx1]      = bestIdx2;
                        ++nmatches;
                    }
                }
            }
            ++a;
            ++b;
        }
        else if (ida < idb)
        {
            while (a < fv1->n_nodes && fv1->node_ids[a] < idb) ++a;  // std::lower_bound on a sorted vector
        }
        else
        {
            while (b < fv2->n_nodes && fv2->node_ids[b] < ida) ++b;
        }
    }
    if (pair_ops) *pair_ops = ops;
    return nmatches;
}
// ---------------------------------------------------------------------------------------------
// Local mapping / loop closing searches, pinned like the rest of this file on the reference's own
// compiled MappingORBMatcher.cpp / LoopORBMatcher.cpp / MonoInitializer.cpp (oracle/_ref/libsnake_ref.so).
// ---------------------------------------------------------------------------------------------
// Saiga::EpipolarDistanceSquared (saiga/vision/reconstruction/Epipolar.cpp:47-56). F * (x, y, 1) as Eigen evaluates a
// fixed 3x3 * 3x1 product with its default SSE2 packets: rows 0 and 1 accumulate left to right, the scalar tail row
// sums as t0 + (t1 + t2) (checked against the compiled reference function in tests/test_mapping_cpu.py).
double epipolar_distance_squared(const double* p1, const double* p2, const double* F /* row-major */)
{
    const double l0 = (F[0] * p1[0] + F[1] * p1[1]) + F[2] * 1.0;
    const double l1 = (F[3] * p1[0] + F[4] * p1[1]) + F[5] * 1.0;
    const double l2 = F[6] * p1[0] + (F[7] * p1[1] + F[8] * 1.0);
    const double d       = (p2[0] * l0 + p2[1] * l1) + 1.0 * l2;
    const double lengSqr = l0 * l0 + l1 * l1;
    return d * d / lengSqr;
}

// MappingORBMatcher::Fuse, LocalMap<FusionPoint> overload (Snake/LocalMapping/MappingORBMatcher.cpp:359-482)
int fuse_search(const snk_frame_view* kf, const uint8_t* point_mask, const snk_point_fusion* pts, int n,
                const snk_camera* K, const snk_scale_pyramid* sp, float th, float obs_factor, int feature_th,
                int32_t* cand_kp, int32_t* cand_id, long long* pair_ops)
{
    const float th_squared = th * th;
    const V3 cam           = camera_position(kf->pose_q, kf->pose_t);
    const double logsf     = std::log(sp->scale_factor);
    int fused              = 0;
    long long ops          = 0;
    for (int i = 0; i < n; ++i)
    {
        const snk_point_fusion& lmp = pts[i];
        if (point_mask && !point_mask[i]) continue;
        V3 wp{lmp.pos[0], lmp.pos[1], lmp.pos[2]};
        V3 np = se3_apply(kf->pose_q, kf->pose_t, wp);
        if (np.z <= 0) continue;
        double u, v;
        project(K, np, u, v);
        if (!in_image(kf->grid, u, v)) continue;
        V3 PO{cam.x - wp.x, cam.y - wp.y, cam.z - wp.z};
        const double dist = std::sqrt(PO.x * PO.x + PO.y * PO.y + PO.z * PO.z);
        // EstimateMinMaxDistance (ScalePyramid.h:26-41)
        double max_d = (double)lmp.ref_depth * sp->scale[lmp.ref_level];
        double min_d = max_d / sp->scale[sp->nlevels - 1];
        min_d /= sp->scale_factor;
        max_d *= sp->scale_factor;
        if (dist < min_d || dist > max_d) continue;
        if (PO.x * lmp.normal[0] + PO.y * lmp.normal[1] + PO.z * lmp.normal[2] < 0.5 * dist) continue;
        float observationFactor = 1.0;
        if (lmp.observations <= 2) observationFactor = obs_factor;
        const float radius = observationFactor * th;
        // PredictScaleLevel (ScalePyramid.h:44-56)
        const double pmax       = (double)lmp.ref_depth * sp->scale[lmp.ref_level];
        const double prediction = std::log(pmax / dist) / logsf;
        const double r = radius, r2 = r * r;
        int bestDist = 256, bestIdx = -1;
        bool any = false;
        // the stereo prediction (projectStereo, Intrinsics4.h:221-230) does not depend on the candidate
        const double invz = 1.0 / np.z;
        const double sx = K->fx * (np.x * invz) + K->cx, sy = K->fy * (np.y * invz) + K->cy;
        const double sr = sx - K->bf * invz;
        for_candidates(kf, u, v, radius, r2, [&](int pid, bool inside) {
            const snk_keypoint_f64& kp = kf->kp_un[pid];
            const double pe            = prediction - kp.octave;  // PredictionConsistent (ScalePyramid.h:58-64)
            if (!(pe * pe < 1.2 * 1.2)) return;
            if (!inside) return;
            any = true;
            double e2;
            if (kf->depth[pid] > 0)
            {
                const double d0 = sx - kp.x, d1 = sy - kp.y, d2 = sr - (double)kf->right_points[pid];
                e2 = d0 * d0 + d1 * d1 + d2 * d2;
            }
            else
            {
                const double d0 = u - kp.x, d1 = v - kp.y;
                e2 = d0 * d0 + d1 * d1;
            }
            if (e2 > th_squared * observationFactor) return;
            ++ops;
            const int d = hamming(lmp.desc, kf->desc + 4 * (size_t)pid);
            if (d < bestDist) bestDist = d, bestIdx = pid;
        });
        if (!any) continue;
        if (bestDist <= feature_th)
        {
            cand_kp[fused] = bestIdx;
            cand_id[fused] = lmp.id;
            ++fused;
        }
    }
    if (pair_ops) *pair_ops = ops;
    return fused;
}

// LoopORBmatcher::Fuse, the search (Snake/LoopClosing/LoopORBMatcher.cpp:13-72)
void fuse_search_loop(const snk_frame_view* kf, const uint8_t* skip, const snk_point_fusion* pts, int n,
                      const snk_camera* K, float th, int32_t* best_idx, int32_t* best_dist)
{
    const V3 cam = camera_position(kf->pose_q, kf->pose_t);
    for (int i = 0; i < n; ++i)
    {
        best_idx[i]  = -1;
        best_dist[i] = 256;
        if (skip && skip[i]) continue;
        const snk_point_fusion& mp = pts[i];
        V3 wp{mp.pos[0], mp.pos[1], mp.pos[2]};
        V3 np = se3_apply(kf->pose_q, kf->pose_t, wp);
        if (np.z <= 0) continue;
        double u, v;
        project(K, np, u, v);
        if (!in_image(kf->grid, u, v)) continue;
        V3 PO{cam.x - wp.x, cam.y - wp.y, cam.z - wp.z};
        const double dist3D = std::sqrt(PO.x * PO.x + PO.y * PO.y + PO.z * PO.z);
        if (PO.x * mp.normal[0] + PO.y * mp.normal[1] + PO.z * mp.normal[2] < 0.5 * dist3D) continue;
        const float radius = th * 1;
        const float r2     = radius * radius;  // GetFeaturesInArea(…, float r): r*r in float (Features.cpp:17)
        int bestDist = 256, bestIdx = -1;
        for_candidates(kf, u, v, radius, (double)r2, [&](int pid, bool inside) {
            if (!inside) return;
            const int d = hamming(mp.desc, kf->desc + 4 * (size_t)pid);
            if (d < bestDist) bestDist = d, bestIdx = pid;
        });
        best_idx[i]  = bestIdx;
        best_dist[i] = bestDist;
    }
}

// MappingORBMatcher::SearchForTriangulation2 (Snake/LocalMapping/MappingORBMatcher.cpp:14-101)
int search_triangulation_bow(const snk_frame_view* kf1, const double* np1, const snk_featvec_view* fv1,
                             const snk_frame_view* kf2, const double* np2, const snk_featvec_view* fv2, const double* E,
                             const snk_camera* K, float epipolarDistance, int featureDistance, int32_t* pairs,
                             long long* pair_ops)
{
    const double th_chi1 = epipolarDistance * 2 / K->fx;  // float * int in float, then / double
    const double th_chi2 = th_chi1 * th_chi1;
    int nmatches = 0, a = 0, b = 0;
    long long ops = 0;
    while (a < fv1->n_nodes && b < fv2->n_nodes)
    {
        if (fv1->node_ids[a] == fv2->node_ids[b])
        {
            for (int ia = fv1->offsets[a]; ia < fv1->offsets[a + 1]; ++ia)
            {
                const int idx1 = fv1->indices[ia];
                if (kf1->kp_to_point[idx1] >= 0) continue;
                int bestDist = 50, bestIdx2 = -1;  // TH_LOW
                for (int ib = fv2->offsets[b]; ib < fv2->offsets[b + 1]; ++ib)
                {
                    const int idx2 = fv2->indices[ib];
                    if (kf2->kp_to_point[idx2] >= 0) continue;
                    ++ops;
                    const int dist = hamming(kf1->desc + 4 * (size_t)idx1, kf2->desc + 4 * (size_t)idx2);
                    if (dist > featureDistance || dist > bestDist) continue;
                    if (epipolar_distance_squared(np1 + 2 * idx1, np2 + 2 * idx2, E) < th_chi2) bestIdx2 = idx2, bestDist = dist;
                }
                if (bestIdx2 >= 0)
                {
                    pairs[2 * nmatches]     = idx1;
                    pairs[2 * nmatches + 1] = bestIdx2;
                    ++nmatches;
                }
            }
            ++a, ++b;
        }
        else if (fv1->node_ids[a] < fv2->node_ids[b])
            a = (int)(std::lower_bound(fv1->node_ids + a, fv1->node_ids + fv1->n_nodes, fv2->node_ids[b]) - fv1->node_ids);
        else
            b = (int)(std::lower_bound(fv2->node_ids + b, fv2->node_ids + fv2->n_nodes, fv1->node_ids[a]) - fv2->node_ids);
    }
    if (pair_ops) *pair_ops = ops;
    return nmatches;
}

// MappingORBMatcher::SearchForTriangulationProject (Snake/LocalMapping/MappingORBMatcher.cpp:167-247)
int search_triangulation_project(const snk_frame_view* kf1, const double* np1, const snk_frame_view* kf2,
                                 const double* np2, const double* grid, int grid_rows, int grid_cols, const double* E12,
                                 const snk_camera* K, float epipolarDistance, int featureDistance, int32_t* pairs,
                                 long long* pair_ops)
{
    (void)grid_rows;
    const double th_chi1 = epipolarDistance / K->fx;
    const double th_chi2 = th_chi1 * th_chi1;
    // Sophus SE3::inverse(): (q^-1, q^-1 * (t * -1))
    const double qi[4] = {-kf1->pose_q[0], -kf1->pose_q[1], -kf1->pose_q[2], kf1->pose_q[3]};
    const V3 ti        = camera_position(kf1->pose_q, kf1->pose_t);
    const double tinv[3] = {ti.x, ti.y, ti.z};
    int nmatches = 0;
    long long ops = 0;
    for (int idx1 = 0; idx1 < kf1->n; ++idx1)
    {
        if (kf1->kp_to_point[idx1] >= 0) continue;
        const snk_keypoint_f64& kp1 = kf1->kp_un[idx1];
        const Cell cell = cell_clamped(kf1->grid, kp1.x, kp1.y);
        const double z  = grid[(size_t)(cell.cy / 4) * grid_cols + cell.cx / 4];
        // K.unproject(ip, z) (Intrinsics4.h:70-74)
        V3 p{(kp1.x - K->cx) / K->fx, (kp1.y - K->cy) / K->fy, 1.0};
        p = V3{p.x * z, p.y * z, p.z * z};
        const V3 wp = se3_apply(qi, tinv, p);
        const V3 c2 = se3_apply(kf2->pose_q, kf2->pose_t, wp);
        double u, v;
        project(K, c2, u, v);
        if (!in_image(kf1->grid, u, v)) continue;
        int bestDist = 50, bestIdx2 = -1;
        for_candidates(kf2, u, v, 20.f, 400.0, [&](int idx2, bool inside) {
            if (!inside) return;
            if (kf2->kp_to_point[idx2] >= 0) return;
            if (epipolar_distance_squared(np1 + 2 * idx1, np2 + 2 * idx2, E12) > th_chi2) return;
            ++ops;
            const int dist = hamming(kf1->desc + 4 * (size_t)idx1, kf2->desc + 4 * (size_t)idx2);
            if (dist > featureDistance || dist > bestDist) return;
            bestIdx2 = idx2, bestDist = dist;
        });
        if (bestIdx2 >= 0)
        {
            pairs[2 * nmatches]     = idx1;
            pairs[2 * nmatches + 1] = bestIdx2;
            ++nmatches;
        }
    }
    if (pair_ops) *pair_ops = ops;
    return nmatches;
}

// MonoInitializer::ComputeMatchesPrediction (Snake/Tracking/Initialization/MonoInitializer.cpp:619-697)
int match_prediction(const uint64_t* desc_left, int n_left, double* last_positions, const snk_frame_view* right,
                     float radius, int featureError, int32_t* matches)
{
    int count = 0;
    const float r2 = radius * radius;
    for (int i = 0; i < n_left; ++i)
    {
        int bestDist = 256, bestIdx2 = -1;
        for_candidates(right, last_positions[2 * i], last_positions[2 * i + 1], radius, (double)r2, [&](int i2, bool inside) {
            if (!inside) return;
            const int d = hamming(desc_left + 4 * (size_t)i, right->desc + 4 * (size_t)i2);
            if (d < bestDist) bestDist = d, bestIdx2 = i2;
        });
        if (bestDist < featureError)
        {
            last_positions[2 * i]     = right->kp_un[bestIdx2].x;
            last_positions[2 * i + 1] = right->kp_un[bestIdx2].y;
            matches[i]                = bestIdx2;
            ++count;
        }
        else
            matches[i] = -1;
    }
    return count;
}

// Features::GetFeaturesInArea (Snake/Map/Features.cpp:13-30) as every search above enumerates its candidates
int features_in_area(const snk_frame_view* f, double px, double py, float r, int32_t* out)
{
    int c = 0;
    const float r2 = r * r;
    for_candidates(f, px, py, r, (double)r2, [&](int pid, bool inside) {
        if (inside) out[c++] = pid;
    });
    return c;
}

// FeatureGridBounds2<double,20>::computeFromIntrinsicsDist (saiga/vision/util/FeatureGrid2.h:63-104) with
// undistortAll (Distortion.h:326-338): the undistorted image border decides the grid the searches run on
void grid_bounds(int w, int h, const double* K /* fx fy cx cy s */, const double* dist, int resolution, snk_grid_view* out)
{
    std::vector<double> px, py;
    for (int i = 0; i < resolution + 1; ++i)
    {
        const float alpha = float(i) / resolution * w;
        px.push_back(alpha), py.push_back(0);
        px.push_back(alpha), py.push_back(h - 1);
        const float beta = float(i) / resolution * h;
        px.push_back(0), py.push_back(beta);
        px.push_back(w - 1), py.push_back(beta);
    }
    double bminx = 0, bminy = 0, bmaxx = 0, bmaxy = 0;
    for (size_t k = 0; k < px.size(); ++k)
    {
        const double nx0 = (px[k] - K[2]) / K[0], ny0 = (py[k] - K[3]) / K[1];  // unproject2
        double nx, ny;
        undistort_gn(dist, nx0, ny0, nx, ny);
        const double ux = K[0] * nx + K[4] * ny + K[2], uy = K[1] * ny + K[3];  // normalizedToImage
        if (k == 0) bminx = bmaxx = ux, bminy = bmaxy = uy;
        bminx = std::min(ux, bminx), bminy = std::min(uy, bminy);
        bmaxx = std::max(ux, bmaxx), bmaxy = std::max(uy, bmaxy);
    }
    bminx -= 0.1, bminy -= 0.1, bmaxx += 0.1, bmaxy += 0.1;
    out->bmin_x = bminx, out->bmin_y = bminy, out->bmax_x = bmaxx, out->bmax_y = bmaxy;
    out->cell_size_inv_x = out->cell_size_inv_y = 1.0 / 20.0;
    out->rows  = (int)std::ceil((bmaxy - bminy) / 20.0);
    out->cols  = (int)std::ceil((bmaxx - bminx) / 20.0);
    out->cells = nullptr;
}
}  // namespace orbref
