// Hamming matching kernels. A descriptor is 256 bits = 8 x u32 held in registers; one distance is 8 LOP3 (xor) +
// 8 POPC + 7 IADD. Every matcher reproduces the reference's tie-breaking: strict '<' keeps the first candidate in
// the reference's iteration order.
#pragma once
#include "common.cuh"

namespace snk
{
struct Desc
{
    uint32_t w[8];
};

__device__ __forceinline__ Desc load_desc(const uint64_t* __restrict__ p)
{
    const uint4* q = reinterpret_cast<const uint4*>(p);
    const uint4 a = q[0], b = q[1];
    return Desc{{a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w}};
}
__device__ __forceinline__ int hamming(const Desc& a, const Desc& b)
{
    int d = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) d += __popc(a.w[i] ^ b.w[i]);
    return d;
}

// Saiga::distance for n independent pairs (Features.h:124-138)
__global__ void k_hamming_pairs(const uint64_t* __restrict__ a, const uint64_t* __restrict__ b, int n, int* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = hamming(load_desc(a + 4 * (size_t)i), load_desc(b + 4 * (size_t)i));
}

// --------------------------------------------------------------------------------------------------------------
// BruteForceMatcher::matchKnn2 (Features.h:323-354). One thread per query keeps the query in registers and walks the
// train set in ascending j (tiles staged in shared memory, read as broadcasts), so strict '<' picks exactly the
// reference's first minimum and its second-best.
// --------------------------------------------------------------------------------------------------------------
constexpr int KNN_NT = 128, KNN_TILE = 128;

__global__ void __launch_bounds__(KNN_NT) k_knn2(const uint64_t* __restrict__ d1, const int* __restrict__ n_arr, int n_fixed,
                                                 int stride1, const uint64_t* __restrict__ d2,
                                                 const int* __restrict__ m_arr, int m_fixed, int stride2,
                                                 snk_knn2* __restrict__ out)
{
    __shared__ uint4 tile[KNN_TILE][2];
    const int prob = blockIdx.y;
    const int n    = n_arr ? n_arr[prob] : n_fixed;
    const int m    = m_arr ? m_arr[prob] : m_fixed;
    const int i    = blockIdx.x * KNN_NT + threadIdx.x;
    if (blockIdx.x * KNN_NT >= n) return;
    const uint64_t* q1 = d1 + 4 * (size_t)prob * stride1;
    const uint64_t* q2 = d2 + 4 * (size_t)prob * stride2;
    Desc q{};
    if (i < n) q = load_desc(q1 + 4 * (size_t)i);
    int b0 = 1000, i0 = -1, b1 = 1000, i1 = -1;
    for (int base = 0; base < m; base += KNN_TILE)
    {
        const int cnt = min(KNN_TILE, m - base);
        __syncthreads();
        for (int k = threadIdx.x; k < 2 * cnt; k += KNN_NT)
            tile[k >> 1][k & 1] = reinterpret_cast<const uint4*>(q2 + 4 * (size_t)base)[k];
        __syncthreads();
        for (int j = 0; j < cnt; ++j)
        {
            const uint4 a = tile[j][0], b = tile[j][1];
            const int d = __popc(q.w[0] ^ a.x) + __popc(q.w[1] ^ a.y) + __popc(q.w[2] ^ a.z) + __popc(q.w[3] ^ a.w) +
                          __popc(q.w[4] ^ b.x) + __popc(q.w[5] ^ b.y) + __popc(q.w[6] ^ b.z) + __popc(q.w[7] ^ b.w);
            if (d < b0)
            {
                b1 = b0, i1 = i0;
                b0 = d, i0 = base + j;
            }
            else if (d < b1)
            {
                b1 = d, i1 = base + j;
            }
        }
    }
    if (i < n) out[(size_t)prob * stride1 + i] = snk_knn2{b0, i0, b1, i1};
}

// --------------------------------------------------------------------------------------------------------------
// The same matchKnn2 for ONE problem that would not fill the machine with a thread per query (1000 x 1000 is 8 blocks
// of k_knn2 on 148 SMs): the train set is split across warps instead. A warp keeps 32 train descriptors in registers
// (one per lane) and walks a group of queries; per query every lane computes one distance and the warp reduces to the
// two smallest keys with two REDUX min operations. A key is distance << 22 | train index, so the unsigned minimum is
// the reference's choice: smallest distance, and among equals the first in train order (strict '<', Features.h:335-348);
// the second key is the minimum over the remaining lanes — exactly the (best, second best) pair the serial loop leaves.
// The partial pairs of all train chunks are merged by k_knn2_merge in the same key order.
// --------------------------------------------------------------------------------------------------------------
constexpr int KNS_WARPS = 4;                  // train chunks (of 32) per block
constexpr unsigned KNS_NONE = 0xFFFFFFFFu;    // "no candidate" key
constexpr int KNS_MAX_M = 1 << 22;            // train indices must fit the key

__global__ void __launch_bounds__(KNS_WARPS * 32) k_knn2_split(const uint64_t* __restrict__ d1, int n, const uint64_t* __restrict__ d2,
                                                               int m, int qgroup, uint2* __restrict__ partial)
{
    extern __shared__ uint4 kns_q[];  // the block's query group, 2 x uint4 per descriptor
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int q0 = blockIdx.y * qgroup, nq = min(qgroup, n - q0);
    const int chunk = blockIdx.x * KNS_WARPS + wid;  // train descriptors [32 chunk, 32 chunk + 32)
    const int j = chunk * 32 + lane;
    for (int k = threadIdx.x; k < 2 * nq; k += KNS_WARPS * 32) kns_q[k] = reinterpret_cast<const uint4*>(d1 + 4 * (size_t)q0)[k];
    Desc t{};
    if (j < m) t = load_desc(d2 + 4 * (size_t)j);
    const unsigned jkey = (j < m) ? (unsigned)j : KNS_NONE;
    __syncthreads();
    if (chunk * 32 >= m) return;
    uint2* out = partial + (size_t)chunk * n + q0;
    unsigned r1 = KNS_NONE, r2 = KNS_NONE;
    for (int q = 0; q < nq; ++q)
    {
        const uint4 a = kns_q[2 * q], b = kns_q[2 * q + 1];
        const int d = __popc(t.w[0] ^ a.x) + __popc(t.w[1] ^ a.y) + __popc(t.w[2] ^ a.z) + __popc(t.w[3] ^ a.w) +
                      __popc(t.w[4] ^ b.x) + __popc(t.w[5] ^ b.y) + __popc(t.w[6] ^ b.z) + __popc(t.w[7] ^ b.w);
        const unsigned key = (j < m) ? (((unsigned)d << 22) | jkey) : KNS_NONE;
        const unsigned k1 = __reduce_min_sync(0xffffffffu, key);
        const unsigned k2 = __reduce_min_sync(0xffffffffu, key == k1 ? KNS_NONE : key);  // keys of valid lanes are distinct
        if (lane == (q & 31)) r1 = k1, r2 = k2;
        if ((q & 31) == 31 || q == nq - 1)
        {
            const int qq = (q & ~31) + lane;  // lane l holds query (q & ~31) + l: one coalesced store per 32 queries
            if (qq <= q) out[qq] = make_uint2(r1, r2);
        }
    }
}

__global__ void k_knn2_merge(const uint2* __restrict__ partial, int n, int n_chunks, snk_knn2* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned a = KNS_NONE, b = KNS_NONE;  // the two smallest keys so far, a < b (or both "none")
    for (int c = 0; c < n_chunks; ++c)
    {
        const uint2 p = partial[(size_t)c * n + i];  // p.x < p.y
        if (p.x < a)
        {
            b = min(a, p.y);
            a = p.x;
        }
        else
            b = min(b, p.x);
    }
    snk_knn2 r;
    r.dist0 = (a == KNS_NONE) ? 1000 : (int)(a >> 22);
    r.idx0  = (a == KNS_NONE) ? -1 : (int)(a & 0x3fffffu);
    r.dist1 = (b == KNS_NONE) ? 1000 : (int)(b >> 22);
    r.idx1  = (b == KNS_NONE) ? -1 : (int)(b & 0x3fffffu);
    out[i]  = r;
}

// --------------------------------------------------------------------------------------------------------------
// Preprocess::StereoMatching (Snake/Preprocess/Preprocess.cpp:152-239). One CTA per stereo pair.
// The reference buckets right keypoints by iRound(y) and scans rows y-r..y+r, each bucket in ascending index: the
// candidate order is (row, index). Here the right keypoints are sorted by that key in shared memory (bitonic), each
// left keypoint binary-searches its first row and walks the sorted run — same candidates, same order.
// --------------------------------------------------------------------------------------------------------------
constexpr int ST_NT = 256;
constexpr int ST_ROW_BINS = 4096;  // image rows the counting sort of k_stereo covers (levels are at most 4000 rows)

__device__ __forceinline__ int iround_f(float v) { return (int)__fadd_rn(v, v >= 0.f ? 0.5f : -0.5f); }

template <typename KP>
__global__ void __launch_bounds__(ST_NT) k_stereo(const KP* __restrict__ kps, const uint64_t* __restrict__ desc,
                                                  const int* __restrict__ counts, int stride, int nl_fixed, int nr_fixed,
                                                  snk_stereo_params prm, int sort_cap, float* __restrict__ right_points,
                                                  float* __restrict__ depth, int* __restrict__ n_matches,
                                                  int* __restrict__ err, int nbins, int staged)
{
    // nbins: rows the counting sort covers, a multiple of ST_NT, <= ST_ROW_BINS (the image height rounded up when the
    // caller knows it: clearing, scanning and ordering 4096 rows for a 480-row image was most of the sort)
    extern __shared__ uint32_t keys[];  // sort_cap entries: (iRound(y) + 32768) << 16 | index, then ST_ROW_BINS + 1 row counters
    __shared__ int s_matches;
    const int pair = blockIdx.x;
    // layout: left image = entry 2*pair, right image = entry 2*pair+1 of [..][stride] arrays
    const KP* L           = kps + (size_t)(2 * pair) * stride;
    const KP* R           = kps + (size_t)(2 * pair + 1) * stride;
    const uint64_t* DL    = desc + 4 * (size_t)(2 * pair) * stride;
    const uint64_t* DR    = desc + 4 * (size_t)(2 * pair + 1) * stride;
    const int nl          = counts ? counts[2 * pair] : nl_fixed;
    const int nr          = counts ? counts[2 * pair + 1] : nr_fixed;
    float* rp             = right_points + (size_t)pair * stride;
    float* dp             = depth + (size_t)pair * stride;
    const int tid         = threadIdx.x;
    // gridDim.y blocks share one pair: block `part` owns the left keypoints of every gridDim.y-th group of ST_NT (each
    // block sorts the right image's rows for itself — a few microseconds — so the blocks stay independent)
    const int part = blockIdx.y, step = ST_NT * gridDim.y;
    if (tid == 0) s_matches = 0;
    for (int i = part * ST_NT + tid; i < nl; i += step) rp[i] = dp[i] = -1000.f;  // Frame::allocateTmp (Snake/Map/Frame.cpp:25-26)
    if (nr > sort_cap)
    {
        if (tid == 0) atomicExch(err, 1);
        return;
    }
    // ---- right keypoints ordered by (iRound(y), index) ----
    // Rows inside [0, ST_ROW_BINS): counting sort (histogram, scan, scatter, then every row orders its handful of
    // entries by index) — a few passes over the keys. Any row outside that range: the general bitonic sort.
    int* bins = reinterpret_cast<int*>(keys + sort_cap);  // ST_ROW_BINS + 1 counters
    __shared__ int s_wide;
    if (tid == 0) s_wide = 0;
    for (int i = tid; i <= nbins; i += ST_NT) bins[i] = 0;
    __syncthreads();
    for (int i = tid; i < nr; i += ST_NT)
    {
        const int row = iround_f((float)R[i].y);
        if (row < -32768 || row > 32766) atomicExch(err, 2);
        if (row < 0 || row >= nbins)
            s_wide = 1;
        else
            atomicAdd(&bins[row], 1);
    }
    __syncthreads();
    if (!s_wide)
    {
        // exclusive scan of the row counts: every thread owns nbins / ST_NT consecutive bins
        const int PER = nbins / ST_NT;
        int sum = 0;
        for (int k = 0; k < PER; ++k) sum += bins[tid * PER + k];
        const int lane = tid & 31, wid = tid >> 5;
        const int inc  = warp_incl_scan(sum, lane);
        __shared__ int s_wsum[ST_NT / 32];
        if (lane == 31) s_wsum[wid] = inc;
        __syncthreads();
        int base = inc - sum;
#pragma unroll
        for (int k = 0; k < ST_NT / 32; ++k)
            if (k < wid) base += s_wsum[k];
        // bins[r] = first slot of row r; it doubles as the scatter cursor and ends up at the first slot of row r + 1
        for (int k = 0; k < PER; ++k)
        {
            const int cnt       = bins[tid * PER + k];
            bins[tid * PER + k] = base;
            base += cnt;
        }
        if (tid == ST_NT - 1) bins[nbins] = base;
        __syncthreads();
        // scatter (order inside a row arbitrary for now); the cursor of row r ends at the start of row r + 1
        for (int i = tid; i < nr; i += ST_NT)
        {
            const int row = iround_f((float)R[i].y);
            const int pos = atomicAdd(&bins[row], 1);
            keys[pos]     = ((uint32_t)(row + 32768) << 16) | (uint32_t)i;
        }
        __syncthreads();
        // row r now occupies [end of row r-1, bins[r]): order its entries by index (insertion sort, rows are short)
        for (int r = tid; r < nbins; r += ST_NT)
        {
            const int e = bins[r], b0 = (r == 0) ? 0 : bins[r - 1];
            for (int i = b0 + 1; i < e; ++i)
            {
                const uint32_t v = keys[i];
                int j = i - 1;
                while (j >= b0 && keys[j] > v)
                {
                    keys[j + 1] = keys[j];
                    --j;
                }
                keys[j + 1] = v;
            }
        }
        __syncthreads();
    }
    else
    {
        int np2 = 1;
        while (np2 < nr) np2 <<= 1;
        for (int i = tid; i < np2; i += ST_NT)
        {
            uint32_t k = 0xffffffffu;
            if (i < nr) k = ((uint32_t)(iround_f((float)R[i].y) + 32768) << 16) | (uint32_t)i;
            keys[i] = k;
        }
        __syncthreads();
        for (int k = 2; k <= np2; k <<= 1)
            for (int j = k >> 1; j > 0; j >>= 1)
            {
                for (int i = tid; i < np2; i += ST_NT)
                {
                    const int ixj = i ^ j;
                    if (ixj > i)
                    {
                        const uint32_t a = keys[i], b = keys[ixj];
                        const bool up = (i & k) == 0;
                        if ((a > b) == up)
                        {
                            keys[i]   = b;
                            keys[ixj] = a;
                        }
                    }
                }
                __syncthreads();
            }
    }

    // staged != 0: x and octave of the right keypoints in sorted order behind the counters, so that the candidate walk
    // below touches shared memory only (its loop is a serial chain per thread; global loads in it cost a round trip each)
    using XT = decltype(KP::x);
    XT* sx   = reinterpret_cast<XT*>(reinterpret_cast<unsigned char*>(keys) + (((size_t)(sort_cap + nbins + 4) * 4 + 7) & ~(size_t)7));
    int* so  = reinterpret_cast<int*>(sx + sort_cap);
    if (staged)
    {
        for (int k = tid; k < nr; k += ST_NT)
        {
            const int o = (int)(keys[k] & 0xffffu);
            sx[k] = R[o].x;
            so[k] = R[o].octave;
        }
        __syncthreads();
    }

    const float min_disp = 0.f;
    const float max_disp = (float)(prm.bf * 0.5);
    const int th_dist    = prm.relaxed ? 75 : 40;
    const double ratio   = prm.relaxed ? 0.9 : 0.7;
    const float th_rot   = prm.relaxed ? 25.f : 5.f;
    int local = 0;
    for (int i = part * ST_NT + tid; i < nl; i += step)
    {
        const double lx = (double)L[i].x;
        const int oct   = L[i].octave;
        const int y     = iround_f((float)L[i].y);
        const float r   = (float)ceil(2.0f * prm.scale[oct]);  // float * double -> double, ceil, narrowed to float
        const int row_lo = y - (int)r, row_hi = y + (int)r;
        // lower_bound of (row_lo, 0)
        const uint32_t klo = (uint32_t)(max(row_lo, -32768) + 32768) << 16;
        int lo = 0, hi = nr;
        while (lo < hi)
        {
            const int mid = (lo + hi) >> 1;
            if (keys[mid] < klo)
                lo = mid + 1;
            else
                hi = mid;
        }
        const Desc q = load_desc(DL + 4 * (size_t)i);
        int best_id = -1, best = 250, second = 250;
        for (int k = lo; k < nr; ++k)
        {
            const uint32_t key = keys[k];
            if ((int)(key >> 16) - 32768 > row_hi) break;
            const int o = (int)(key & 0xffffu);
            const double disparity = lx - (double)(staged ? sx[k] : R[o].x);
            if (disparity < min_disp || disparity > max_disp) continue;
            if (abs(oct - (staged ? so[k] : R[o].octave)) > 1) continue;
            const int d = hamming(q, load_desc(DR + 4 * (size_t)o));
            if (d < best)
            {
                second  = best;
                best    = d;
                best_id = o;
            }
            else if (d < second)
                second = d;
        }
        if (best > th_dist) continue;
        if ((double)best > ratio * (double)second) continue;
        const double a1 = (double)L[i].angle, a2 = (double)R[best_id].angle;
        // the reference really adds 365 (sic), Preprocess.cpp:216-217
        const float rot = (float)fmin(fabs(a1 - a2), fmin(fabs((a1 + 365) - a2), fabs(a1 - (a2 + 365))));
        if (rot > th_rot) continue;
        double right_point = (double)R[best_id].x;
        double disparity   = lx - right_point;
        if (disparity <= 0.001)
        {
            disparity   = 0.001;
            right_point = lx - disparity;
        }
        rp[i] = (float)right_point;
        dp[i] = (float)(prm.bf / disparity);
        ++local;
    }
    local = warp_sum(local);
    if ((tid & 31) == 0 && local) atomicAdd(&s_matches, local);
    __syncthreads();
    if (tid == 0 && s_matches) atomicAdd(&n_matches[pair], s_matches);  // n_matches is zeroed by the launcher
}

// --------------------------------------------------------------------------------------------------------------
// Preprocess::undistortKeypoints (Snake/Preprocess/Preprocess.cpp:55-77), one thread per keypoint. IEEE double in the
// reference's operation order (-fmad=false): 5 Gauss-Newton steps of undistortPointGN (Distortion.h:241-285) with
// the 2x2 pivoted LDLT solve Eigen performs (Eigen/src/Cholesky/LDLT.h:300-400), quaternion rotation, K_dst.
// --------------------------------------------------------------------------------------------------------------
struct DistortEval
{
    double xd, yd, j00, j01, j10, j11;
};
template <bool WITH_J>
__host__ __device__ __forceinline__ DistortEval distort_point(const double* __restrict__ D, double x, double y)
{
    DistortEval e;
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
    e.xd = x * radial + tx;
    e.yd = y * radial + ty;
    if (r2 > 100000.0 * 100000.0) e.xd = e.yd = 100000;
    if (WITH_J)
    {
        const double cx3 = 4 * x * x * x + 4 * x * y * y, cy3 = 4 * y * y * y + 4 * y * x * x;
        const double cx5 = 6 * x2 * x2 * x + 12 * x2 * x * y2 + 6 * x * y2 * y2;
        const double cy5 = 6 * y2 * y2 * y + 12 * y2 * y * x2 + 6 * y * x2 * x2;
        double u0 = k1 * (2 * x), u1 = k1 * (2 * y);
        u0 += k2 * cx3;
        u1 += k2 * cy3;
        u0 += k3 * cx5;
        u1 += k3 * cy5;
        double v0 = k4 * (2 * x), v1 = k4 * (2 * y);
        v0 += k5 * cx3;
        v1 += k5 * cy3;
        v0 += k6 * cx5;
        v1 += k6 * cy5;
        const double jr0 = (u0 * radial_v - v0 * radial_u) * radial_iv * radial_iv;
        const double jr1 = (u1 * radial_v - v1 * radial_u) * radial_iv * radial_iv;
        e.j00 = (x * jr0 + radial) + (2 * p1 * y + 6 * p2 * x);
        e.j01 = (x * jr1) + (2 * p1 * x + 2 * p2 * y);
        e.j10 = (y * jr0) + (2 * p2 * y + 2 * p1 * x);
        e.j11 = (y * jr1 + radial) + (2 * p2 * x + 6 * p1 * y);
    }
    return e;
}

// undistortPointGN (Distortion.h:241-285) with the point as its own start value. Host and device: the host uses it
// for the feature-grid bounds (snk_grid_bounds), the device for the keypoints.
__host__ __device__ inline void undistort_gn_hd(const double* __restrict__ D, double px, double py, double& ox, double& oy)
{
    double x = px, y = py, lx = px, ly = py;
    double last_chi2 = (double)INFINITY;
    for (int it = 0; it < 5; ++it)
    {
        const DistortEval e = distort_point<true>(D, x, y);
        const double r0 = e.xd - px, r1 = e.yd - py;
        const double chi2 = r0 * r0 + r1 * r1;
        if (chi2 > last_chi2)
        {
            x = lx;
            y = ly;
            continue;
        }
        // JtJ (lower part) and -Jt res
        const double a = e.j00 * e.j00 + e.j10 * e.j10, b = e.j01 * e.j00 + e.j11 * e.j10,
                     c = e.j01 * e.j01 + e.j11 * e.j11;
        const double g0 = (-e.j00) * r0 + (-e.j10) * r1, g1 = (-e.j01) * r0 + (-e.j11) * r1;
        lx        = x;
        ly        = y;
        last_chi2 = chi2;
        // LDLT with the larger diagonal entry as first pivot (the first one on ties)
        const bool swp  = !(fabs(a) >= fabs(c));
        const double d0 = swp ? c : a, rest = swp ? a : c;
        double s0 = 0, s1 = 0;
        if (fabs(d0) > 0)
        {
            const double l  = b / d0;
            const double d1 = rest - l * (d0 * l);
            double y0 = swp ? g1 : g0, y1 = swp ? g0 : g1;
            y1 -= l * y0;
            const double tol = 2.2250738585072014e-308;  // numeric_limits<double>::min(): Eigen's pivot cut-off
            y0 = (fabs(d0) > tol) ? y0 / d0 : 0.0;
            y1 = (fabs(d1) > tol) ? y1 / d1 : 0.0;
            y0 -= l * y1;
            s0 = swp ? y1 : y0;
            s1 = swp ? y0 : y1;
        }
        x += s0;
        y += s1;
    }
    {
        const DistortEval e = distort_point<false>(D, x, y);
        const double r0 = e.xd - px, r1 = e.yd - py;
        if (r0 * r0 + r1 * r1 > last_chi2)
        {
            x = lx;
            y = ly;
        }
    }
    ox = x;
    oy = y;
}

__global__ void k_undistort(snk_rectify r, const snk_keypoint_f64* __restrict__ kps, int n,
                            snk_keypoint_f64* __restrict__ kp_un, double* __restrict__ normalized)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    snk_keypoint_f64 k = kps[i];
    const double px = (k.x - r.cx) / r.fx, py = (k.y - r.cy) / r.fy;
    double x, y;
    undistort_gn_hd(r.dist, px, py, x, y);
    // Quaternion * Vector3 as Eigen evaluates it: uv = 2 (q.vec x v); v + w uv + q.vec x uv
    const double qx = r.quat[0], qy = r.quat[1], qz = r.quat[2], qw = r.quat[3];
    double u0 = qy * 1.0 - qz * y, u1 = qz * x - qx * 1.0, u2 = qx * y - qy * x;
    u0 += u0;
    u1 += u1;
    u2 += u2;
    const double c0 = qy * u2 - qz * u1, c1 = qz * u0 - qx * u2, c2 = qx * u1 - qy * u0;
    const double p0 = (x + qw * u0) + c0, p1 = (y + qw * u1) + c1, p2 = (1.0 + qw * u2) + c2;
    const double mx = p0 / p2, my = p1 / p2;
    normalized[2 * i]     = mx;
    normalized[2 * i + 1] = my;
    k.x      = r.dfx * mx + r.ds * my + r.dcx;
    k.y      = r.dfy * my + r.dcy;
    kp_un[i] = k;
}

// --------------------------------------------------------------------------------------------------------------
// Preprocess::ComputeStereoFromRGBD (Snake/Preprocess/Preprocess.cpp:79-120), one thread per keypoint, double math in
// the reference's operation order (library is built with -fmad=false)
// --------------------------------------------------------------------------------------------------------------
__global__ void k_stereo_rgbd(snk_rgbd_params p, const snk_keypoint_f64* __restrict__ kp, int n,
                              const float* __restrict__ depth_img, int rows, int cols, int pitch_f,
                              float* __restrict__ right_points, float* __restrict__ depth, int* __restrict__ n_matches)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    int hit     = 0;
    if (i < n)
    {
        // K.unproject2
        const double x = (kp[i].x - p.cx) / p.fx, y = (kp[i].y - p.cy) / p.fy;
        // distortNormalizedPoint (Distortion.h:128-170)
        const double k1 = p.dist[0], k2 = p.dist[1], k3 = p.dist[2], k4 = p.dist[3], k5 = p.dist[4], k6 = p.dist[5];
        const double p1 = p.dist[6], p2 = p.dist[7];
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
        // depthModel.K.normalizedToImage
        const double u = p.dfx * xd + p.ds * yd + p.dcx, v = p.dfy * yd + p.dcy;
        const int px = (int)(u + 0.5), py = (int)(v + 0.5);
        float d = 0.f;
        if (px >= 0 && px < cols && py >= 0 && py < rows) d = depth_img[(size_t)py * pitch_f + px];
        if (d > 0)
        {
            depth[i]        = d;
            right_points[i] = (float)(kp[i].x - p.bf / (double)d);
            hit             = 1;
        }
        else
        {
            depth[i]        = -1.f;
            right_points[i] = -1.f;
        }
    }
    hit = warp_sum(hit);
    if ((threadIdx.x & 31) == 0 && hit) atomicAdd(n_matches, hit);
}

// --------------------------------------------------------------------------------------------------------------
// popc issue-rate probe: 256-bit distances between register-resident operands (denominator of the matcher roofline)
// --------------------------------------------------------------------------------------------------------------
__global__ void k_popc_peak(int iters, unsigned seed, int* __restrict__ sink)
{
    uint32_t a[8], b[8];
#pragma unroll
    for (int k = 0; k < 8; ++k)
    {
        a[k] = seed * (2654435761u + k) + threadIdx.x * 40503u + blockIdx.x;
        b[k] = a[k] ^ (0x9e3779b9u * (k + 1));
    }
    int acc = 0;
    for (int it = 0; it < iters; ++it)
    {
#pragma unroll
        for (int k = 0; k < 8; ++k)
        {
            b[k] ^= b[(k + 1) & 7] + it;  // new operand every iteration, stays in registers
            acc += __popc(a[k] ^ b[k]);
        }
    }
    if (acc == 0x7fffffff) sink[0] = acc;
}
}  // namespace snk
