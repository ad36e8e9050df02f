// Device side of the ORB extractor: data layout descriptors and the kernels.
//
//   k_resize      level l from level l-1        cv::resize INTER_LINEAR fixed point    ORBExtractor.cpp:282
//   k_fast        FAST-9/16 + NMS per cell      cv::FAST per cell, two thresholds      ORBExtractor.cpp:114-157
//   k_distribute  quadtree selection per level  QuadtreeFeatureDistributor::Distribute FeatureDistribution.cpp:72-189
//   k_blur        7x7 Gaussian, reflect-101     cv::GaussianBlur                       ORBExtractor.cpp:213
//   k_describe    orientation + rBRIEF + pack   ORB::ComputeAngle / ComputeDescriptor  OrbDescriptors.cpp:26-125,
//                                               and the Detect tail                    ORBExtractor.cpp:163-171,223-232
//
// All kernels are batched: blockIdx.y (or .z) is the image of the batch; images never interact.
#pragma once
#include "common.cuh"
#include "glibc_math.cuh"
#include "gnu_sort.cuh"

#define SNK_MAX_LEVELS 16

namespace snk
{
struct LevelDev
{
    uint8_t* img;          // level image of batch entry 0 (level 0 may alias the caller's device buffer)
    uint8_t* gauss;        // blurred copy
    size_t img_stride;     // bytes between consecutive batch entries
    size_t gauss_stride;
    int pitch, gauss_pitch;
    int w, h;
    int quota;             // ScalePyramid::Features(level)
    int cell_base;         // first cell slot of this level inside an image's cell list
    int n_cells;
    int cand_base;         // first entry of this level inside an image's work arrays (n_cells * cell_cap entries)
    int sel_base;          // first entry of this level inside an image's selection array (quota + 4 entries)
    float scale;           // (float)ScalePyramid::Scale(level)
    float size;            // (float)(int)(31 * Scale(level))
    int _pad;
};

struct PyrDev
{
    LevelDev lv[SNK_MAX_LEVELS];
    int nlevels;
    int cells_per_img;  // number of FAST cells over all levels
    int cell_cap;       // candidate slots per cell
    int cand_per_img;   // cells_per_img * cell_cap
    int sel_per_img;    // sum over levels of (quota + 4)
    int kp_stride;      // output records per image
    int th_ini, th_min;
};

// One FAST cell == one cv::FAST call of the reference: ROI origin/size inside its level and the offset the reference
// adds to the ROI-relative coordinates (ORBExtractor.cpp:152-153).
struct CellDev
{
    short level;
    short x0, y0;  // iniX, iniY
    short w, h;    // maxX-iniX, maxY-iniY
    short off_x, off_y;
    short _pad;
};

// --------------------------------------------------------------------------------------------------------------
// Pyramid
// --------------------------------------------------------------------------------------------------------------
constexpr int RS_TW = 128, RS_TH = 16, RS_SW = 168, RS_SH = 24;

__global__ void __launch_bounds__(256) k_resize(const uint8_t* __restrict__ src, size_t src_stride, int spitch, int sw,
                                                int sh, uint8_t* __restrict__ dst, size_t dst_stride, int dpitch,
                                                int dw, int dh, const int* __restrict__ xofs,
                                                const short* __restrict__ xa, const int* __restrict__ yofs,
                                                const short* __restrict__ yb)
{
    __shared__ uint8_t tile[RS_SH][RS_SW];
    const int tx0 = blockIdx.x * RS_TW, ty0 = blockIdx.y * RS_TH;
    const int tx1 = min(tx0 + RS_TW, dw) - 1, ty1 = min(ty0 + RS_TH, dh) - 1;
    const uint8_t* S = src + (size_t)blockIdx.z * src_stride;
    uint8_t* D       = dst + (size_t)blockIdx.z * dst_stride;
    // source window covered by this output tile (taps sx, sx+1 and rows sy, sy+1; indices clamp at the border)
    const int sx_lo = xofs[tx0], sx_hi = min(xofs[tx1] + 1, sw - 1);
    const int sy_lo = max(yofs[ty0], 0), sy_hi = min(yofs[ty1] + 1, sh - 1);
    const int cw = sx_hi - sx_lo + 1, ch = sy_hi - sy_lo + 1;
    for (int i = threadIdx.x; i < cw * ch; i += 256)
    {
        int r = i / cw, c = i - r * cw;
        tile[r][c] = S[(size_t)(sy_lo + r) * spitch + sx_lo + c];
    }
    __syncthreads();
    // 256 threads: 32 quads across x, 8 rows at a time
    const int qx = threadIdx.x & 31, qy = threadIdx.x >> 5;
    for (int ry = qy; ry < RS_TH; ry += 8)
    {
        const int y = ty0 + ry;
        if (y >= dh) break;
        const int r0 = min(max(yofs[y], 0), sh - 1) - sy_lo;
        const int r1 = min(max(yofs[y] + 1, 0), sh - 1) - sy_lo;
        const int b0 = yb[2 * y], b1 = yb[2 * y + 1];
        uint32_t packed = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k)
        {
            const int x = tx0 + qx * 4 + k;
            if (x < dw)
            {
                const int c0 = xofs[x] - sx_lo;
                const int c1 = min(xofs[x] + 1, sw - 1) - sx_lo;
                const int a0 = xa[2 * x], a1 = xa[2 * x + 1];
                const int h0 = tile[r0][c0] * a0 + tile[r0][c1] * a1;
                const int h1 = tile[r1][c0] * a0 + tile[r1][c1] * a1;
                int v        = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
                v            = min(max(v, 0), 255);
                packed |= (uint32_t)v << (8 * k);
            }
        }
        const int x = tx0 + qx * 4;
        if (x + 3 < dw)
            *reinterpret_cast<uint32_t*>(D + (size_t)y * dpitch + x) = packed;  // dpitch and x are multiples of 4
        else
            for (int k = 0; k < 4 && x + k < dw; ++k) D[(size_t)y * dpitch + x + k] = (uint8_t)(packed >> (8 * k));
    }
}

// --------------------------------------------------------------------------------------------------------------
// FAST
// --------------------------------------------------------------------------------------------------------------
constexpr int FAST_DIM = 66;  // wCell < 60  =>  ROI side <= 66
constexpr int FAST_PITCH = 68;
constexpr int FAST_NT = 128;

// arc score: max over the 16 arcs of 9 contiguous ring pixels of min(v - ring) and min(ring - v).
// The differences are biased by +256 so that every intermediate is a positive integer: nvcc 12.9 narrows the
// min/max network to 16-bit *unsigned* PTX ops (min.u16/max.u16), which is only correct for non-negative operands
// (the signed formulation returned wrong scores on sm_100a; tools/dev_check.cu guards this on the device).
__device__ __forceinline__ int fast_arc_score(const uint8_t (*t)[FAST_PITCH], int y, int x)
{
    const unsigned v = 256u + t[y][x];
    unsigned d[16];
    d[0]  = v - t[y + 3][x];
    d[1]  = v - t[y + 3][x + 1];
    d[2]  = v - t[y + 2][x + 2];
    d[3]  = v - t[y + 1][x + 3];
    d[4]  = v - t[y][x + 3];
    d[5]  = v - t[y - 1][x + 3];
    d[6]  = v - t[y - 2][x + 2];
    d[7]  = v - t[y - 3][x + 1];
    d[8]  = v - t[y - 3][x];
    d[9]  = v - t[y - 3][x - 1];
    d[10] = v - t[y - 2][x - 2];
    d[11] = v - t[y - 1][x - 3];
    d[12] = v - t[y][x - 3];
    d[13] = v - t[y + 1][x - 3];
    d[14] = v - t[y + 2][x - 2];
    d[15] = v - t[y + 3][x - 1];
    // sliding-window min and max over 9 consecutive (circular) entries by doubling: 2, 4, 8, then +1
    unsigned lo[16], hi[16];
#pragma unroll
    for (int k = 0; k < 16; ++k)
    {
        lo[k] = min(d[k], d[(k + 1) & 15]);
        hi[k] = max(d[k], d[(k + 1) & 15]);
    }
    unsigned lo4[16], hi4[16];
#pragma unroll
    for (int k = 0; k < 16; ++k)
    {
        lo4[k] = min(lo[k], lo[(k + 2) & 15]);
        hi4[k] = max(hi[k], hi[(k + 2) & 15]);
    }
    unsigned best_lo = 0, best_hi = 1024;  // max of window minima, min of window maxima
#pragma unroll
    for (int k = 0; k < 16; ++k)
    {
        best_lo = max(best_lo, min(min(lo4[k], lo4[(k + 4) & 15]), d[(k + 8) & 15]));
        best_hi = min(best_hi, max(max(hi4[k], hi4[(k + 4) & 15]), d[(k + 8) & 15]));
    }
    // un-bias: darker arc = best_lo - 256, brighter arc = 256 - best_hi
    return max((int)best_lo - 256, 256 - (int)best_hi);
}

// One CTA == one cell == the (up to two) cv::FAST calls of ORBExtractor.cpp:133-140.
// The arc score is threshold independent, and for a pixel that is a corner at threshold t every 8-neighbour that is
// not a corner at t scores <= t < its own score, so "survives NMS at t" == "arc > t and arc > all neighbours' arcs
// inside the ROI". The cell is therefore scored once; the threshold is chosen afterwards from the NMS survivors.
__global__ void __launch_bounds__(FAST_NT) k_fast(PyrDev P, const CellDev* __restrict__ cells, uint32_t* __restrict__ cand,
                                                  int* __restrict__ cell_count)
{
    __shared__ uint8_t tile[FAST_DIM][FAST_PITCH];
    __shared__ uint8_t score[FAST_DIM][FAST_PITCH];
    __shared__ uint16_t list[(FAST_DIM - 6) * (FAST_DIM - 6)];
    __shared__ int s_nlist;
    __shared__ int s_warp[2][FAST_NT / 32];

    const CellDev c   = cells[blockIdx.x];
    const LevelDev& L = P.lv[c.level];
    const int img     = blockIdx.y;
    const uint8_t* src = L.img + (size_t)img * L.img_stride + (size_t)c.y0 * L.pitch + c.x0;
    const int w = c.w, h = c.h;
    const int tid = threadIdx.x;

    if (tid == 0) s_nlist = 0;
    for (int i = tid; i < h * w; i += FAST_NT)
    {
        int y = i / w, x = i - y * w;
        tile[y][x]  = src[(size_t)y * L.pitch + x];
        score[y][x] = 0;
    }
    __syncthreads();

    const int zw = w - 6, zh = h - 6;  // detection zone: ROI minus the 3-px frame
    const int nz = (zw > 0 && zh > 0) ? zw * zh : 0;
    const int t  = P.th_min < P.th_ini ? P.th_min : P.th_ini;
    // phase A: cheap rejection (any arc of 9 contains a pixel of every opposite pair)
    for (int i = tid; i < nz; i += FAST_NT)
    {
        const int y = 3 + i / zw, x = 3 + i % zw;
        const int v = tile[y][x];
        const int a = abs(tile[y + 3][x] - v), b = abs(tile[y - 3][x] - v);
        const int e = abs(tile[y][x + 3] - v), f = abs(tile[y][x - 3] - v);
        if (!((a <= t && b <= t) || (e <= t && f <= t)))
        {
            int pos   = atomicAdd(&s_nlist, 1);
            list[pos] = (uint16_t)((y << 8) | x);
        }
    }
    __syncthreads();
    // phase B: full score of the survivors (balanced across the CTA)
    const int nl = s_nlist;
    for (int i = tid; i < nl; i += FAST_NT)
    {
        const int y = list[i] >> 8, x = list[i] & 255;
        const int arc = fast_arc_score(tile, y, x);
        score[y][x]   = (arc > t) ? (uint8_t)arc : 0;
#ifdef SNK_DEBUG_FAST
        if (blockIdx.x == 312 && blockIdx.y == 0 && y == 16 && x == 3)
            printf("DBG cell x0=%d y0=%d w=%d h=%d off=%d,%d pitch=%d | v=%d ring0=%d ring4=%d ring8=%d ring12=%d arc=%d t=%d nl=%d\n", c.x0, c.y0, c.w,
                   c.h, c.off_x, c.off_y, L.pitch, tile[y][x], tile[y + 3][x], tile[y][x + 3], tile[y - 3][x], tile[y][x - 3], arc, t, nl);
#endif
    }
    __syncthreads();

    // phase C: NMS + threshold choice + row-major compaction
    const int per = (nz + FAST_NT - 1) / FAST_NT;  // <= 29
    uint32_t m_min = 0, m_ini = 0;
    for (int k = 0; k < per; ++k)
    {
        const int i = tid * per + k;
        if (i >= nz) break;
        const int y = 3 + i / zw, x = 3 + i % zw;
        const int s = score[y][x];
        // cv response is s-1 and must beat 0 for empty neighbours: s >= 2
        if (s > 1 && s > score[y - 1][x - 1] && s > score[y - 1][x] && s > score[y - 1][x + 1] && s > score[y][x - 1] &&
            s > score[y][x + 1] && s > score[y + 1][x - 1] && s > score[y + 1][x] && s > score[y + 1][x + 1])
        {
            m_min |= 1u << k;
            if (s > P.th_ini) m_ini |= 1u << k;
        }
    }
    const int lane = tid & 31, wid = tid >> 5;
    int c_min = __popc(m_min), c_ini = __popc(m_ini);
    int i_min = warp_incl_scan(c_min, lane), i_ini = warp_incl_scan(c_ini, lane);
    if (lane == 31)
    {
        s_warp[0][wid] = i_min;
        s_warp[1][wid] = i_ini;
    }
    __syncthreads();
    int base_min = 0, base_ini = 0, tot_min = 0, tot_ini = 0;
#pragma unroll
    for (int k = 0; k < FAST_NT / 32; ++k)
    {
        if (k < wid)
        {
            base_min += s_warp[0][k];
            base_ini += s_warp[1][k];
        }
        tot_min += s_warp[0][k];
        tot_ini += s_warp[1][k];
    }
    const bool use_ini = tot_ini > 0;  // "if (cv_KeysCell.empty()) retry with th_fast_min"
    uint32_t m         = use_ini ? m_ini : m_min;
    int pos            = use_ini ? (base_ini + i_ini - c_ini) : (base_min + i_min - c_min);
    const int total    = use_ini ? tot_ini : tot_min;
    const size_t slot  = (size_t)img * P.cells_per_img + blockIdx.x;
    uint32_t* out      = cand + slot * P.cell_cap;
    while (m)
    {
        const int k = __ffs(m) - 1;
        m &= m - 1;
        const int i = tid * per + k;
        const int y = 3 + i / zw, x = 3 + i % zw;
        if (pos < P.cell_cap)
            out[pos] = (uint32_t)(x + c.off_x) | ((uint32_t)(y + c.off_y) << 12) | ((uint32_t)(score[y][x] - 1) << 24);
        ++pos;
    }
    if (tid == 0) cell_count[slot] = total;
}

// --------------------------------------------------------------------------------------------------------------
// Quadtree distribution
// --------------------------------------------------------------------------------------------------------------
struct QNode
{
    float cx, cy, sx, sy;  // corner, size
    int from, to;
};
constexpr int QT_NT = 128;

__device__ __forceinline__ float kp_x(uint32_t p) { return (float)(p & 0xfffu); }
__device__ __forceinline__ float kp_y(uint32_t p) { return (float)((p >> 12) & 0xfffu); }

// dynamic shared memory: 3 node lists + sort keys + per-node child counts/positions = QT_NODE_BYTES per node
constexpr int QT_NODE_BYTES = 3 * 24 + 4 + 16 + 8;
__global__ void __launch_bounds__(QT_NT) k_distribute(PyrDev P, const uint32_t* __restrict__ cand,
                                                      const int* __restrict__ cell_count, int* __restrict__ cell_off,
                                                      uint32_t* __restrict__ workA, uint32_t* __restrict__ workB,
                                                      uint32_t* __restrict__ sel, int* __restrict__ sel_count,
                                                      int node_cap, int* __restrict__ err)
{
    extern __shared__ __align__(16) unsigned char smem[];
    QNode* listA    = reinterpret_cast<QNode*>(smem);
    QNode* listB    = listA + node_cap;
    QNode* leaves   = listB + node_cap;
    uint32_t* keys  = reinterpret_cast<uint32_t*>(leaves + node_cap);
    int* ccnt_big   = reinterpret_cast<int*>(keys + node_cap);           // child sizes of sorted node i (4 per node)
    int2* cpos      = reinterpret_cast<int2*>(ccnt_big + 4 * node_cap);  // (leaf position, next position) of node i
    __shared__ int s_warp[QT_NT / 32], s_ctrl[8];

    const int level = blockIdx.x, img = blockIdx.y;
    const LevelDev& L = P.lv[level];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const size_t wbase = (size_t)img * P.cand_per_img + L.cand_base;
    uint32_t* A = workA + wbase;
    uint32_t* B = workB + wbase;
    uint32_t* out = sel + (size_t)img * P.sel_per_img + L.sel_base;
    const int target = L.quota;
    const int out_cap = target + 4;

    // ---- gather the level's candidates in cell order (== the order DetectKeypoints pushes them) ----
    const int* cc = cell_count + (size_t)img * P.cells_per_img + L.cell_base;
    int* co       = cell_off + (size_t)img * P.cells_per_img + L.cell_base;
    int running   = 0;
    for (int base = 0; base < L.n_cells; base += QT_NT)
    {
        const int i = base + tid;
        int v       = (i < L.n_cells) ? min(cc[i], P.cell_cap) : 0;
        if (i < L.n_cells && cc[i] > P.cell_cap) atomicExch(err, 1);  // cannot happen (cap is the NMS bound); guarded
        int inc = warp_incl_scan(v, lane);
        if (lane == 31) s_warp[wid] = inc;
        __syncthreads();
        int add = 0, tot = 0;
#pragma unroll
        for (int k = 0; k < QT_NT / 32; ++k)
        {
            if (k < wid) add += s_warp[k];
            tot += s_warp[k];
        }
        if (i < L.n_cells) co[i] = running + add + inc - v;
        running += tot;
        __syncthreads();
    }
    const int n = running;
    for (int cidx = wid; cidx < L.n_cells; cidx += QT_NT / 32)
    {
        const int cnt = min(cc[cidx], P.cell_cap), off = co[cidx];
        const uint32_t* s = cand + ((size_t)img * P.cells_per_img + L.cell_base + cidx) * P.cell_cap;
        for (int k = lane; k < cnt; k += 32) A[off + k] = s[k];
    }
    __syncthreads();

    if (n <= target)
    {  // FeatureDistribution.cpp:83-86
        for (int i = tid; i < n; i += QT_NT) out[i] = A[i];
        if (tid == 0) sel_count[img * P.nlevels + level] = n;
        return;
    }

    // ---- root node, FeatureDistribution.cpp:96-103 with min=(16,16), max=(cols-16, rows-16) ----
    QNode* inner = listA;
    QNode* next  = listB;
    if (tid == 0)
    {
        const float minx = 16.f, miny = 16.f, maxx = (float)(L.w - 16), maxy = (float)(L.h - 16);
        const float ccx = __fmul_rn(__fadd_rn(minx, maxx), 0.5f), ccy = __fmul_rn(__fadd_rn(miny, maxy), 0.5f);
        const float side = fmaxf(__fsub_rn(maxx, minx), __fsub_rn(maxy, miny));
        const float half = __fmul_rn(side, 0.5f);
        next[0]          = QNode{__fsub_rn(ccx, half), __fsub_rn(ccy, half), side, side, 0, n};
        s_ctrl[0] = 1;   // n_next
        s_ctrl[1] = 0;   // n_leaves
        s_ctrl[2] = 0;   // last_node_count
        s_ctrl[3] = -1;  // last_processed_inner
        s_ctrl[4] = 0;   // finished
        s_ctrl[5] = 0;   // m: size of `inner` in the last round
        s_ctrl[6] = 0;   // stop (node count unchanged)
    }
    __syncthreads();

    while (true)
    {
        const int n_next = s_ctrl[0];
        if (s_ctrl[4] || n_next == 0) break;
        // inner_nodes.swap(new_inner_nodes); new_inner_nodes.clear();
        QNode* tsw = inner;
        inner      = next;
        next       = tsw;
        const int m        = n_next;
        const int n_leaves = s_ctrl[1];
        __syncthreads();
        if (tid == 0)
        {
            s_ctrl[5] = m;
            s_ctrl[0] = 0;
            if (n_leaves + m == s_ctrl[2])
                s_ctrl[6] = 1;
            else
                s_ctrl[2] = n_leaves + m;
        }
        __syncthreads();
        if (s_ctrl[6]) break;

        // std::sort by NumKeypoints descending (FeatureDistribution.cpp:124) — libstdc++ permutation
        for (int i = tid; i < m; i += QT_NT) keys[i] = ((uint32_t)(inner[i].to - inner[i].from) << gnusort::SHIFT) | (uint32_t)i;
        __syncthreads();
        if (tid == 0) gnusort::sort(keys, m);
        __syncthreads();

        // child sizes of every inner node (the split itself is only applied to the nodes that get processed)
        for (int i = wid; i < m; i += QT_NT / 32)
        {
            const QNode nd = inner[keys[i] & ((1u << gnusort::SHIFT) - 1)];
            const float mx = __fadd_rn(nd.cx, __fmul_rn(nd.sx, 0.5f)), my = __fadd_rn(nd.cy, __fmul_rn(nd.sy, 0.5f));
            int c0 = 0, c1 = 0, c2 = 0, c3 = 0;
            for (int k0 = nd.from; k0 < nd.to; k0 += 32)
            {
                const int k = k0 + lane;
                int q       = -1;
                if (k < nd.to)
                {
                    const uint32_t p = A[k];
                    q                = ((kp_y(p) > my) ? 2 : 0) + ((kp_x(p) > mx) ? 1 : 0);
                }
                c0 += __popc(__ballot_sync(0xffffffffu, q == 0));
                c1 += __popc(__ballot_sync(0xffffffffu, q == 1));
                c2 += __popc(__ballot_sync(0xffffffffu, q == 2));
                c3 += __popc(__ballot_sync(0xffffffffu, q == 3));
            }
            if (lane == 0)
            {
                ccnt_big[4 * i + 0] = c0;
                ccnt_big[4 * i + 1] = c1;
                ccnt_big[4 * i + 2] = c2;
                ccnt_big[4 * i + 3] = c3;
            }
        }
        __syncthreads();

        // serial bookkeeping of the split loop (:127-159): where each node's children land and where it stops
        if (tid == 0)
        {
            int nl = n_leaves, nn = 0, processed = m, fin = 0, lpi = -1;
            for (int i = 0; i < m; ++i)
            {
                cpos[i] = make_int2(nl, nn);
#pragma unroll
                for (int q = 0; q < 4; ++q)
                {
                    const int c = ccnt_big[4 * i + q];
                    nl += (c == 1);
                    nn += (c > 1);
                }
                if (nl + nn + (m - i - 1) >= target)
                {
                    lpi       = i;
                    fin       = 1;
                    processed = i + 1;
                    break;
                }
            }
            if (nl > node_cap || nn > node_cap) atomicExch(err, 2);
            s_ctrl[0] = nn;
            s_ctrl[1] = nl;
            s_ctrl[4] = fin;
            if (fin) s_ctrl[3] = lpi;
            s_ctrl[7] = processed;
        }
        __syncthreads();
        const int processed = s_ctrl[7];

        // splitAndSort (:15-70) of the processed nodes: stable 4-way partition, children appended in order
        for (int i = wid; i < processed; i += QT_NT / 32)
        {
            const QNode nd = inner[keys[i] & ((1u << gnusort::SHIFT) - 1)];
            const float hx = __fmul_rn(nd.sx, 0.5f), hy = __fmul_rn(nd.sy, 0.5f);
            const float mx = __fadd_rn(nd.cx, hx), my = __fadd_rn(nd.cy, hy);
            const int c0 = ccnt_big[4 * i], c1 = ccnt_big[4 * i + 1], c2 = ccnt_big[4 * i + 2], c3 = ccnt_big[4 * i + 3];
            int o0 = nd.from, o1 = o0 + c0, o2 = o1 + c1, o3 = o2 + c2;
            const int len = nd.to - nd.from;
            for (int k0 = 0; k0 < len; k0 += 32)
            {
                const int k = nd.from + k0 + lane;
                int q       = -1;
                uint32_t p  = 0;
                if (k < nd.to)
                {
                    p = A[k];
                    q = ((kp_y(p) > my) ? 2 : 0) + ((kp_x(p) > mx) ? 1 : 0);
                }
                const uint32_t b0 = __ballot_sync(0xffffffffu, q == 0), b1 = __ballot_sync(0xffffffffu, q == 1);
                const uint32_t b2 = __ballot_sync(0xffffffffu, q == 2), b3 = __ballot_sync(0xffffffffu, q == 3);
                const uint32_t lt = (1u << lane) - 1;
                if (q == 0) B[o0 + __popc(b0 & lt)] = p;
                if (q == 1) B[o1 + __popc(b1 & lt)] = p;
                if (q == 2) B[o2 + __popc(b2 & lt)] = p;
                if (q == 3) B[o3 + __popc(b3 & lt)] = p;
                o0 += __popc(b0);
                o1 += __popc(b1);
                o2 += __popc(b2);
                o3 += __popc(b3);
            }
            __syncwarp();
            for (int k = nd.from + lane; k < nd.to; k += 32) A[k] = B[k];
            if (lane == 0)
            {
                int2 pos = cpos[i];
                int from = nd.from;
                const int cs[4] = {c0, c1, c2, c3};
#pragma unroll
                for (int q = 0; q < 4; ++q)
                {
                    const int c = cs[q];
                    QNode ch;
                    ch.cx   = (q & 1) ? mx : nd.cx;
                    ch.cy   = (q & 2) ? my : nd.cy;
                    ch.sx   = hx;
                    ch.sy   = hy;
                    ch.from = from;
                    ch.to   = from + c;
                    from += c;
                    if (c == 1)
                    {
                        if (pos.x < node_cap) leaves[pos.x] = ch;
                        ++pos.x;
                    }
                    else if (c > 1)
                    {
                        if (pos.y < node_cap) next[pos.y] = ch;
                        ++pos.y;
                    }
                }
            }
        }
        __syncthreads();
    }
    __syncthreads();

    // ---- emit the best keypoint of every leaf, every new inner node and inner[last_processed..] (:163-187) ----
    const int n_leaves = s_ctrl[1], n_next = s_ctrl[0], m = s_ctrl[5], lpi = s_ctrl[3];
    const int n_tail   = (lpi >= 0) ? (m - lpi) : 0;
    const int total    = n_leaves + n_next + n_tail;
    if (total > out_cap && tid == 0) atomicExch(err, 3);
    for (int o = wid; o < total; o += QT_NT / 32)
    {
        QNode nd;
        if (o < n_leaves)
            nd = leaves[o];
        else if (o < n_leaves + n_next)
            nd = next[o - n_leaves];
        else
            nd = inner[keys[lpi + (o - n_leaves - n_next)] & ((1u << gnusort::SHIFT) - 1)];
        // std::max_element on response: first maximum in range order
        uint32_t bestkey = 0;  // (response+1) << 20 | (0xfffff - relative index)
        for (int k = nd.from + lane; k < nd.to; k += 32)
        {
            const uint32_t key = (((A[k] >> 24) + 1u) << 20) | (0xfffffu - (uint32_t)(k - nd.from));
            bestkey            = max(bestkey, key);
        }
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) bestkey = max(bestkey, __shfl_xor_sync(0xffffffffu, bestkey, s));
        if (lane == 0 && o < out_cap) out[o] = A[nd.from + (int)(0xfffffu - (bestkey & 0xfffffu))];
    }
    if (tid == 0) sel_count[img * P.nlevels + level] = min(total, out_cap);
}

// --------------------------------------------------------------------------------------------------------------
// Gaussian blur 7x7, sigma 2, kernel {18,34,48,56,48,34,18}/256 in 8.8 fixed point, BORDER_REFLECT_101
// --------------------------------------------------------------------------------------------------------------
constexpr int BL_TW = 128, BL_TH = 32;
struct TileDev
{
    short level, tx, ty, _pad;
};

__device__ __forceinline__ int reflect101(int p, int len)
{
    // valid for -len < p < 2*len-1 (halo of 3, levels are far larger)
    p = p < 0 ? -p : p;
    return p >= len ? 2 * (len - 1) - p : p;
}

__global__ void __launch_bounds__(256) k_blur(PyrDev P, const TileDev* __restrict__ tiles)
{
    __shared__ uint8_t src[BL_TH + 6][BL_TW + 8];
    __shared__ uint16_t hor[BL_TH + 6][BL_TW];
    const TileDev t   = tiles[blockIdx.x];
    const LevelDev& L = P.lv[t.level];
    const uint8_t* S  = L.img + (size_t)blockIdx.y * L.img_stride;
    uint8_t* D        = L.gauss + (size_t)blockIdx.y * L.gauss_stride;
    const int x0 = t.tx * BL_TW, y0 = t.ty * BL_TH;
    const int tw = min(BL_TW, L.w - x0), th = min(BL_TH, L.h - y0);
    const int tid = threadIdx.x;

    for (int i = tid; i < (th + 6) * (tw + 6); i += 256)
    {
        const int r = i / (tw + 6), c = i - r * (tw + 6);
        const int gy = reflect101(y0 + r - 3, L.h), gx = reflect101(x0 + c - 3, L.w);
        src[r][c] = S[(size_t)gy * L.pitch + gx];
    }
    __syncthreads();
    for (int i = tid; i < (th + 6) * tw; i += 256)
    {
        const int r = i / tw, c = i - r * tw;
        const uint8_t* p = &src[r][c];
        hor[r][c] = (uint16_t)(18 * (p[0] + p[6]) + 34 * (p[1] + p[5]) + 48 * (p[2] + p[4]) + 56 * p[3]);
    }
    __syncthreads();
    for (int i = tid; i < th * tw; i += 256)
    {
        const int r = i / tw, c = i - r * tw;
        const uint32_t v = 18u * (hor[r][c] + hor[r + 6][c]) + 34u * (hor[r + 1][c] + hor[r + 5][c]) +
                           48u * (hor[r + 2][c] + hor[r + 4][c]) + 56u * hor[r + 3][c];
        D[(size_t)(y0 + r) * L.gauss_pitch + x0 + c] = (uint8_t)((v + 32768u) >> 16);
    }
}

// --------------------------------------------------------------------------------------------------------------
// Orientation + rBRIEF + output record. One warp per selected keypoint.
// --------------------------------------------------------------------------------------------------------------
__constant__ signed char c_pattern[1024];  // 512 points (x,y) — include/snakeorb_pattern.h
__constant__ int c_umax[16];               // ORBPattern::AngleUmax()

constexpr int DS_WARPS = 8;

__global__ void __launch_bounds__(DS_WARPS * 32) k_describe(PyrDev P, const uint32_t* __restrict__ sel,
                                                            const int* __restrict__ sel_count,
                                                            snk_keypoint_f32* __restrict__ kps,
                                                            uint64_t* __restrict__ desc, int* __restrict__ counts)
{
    __shared__ char2 pat[16][32];  // pat[j][lane] = point 16*lane + j  (conflict-free: lanes are consecutive)
    __shared__ int s_pref[SNK_MAX_LEVELS + 1];
    const int img = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (int i = tid; i < 512; i += DS_WARPS * 32)
        pat[i & 15][i >> 4] = make_char2(c_pattern[2 * i], c_pattern[2 * i + 1]);
    if (tid == 0)
    {
        int acc = 0;
        for (int l = 0; l < P.nlevels; ++l)
        {
            s_pref[l] = acc;
            acc += sel_count[img * P.nlevels + l];
        }
        s_pref[P.nlevels] = acc;
        if (blockIdx.x == 0) counts[img] = min(acc, P.kp_stride);
    }
    __syncthreads();
    const int total = min(s_pref[P.nlevels], P.kp_stride);
    const int g     = blockIdx.x * DS_WARPS + wid;
    if (g >= total) return;
    int level = 0;
    while (level + 1 < P.nlevels && g >= s_pref[level + 1]) ++level;
    const LevelDev& L = P.lv[level];
    const uint32_t p  = sel[(size_t)img * P.sel_per_img + L.sel_base + (g - s_pref[level])];
    // kp.point += (minBorderX, minBorderY)   ORBExtractor.cpp:165-166
    const int x = (int)(p & 0xfffu) + 16, y = (int)((p >> 12) & 0xfffu) + 16;
    const float resp = (float)(p >> 24);

    // ---- ORB::ComputeAngle: integer moments over the radius-15 disc of the un-blurred level ----
    const uint8_t* raw = L.img + (size_t)img * L.img_stride + (size_t)y * L.pitch + x;
    const int u        = lane - 15;  // lane 31 idles
    int m10 = 0, m01 = 0;
    if (lane < 31)
    {
        const int au = abs(u);
#pragma unroll 1
        for (int v = -15; v <= 15; ++v)
        {
            if (au <= c_umax[abs(v)])
            {
                const int val = raw[v * L.pitch + u];
                m10 += u * val;
                m01 += v * val;
            }
        }
    }
    m10 = warp_sum(m10);
    m01 = warp_sum(m01);
    // degrees(atan2f(m01, m10)): a * 180.f / pi_f, then wrap negatives by +360 (OrbDescriptors.cpp:52-53)
    float angle = snkmath::atan2f_fdlibm((float)m01, (float)m10);
    angle       = __fdiv_rn(__fmul_rn(angle, 180.0f), 3.14159265358979323846f);
    angle       = __fadd_rn((angle < 0.f) ? 360.f : 0.f, angle);

    // ---- ORB::ComputeDescriptor on the blurred level; lane i produces byte i ----
    const float rad = __fmul_rn(__fdiv_rn(angle, 180.0f), 3.14159265358979323846f);
    const float a = snkmath::cosf_glibc(rad), b = snkmath::sinf_glibc(rad);
    const uint8_t* gs = L.gauss + (size_t)img * L.gauss_stride;
    const float px = (float)x, py = (float)y;
    uint32_t byte = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j)
    {
        int tv[2];
#pragma unroll
        for (int e = 0; e < 2; ++e)
        {
            const char2 q  = pat[2 * j + e][lane];
            const float ux = (float)q.x, uy = (float)q.y;
            const float fx = __fadd_rn(px, __fsub_rn(__fmul_rn(ux, a), __fmul_rn(uy, b)));
            const float fy = __fadd_rn(py, __fadd_rn(__fmul_rn(ux, b), __fmul_rn(uy, a)));
            const int ix   = (int)__fadd_rn(fx, fx >= 0.f ? 0.5f : -0.5f);  // Saiga::iRound
            const int iy   = (int)__fadd_rn(fy, fy >= 0.f ? 0.5f : -0.5f);
            tv[e]          = gs[(size_t)iy * L.gauss_pitch + ix];
        }
        byte |= (uint32_t)(tv[0] < tv[1]) << j;
    }
    // 4 lanes -> one little-endian 32-bit word
    uint32_t word = byte << (8 * (lane & 3));
    word |= __shfl_xor_sync(0xffffffffu, word, 1);
    word |= __shfl_xor_sync(0xffffffffu, word, 2);
    const size_t o = (size_t)img * P.kp_stride + g;
    if ((lane & 3) == 0) reinterpret_cast<uint32_t*>(desc + 4 * o)[lane >> 2] = word;

    // ---- output record (ORBExtractor.cpp:163-171, 223-232) ----
    if (lane < 6)
    {
        float f;
        const float sc = L.scale;
        if (lane == 0)
            f = (level != 0) ? __fmul_rn(px, sc) : px;
        else if (lane == 1)
            f = (level != 0) ? __fmul_rn(py, sc) : py;
        else if (lane == 2)
            f = L.size;
        else if (lane == 3)
            f = angle;
        else if (lane == 4)
            f = resp;
        else
            f = __int_as_float(level);
        reinterpret_cast<float*>(kps + o)[lane] = f;
    }
}
}  // namespace snk
