// Device side of the ORB extractor: data layout descriptors and the kernels.
//
//   k_resize      level l from level l-1        cv::resize INTER_LINEAR fixed point    ORBExtractor.cpp:282
//   k_fast_group  FAST-9/16 + NMS per cell,     cv::FAST per cell, two thresholds      ORBExtractor.cpp:114-157
//                 4 neighbouring cells per CTA  (fast_group.cuh; k_fast = the earlier one-cell kernel, SNK_FAST_PER_CELL=1)
//   k_distribute  quadtree selection per level  QuadtreeFeatureDistributor::Distribute FeatureDistribution.cpp:72-189
//   k_blur        7x7 Gaussian, reflect-101     cv::GaussianBlur                       ORBExtractor.cpp:213
//   k_describe    orientation + rBRIEF + pack   ORB::ComputeAngle / ComputeDescriptor  OrbDescriptors.cpp:26-125,
//                                               and the Detect tail                    ORBExtractor.cpp:163-171,223-232
//
// All kernels are batched: blockIdx.y (or .z) is the image of the batch; images never interact.
#pragma once
#include "common.cuh"
#include "glibc_math.cuh"
#include "gnu_sort_cta.cuh"

#include <cuda.h>  // CUtensorMap (the encoder is fetched through cudaGetDriverEntryPoint, no link dependency)

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

// One TMA descriptor per pyramid level: 3-D tensor (x, y, image) of u8 with box FAST_PITCH x (max cell height) x 1.
struct TmaMaps
{
    CUtensorMap level[SNK_MAX_LEVELS];
};

// One FAST cell == one cv::FAST call of the reference: ROI origin/size inside its level and the offset the reference
// adds to the ROI-relative coordinates (ORBExtractor.cpp:152-153).
struct CellDev
{
    short level;
    short x0, y0;  // iniX, iniY
    short w, h;    // maxX-iniX, maxY-iniY
    short off_x, off_y;
    // lane -> (row, word) splits of the kernel's two word-wise passes, precomputed on the host (no integer division
    // on the device): a warp covers `rows` rows of `words` words per step; lane / words == (lane * inv) >> 16
    unsigned short box_h, ld_inv, ld_rows;      // rows of the level's TMA box (tile bytes = FAST_PITCH * box_h)
    unsigned short zn_words, zn_inv, zn_rows;   // compass pass: words intersecting the zone columns
    unsigned short zn_word0;                    // first such word
    unsigned short _pad[2];
};
static_assert(sizeof(CellDev) == 32, "CellDev layout");

// --------------------------------------------------------------------------------------------------------------
// Pyramid: level l from level l-1, cv::resize INTER_LINEAR fixed point.
//
// One warp per 128 x RS_ROWS output strip, no shared memory. A lane owns 4 adjacent output pixels. Everything that
// depends on x only (source word, byte shift, byte selectors, the four 11-bit coefficient pairs) comes from a per-level
// table built on the host exactly as cv::resize builds xofs/ialpha, and is loaded once per lane. Per source row a lane
// loads 3 aligned words, funnel-shifts them so the window starts at its first tap, permutes the (tap0, tap1) byte pairs
// of its four pixels into two registers and gets the four horizontal sums with dp2a (u16 coefficient pair x u8 pair).
// Both source rows are loaded for every output row: re-using the previous row's sums saved instructions but chained
// the rows of a strip; independent iterations (unrolled by 4) let the loads of the next rows start early, which is
// what this latency-bound kernel needs (measured: 0.61 -> 0.57 ms per 512 images).
// --------------------------------------------------------------------------------------------------------------
constexpr int RS_TW = 128, RS_ROWS = 32;
struct ResizeQuad  // 32 bytes per 4 output pixels
{
    int word;            // first source word of the window
    int shift_bits;      // 8 * (xofs[x] & 3)
    unsigned sel;        // byte selectors: low 16 bits pixels 0,1 / high 16 bits pixels 2,3: [t0 t1 t0 t1] each
    unsigned _pad;
    unsigned coef[4];    // ialpha pair per pixel: a0 | a1 << 16
};

struct ResizeRow  // 16 bytes per output row
{
    int sy0, sy1;      // the two source rows, already clamped to the image
    unsigned b0, b1;   // vertical coefficients << 16 (0..2048: linear interpolation never produces negative ones)
};

// hi32(a * b) + c in one IMAD.HI
__device__ __forceinline__ unsigned mad_hi(unsigned a, unsigned b, unsigned c)
{
    unsigned d;
    asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// The same resize with the source staged by TMA: a CTA's 128 x 32 output pixels read a source window of about 160 x 40
// bytes, fetched as ONE tensor box (start rounded down to 16 bytes, out-of-image bytes zero-filled: they only ever meet a
// zero coefficient) while the threads fetch their table entries; the arithmetic then runs on shared memory — no
// global-load latency in the row loop, no 64-bit address arithmetic (k_resize: 26 thread-instructions per pixel, 60 % of
// the issue slots busy, the rest waiting on loads). Box sizes per level from the host (largest window of any tile).
#ifndef SNK_RT_H
#define SNK_RT_H 64
#endif
#ifndef SNK_RT_UNROLL
#define SNK_RT_UNROLL 4
#endif
constexpr int RT_W = 128, RT_H = SNK_RT_H, RT_UNROLL = SNK_RT_UNROLL;
__global__ void __launch_bounds__(128) k_resize_tma(const __grid_constant__ CUtensorMap map, uint8_t* __restrict__ dst, size_t dst_stride,
                                                    int dpitch, int dw, int dh, const ResizeQuad* __restrict__ quads,
                                                    const ResizeRow* __restrict__ rows, int box_w, int box_h, int* __restrict__ err)
{
    extern __shared__ __align__(128) unsigned char rs_smem[];
    __shared__ __align__(8) unsigned long long rs_bar;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int x0t = blockIdx.x * RT_W, y0t = blockIdx.y * RT_H;
    const int tile_x0 = (__ldg(&quads[x0t >> 2].word) * 4) & ~15, tile_y0 = __ldg(&rows[y0t].sy0);
    const unsigned bar = (unsigned)__cvta_generic_to_shared(&rs_bar);
    if (tid == 0)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((unsigned)(box_w * box_h)) : "memory");
        asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
                (unsigned)__cvta_generic_to_shared(rs_smem)),
            "l"(reinterpret_cast<unsigned long long>(&map)), "r"(tile_x0), "r"(tile_y0), "r"((int)blockIdx.z), "r"(bar)
            : "memory");
    }
    __syncthreads();  // the barrier is initialised before anybody polls it
    const int x = x0t + 4 * lane, y0 = y0t + (RT_H / 4) * wid;
    if (y0 >= dh || x >= dw) return;  // no CTA-wide primitives below
    uint8_t* D         = dst + (size_t)blockIdx.z * dst_stride + (size_t)y0 * dpitch + x;
    const ResizeQuad q = quads[x >> 2];
    const unsigned sel01 = q.sel & 0xffffu, sel23 = q.sel >> 16;
    const int pitch_w = box_w >> 2;
    const uint32_t* T0 = reinterpret_cast<const uint32_t*>(rs_smem) + (q.word - (tile_x0 >> 2)) - tile_y0 * pitch_w;
    {
        unsigned done = 0;
        for (int spin = 0; !done && spin < (1 << 22); ++spin)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, 0x989680;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done)
                         : "r"(bar), "r"(0u)
                         : "memory");
        if (!done && lane == 0) atomicExch(err, 4);
    }
    auto hrow = [&](int sy, unsigned (&h)[4]) {
        const uint32_t* row = T0 + sy * pitch_w;
        const unsigned w0 = row[0], w1 = row[1], w2 = row[2];
        const unsigned v0 = __funnelshift_r(w0, w1, q.shift_bits), v1 = __funnelshift_r(w1, w2, q.shift_bits);
        const unsigned r01 = __byte_perm(v0, v1, sel01), r23 = __byte_perm(v0, v1, sel23);
        h[0] = __dp2a_lo(q.coef[0], r01, 0u) >> 4;
        h[1] = __dp2a_hi(q.coef[1], r01, 0u) >> 4;
        h[2] = __dp2a_lo(q.coef[2], r23, 0u) >> 4;
        h[3] = __dp2a_hi(q.coef[3], r23, 0u) >> 4;
    };
    // consecutive output rows mostly share a source row (sy0 of a row is sy1 of the row above unless the step skips one):
    // its horizontal sums are kept — a uniform branch, and shared-memory latency is short enough for the serial chain
    unsigned ha[4], hb[4] = {0u, 0u, 0u, 0u};
    int have = -1;  // source row whose sums are in hb
    const int y1 = min(y0 + RT_H / 4, dh);
#pragma unroll RT_UNROLL
    for (int y = y0; y < y1; ++y, D += dpitch)
    {
        const int4 rt = __ldg(reinterpret_cast<const int4*>(rows) + y);
        if (rt.x == have)
        {
#pragma unroll
            for (int j = 0; j < 4; ++j) ha[j] = hb[j];
        }
        else
            hrow(rt.x, ha);
        hrow(rt.y, hb);
        have = rt.y;
        unsigned v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = mad_hi(hb[j], (unsigned)rt.w, mad_hi(ha[j], (unsigned)rt.z, 2u)) >> 2;  // as in k_resize
        const unsigned packed = __byte_perm(__byte_perm(v[0], v[1], 0x0040), __byte_perm(v[2], v[3], 0x0040), 0x5410);
        // whole words always: the destination is one of the pyramid's own levels, whose pitch is a multiple of 64 — the
        // bytes of a row's last word beyond the width land in the pitch padding, which nothing reads as pixels (the
        // byte-wise tail of k_resize cost 11 % of this kernel's instructions: every warp at the right edge ran both paths)
        *reinterpret_cast<uint32_t*>(D) = packed;
    }
}

// CLAMP: the source may end with its last row (level 0 aliasing a caller's buffer): the window's second and third word
// are clamped to the row. Otherwise the source is one of the pyramid's own buffers — rows are followed by pitch padding, the
// next row, or the allocation's slack — and the three words are read at fixed offsets from one address (a word past the
// row only ever meets a zero coefficient or a selector that does not pick it).
template <bool CLAMP>
__global__ void __launch_bounds__(128) k_resize(const uint8_t* __restrict__ src, size_t src_stride, int spitch, int sh,
                                                uint8_t* __restrict__ dst, size_t dst_stride, int dpitch, int dw, int dh,
                                                const ResizeQuad* __restrict__ quads, const ResizeRow* __restrict__ rows,
                                                int rows_per_strip)
{
    const int lane  = threadIdx.x & 31;
    const int strip = blockIdx.y * 4 + (threadIdx.x >> 5);
    const int x     = blockIdx.x * RS_TW + 4 * lane;
    const int y0    = strip * rows_per_strip;  // shorter strips on small levels: the rows of a strip are a serial chain
    if (y0 >= dh || x >= dw) return;  // no warp-wide primitives below
    const uint8_t* S = src + (size_t)blockIdx.z * src_stride;
    uint8_t* D       = dst + (size_t)blockIdx.z * dst_stride + (size_t)y0 * dpitch + x;
    const ResizeQuad q = quads[x >> 2];
    const int wlast = (spitch >> 2) - 1;
    const int i0 = min(q.word, wlast), i1 = min(q.word + 1, wlast), i2 = min(q.word + 2, wlast);
    const unsigned sel01 = q.sel & 0xffffu, sel23 = q.sel >> 16;

    // horizontal pass of one source row: the four sums, already >> 4 as the vertical pass consumes them
    const uint32_t* S0 = reinterpret_cast<const uint32_t*>(S) + q.word;
    const int pitch_w  = spitch >> 2;
    auto hrow = [&](int sy, unsigned (&h)[4]) {
        unsigned w0, w1, w2;
        if (CLAMP)
        {
            const uint32_t* row = reinterpret_cast<const uint32_t*>(S + (size_t)sy * spitch);
            w0 = __ldg(row + i0), w1 = __ldg(row + i1), w2 = __ldg(row + i2);
        }
        else
        {
            const uint32_t* row = S0 + sy * pitch_w;
            w0 = __ldg(row), w1 = __ldg(row + 1), w2 = __ldg(row + 2);
        }
        const unsigned v0 = __funnelshift_r(w0, w1, q.shift_bits), v1 = __funnelshift_r(w1, w2, q.shift_bits);
        const unsigned r01 = __byte_perm(v0, v1, sel01), r23 = __byte_perm(v0, v1, sel23);
        h[0] = __dp2a_lo(q.coef[0], r01, 0u) >> 4;
        h[1] = __dp2a_hi(q.coef[1], r01, 0u) >> 4;
        h[2] = __dp2a_lo(q.coef[2], r23, 0u) >> 4;
        h[3] = __dp2a_hi(q.coef[3], r23, 0u) >> 4;
    };

    unsigned ha[4], hb[4];
    const int y1 = min(y0 + rows_per_strip, dh);
#pragma unroll 4
    for (int y = y0; y < y1; ++y, D += dpitch)
    {
        const int4 rt = __ldg(reinterpret_cast<const int4*>(rows) + y);  // same address in every lane: one broadcast load
        const int sy0 = rt.x, sy1 = rt.y;
        const unsigned b0 = (unsigned)rt.z, b1 = (unsigned)rt.w;
        hrow(sy0, ha);
        hrow(sy1, hb);
        // cv's vertical pass: uchar((((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2) — each product is
        // truncated on its own, which is hi32((S >> 4) * (b << 16)): two IMAD.HI per pixel. The sum never exceeds 1022,
        // and the reference narrows with a plain cast, so the low byte is the result.
        unsigned v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = mad_hi(hb[j], b1, mad_hi(ha[j], b0, 2u)) >> 2;
        const unsigned packed = __byte_perm(__byte_perm(v[0], v[1], 0x0040), __byte_perm(v[2], v[3], 0x0040), 0x5410);
        *reinterpret_cast<uint32_t*>(D) = packed;  // whole words always, as in k_resize_tma (the tail lands in the pitch padding)
    }
}

// --------------------------------------------------------------------------------------------------------------
// FAST
// --------------------------------------------------------------------------------------------------------------
constexpr int FAST_DIM = 66;    // wCell < 60  =>  ROI side <= 66
constexpr int FAST_PITCH = 96;  // tile row pitch == TMA box width: ROI (<= 66) + up to 15 bytes of alignment, multiple of 16
constexpr int FAST_NT = 128;
constexpr int FAST_ZMAX = (FAST_DIM - 6) * (FAST_DIM - 6);  // detection-zone pixels
constexpr int FAST_BMW = (FAST_ZMAX + 31) / 32;

// arc score: max over the 16 arcs of 9 contiguous ring pixels of min(v - ring) and min(ring - v).
// Both polarities run through ONE min network on packed 16-bit pairs (VIMNMX3.U16x2): lane 0 carries d = 256 + v - ring
// (always positive), lane 1 carries 1023 - d, so a window minimum of lane 1 is 1023 - (window maximum of d).
// Window of 9 = min3 of three windows of 3. (An earlier signed 32-bit formulation was mis-narrowed by nvcc 12.9 to
// min.u16/max.u16 and returned wrong scores on sm_100a; everything here is explicitly unsigned and non-negative, and
// tools/dev_check.cu checks the function on the device against brute force.)
template <int PITCH>
__device__ __forceinline__ int fast_arc_score(const uint8_t (*t)[PITCH], int y, int x)
{
    // d = 256 + centre - ring pixel; e = d | (1023 - d) << 16 = (256 + centre) * 0xFFFF0001 + 0x03FF0000 + pixel * 0xFFFF
    const unsigned v = (256u + t[y][x]) * 0xFFFF0001u + 0x03FF0000u;
    unsigned e[16];
#define SNK_E(k, yy, xx) e[k] = t[yy][xx] * 0x0000FFFFu + v;
    SNK_E(0, y + 3, x) SNK_E(1, y + 3, x + 1) SNK_E(2, y + 2, x + 2) SNK_E(3, y + 1, x + 3)
    SNK_E(4, y, x + 3) SNK_E(5, y - 1, x + 3) SNK_E(6, y - 2, x + 2) SNK_E(7, y - 3, x + 1)
    SNK_E(8, y - 3, x) SNK_E(9, y - 3, x - 1) SNK_E(10, y - 2, x - 2) SNK_E(11, y - 1, x - 3)
    SNK_E(12, y, x - 3) SNK_E(13, y + 1, x - 3) SNK_E(14, y + 2, x - 2) SNK_E(15, y + 3, x - 1)
#undef SNK_E
    unsigned m3[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) m3[k] = __vimin3_u16x2(e[k], e[(k + 1) & 15], e[(k + 2) & 15]);
    unsigned best = 0;  // per-lane maximum over the 16 window minima
#pragma unroll
    for (int k = 0; k < 16; k += 2)
    {
        const unsigned a = __vimin3_u16x2(m3[k], m3[(k + 3) & 15], m3[(k + 6) & 15]);
        const unsigned b = __vimin3_u16x2(m3[k + 1], m3[(k + 4) & 15], m3[(k + 7) & 15]);
        best             = __vimax3_u16x2(best, a, b);
    }
    // darker arc = lane0 - 256; brighter arc = 256 - min(window max) = lane1 - 767
    return max((int)(best & 0xffffu) - 256, (int)(best >> 16) - 767);
}

// One CTA == one cell == the (up to two) cv::FAST calls of ORBExtractor.cpp:133-140.
// The arc score is threshold independent, and for a pixel that is a corner at threshold t every 8-neighbour that is
// not a corner at t scores <= t < its own score, so "survives NMS at t" == "arc > t and arc > all neighbours' arcs
// inside the ROI". The cell is therefore scored once at the lower threshold; the threshold is chosen afterwards from
// the NMS survivors. Work is filtered in stages so that the expensive steps only see the pixels that need them:
//   A  compass test on every zone pixel, 4 pixels per thread with byte SIMD (necessary condition for any arc of 9)
//   B  arc score of the survivors of A (packed 16-bit min network)  -> corners at the lower threshold
//   C  3x3 NMS of the corners, survivors marked in a bitmap over the zone; popc-prefix of the bitmap gives every
//      survivor its row-major rank (== cv::FAST output order) without sorting
// A CTA walks FAST_CPB consecutive cells; the tile of cell k+1 is fetched by TMA into the other buffer while cell k is
// processed (two tile buffers, two mbarriers), so the load latency of all but the first cell is hidden.
constexpr int FAST_CPB = 4;
__global__ void __launch_bounds__(FAST_NT) k_fast(PyrDev P, const __grid_constant__ TmaMaps maps,
                                                  const CellDev* __restrict__ cells, uint32_t* __restrict__ cand,
                                                  int* __restrict__ cell_count, int* __restrict__ err, int tile_rows, int list_cap)
{
    // dynamic shared memory: 2 tile buffers + score tile (tile_rows x FAST_PITCH each, 128-byte aligned), 2 lists
    extern __shared__ __align__(128) unsigned char fast_smem[];
    const int tile_bytes = tile_rows * FAST_PITCH;
    uint8_t(*score)[FAST_PITCH] = reinterpret_cast<uint8_t(*)[FAST_PITCH]>(fast_smem + 2 * tile_bytes);
    uint16_t* list1 = reinterpret_cast<uint16_t*>(fast_smem + 3 * tile_bytes);
    uint16_t* list2 = list1 + list_cap;
    __shared__ uint32_t bm[2][FAST_BMW];  // [0] survivors at the lower threshold, [1] those also above th_ini
    __shared__ uint16_t pre[FAST_BMW + 1];
    __shared__ int s_n1, s_n2, s_ini, s_warp[FAST_NT / 32];
    __shared__ __align__(8) unsigned long long tma_bar[2];

    const int img = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int cell0 = blockIdx.x * FAST_CPB, ncell = min(FAST_CPB, P.cells_per_img - cell0);
    const unsigned bar_addr0  = (unsigned)__cvta_generic_to_shared(&tma_bar[0]);
    const unsigned tile_addr0 = (unsigned)__cvta_generic_to_shared(fast_smem);

    // one TMA box copy (cp.async.bulk.tensor) per cell, completion on the buffer's mbarrier. TMA tiled loads need a
    // 16-byte aligned innermost start coordinate (an unaligned one raises "illegal instruction" on sm_100a,
    // tools/tma_check.cu): the box starts at the aligned column, the ROI sits `a` = x0 & 15 bytes into the tile rows
    auto issue = [&](const CellDev& cc, int buf) {
        const unsigned bar = bar_addr0 + 8u * buf, dst = tile_addr0 + (unsigned)(buf * tile_bytes);
        const unsigned bytes = (unsigned)(FAST_PITCH * cc.box_h);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
            "l"(reinterpret_cast<unsigned long long>(&maps.level[cc.level])), "r"((int)cc.x0 - (cc.x0 & 15)), "r"((int)cc.y0), "r"(img),
            "r"(bar)
            : "memory");
    };
    if (tid == 0)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_addr0));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_addr0 + 8u));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    CellDev c = cells[cell0];
    if (tid == 0) issue(c, 0);
    for (int i = tid; i < tile_bytes >> 4; i += FAST_NT) reinterpret_cast<uint4*>(&score[0][0])[i] = make_uint4(0u, 0u, 0u, 0u);

#pragma unroll 1
    for (int kc = 0; kc < ncell; ++kc)
    {
    const int buf = kc & 1;
    uint8_t(*tile)[FAST_PITCH] = reinterpret_cast<uint8_t(*)[FAST_PITCH]>(fast_smem + buf * tile_bytes);
    CellDev cnext = c;
    if (kc + 1 < ncell)
    {
        cnext = cells[cell0 + kc + 1];
        if (tid == 0) issue(cnext, buf ^ 1);  // that buffer's last readers passed the barrier that ended iteration kc-1
    }
    const int w = c.w, h = c.h;
    const int a = c.x0 & 15;
    // while the copies are in flight: reset the per-cell state, clear the score tile (16-byte stores) and the bitmaps
    if (tid == 0) s_n1 = s_n2 = s_ini = 0;
    for (int i = tid; i < 2 * FAST_BMW; i += FAST_NT) (&bm[0][0])[i] = 0;
    // the score tile is all zero here: cleared once before the first cell, and every cell resets the entries it wrote
    {
        // bounded wait: a copy that never completes (bad descriptor) must not hang the device; it is reported instead
        const unsigned bar = bar_addr0 + 8u * buf, parity = (unsigned)((kc >> 1) & 1);
        unsigned done = 0;
        for (int spin = 0; !done && spin < (1 << 22); ++spin)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done)
                         : "r"(bar), "r"(parity)
                         : "memory");
        if (!done && tid == 0) atomicExch(err, 4);
    }
    __syncthreads();

    const int zw = w - 6, zh = h - 6;  // detection zone: ROI minus the 3-px frame
    const int nz = (zw > 0 && zh > 0) ? zw * zh : 0;
    const int t  = P.th_min < P.th_ini ? P.th_min : P.th_ini;

    // ---- A: compass test, 4 pixels (one aligned tile word) per thread ----
    // |p - v| per byte with VABSDIFF4; "byte > t" as bit 7 of ((x & 0x7f) + (127 - t)) | x  (no carry between bytes);
    // a corner needs two adjacent compass points beyond the threshold on the same side, so in particular
    // (S or N far) and (E or W far). The sign is checked by the ring test that follows.
    if (nz)
    {
        const int c0 = 3 + a, c1 = 3 + a + zw;              // zone columns [c0, c1) in tile coordinates
        const int wz0 = c.zn_word0, nwz = c.zn_words;  // words that intersect the zone, <= 17
        const int rpw = c.zn_rows, sub = (lane * c.zn_inv) >> 16, wx = wz0 + lane - sub * nwz;
        const unsigned addc = (unsigned)(127 - min(t, 127)) * 0x01010101u;
        unsigned zmask = 0x80808080u;  // bytes of my word that lie inside the zone columns
        if (4 * wx < c0) zmask &= 0xffffffffu << (8 * (c0 - 4 * wx));
        if (4 * wx + 4 > c1) zmask &= 0xffffffffu >> (8 * (4 * wx + 4 - c1));
        const int wl = max(wx - 1, 0), wr = min(wx + 1, FAST_PITCH / 4 - 1);
        if (sub < rpw)
            for (int y = 3 + wid * rpw + sub; y < 3 + zh; y += (FAST_NT / 32) * rpw)
            {
                const uint32_t* rc = reinterpret_cast<const uint32_t*>(&tile[y][0]);
                const unsigned v  = rc[wx];
                const unsigned pS = rc[wx + 3 * (FAST_PITCH / 4)], pN = rc[wx - 3 * (FAST_PITCH / 4)];
                const unsigned pE = __funnelshift_r(v, rc[wr], 24);  // bytes x+3
                const unsigned pW = __funnelshift_r(rc[wl], v, 8);   // bytes x-3
                const unsigned dS = __vabsdiffu4(pS, v), dN = __vabsdiffu4(pN, v), dE = __vabsdiffu4(pE, v), dW = __vabsdiffu4(pW, v);
                const unsigned fS = ((dS & 0x7f7f7f7fu) + addc) | dS, fN = ((dN & 0x7f7f7f7fu) + addc) | dN;
                const unsigned fE = ((dE & 0x7f7f7f7fu) + addc) | dE, fW = ((dW & 0x7f7f7f7fu) + addc) | dW;
                unsigned m4 = (fS | fN) & (fE | fW) & zmask;  // bit 8j+7: pixel j of the word is a candidate
                if (m4)
                {
                    int off = atomicAdd(&s_n1, __popc(m4));  // list order is irrelevant (ranks come from the bitmap)
                    do
                    {
                        const int j = (__ffs(m4) - 1) >> 3;
                        m4 &= m4 - 1;
                        list1[off++] = (uint16_t)((y << 8) | (4 * wx + j - a));
                    } while (m4);
                }
            }
    }
    __syncthreads();
    // ---- B: arc score of every candidate; corners (arc > lower threshold) go to list2 ----
    const int n1 = s_n1;
    for (int base = 0; base < n1; base += FAST_NT)
    {
        const int k = base + tid;
        bool pass   = false;
        uint16_t e  = 0;
        if (k < n1)
        {
            e = list1[k];
            const int y = e >> 8, xa = (e & 255) + a;
            const int arc = fast_arc_score(tile, y, xa);
            pass = arc > t;
            if (pass) score[y][xa] = (uint8_t)arc;
        }
        const unsigned m = __ballot_sync(0xffffffffu, pass);
        int off = 0;
        if (lane == 0 && m) off = atomicAdd(&s_n2, __popc(m));
        off = __shfl_sync(0xffffffffu, off, 0);
        if (pass) list2[off + __popc(m & ((1u << lane) - 1))] = e;
    }
    __syncthreads();
    const int n2 = s_n2;
    // ---- C: NMS, survivors into the bitmaps ----
    for (int k = tid; k < n2; k += FAST_NT)
    {
        const int y = list2[k] >> 8, x = list2[k] & 255, xa = x + a;
        const int s = score[y][xa];
        // cv response is s-1 and must beat 0 for empty neighbours: s >= 2
        if (s > 1 && s > score[y - 1][xa - 1] && s > score[y - 1][xa] && s > score[y - 1][xa + 1] && s > score[y][xa - 1] &&
            s > score[y][xa + 1] && s > score[y + 1][xa - 1] && s > score[y + 1][xa] && s > score[y + 1][xa + 1])
        {
            const int i = (y - 3) * zw + (x - 3);
            atomicOr(&bm[0][i >> 5], 1u << (i & 31));
            if (s > P.th_ini)
            {
                atomicOr(&bm[1][i >> 5], 1u << (i & 31));
                s_ini = 1;
            }
        }
    }
    __syncthreads();
    // ---- threshold choice ("if (cv_KeysCell.empty()) retry with th_fast_min") and row-major ranks ----
    const uint32_t* B = bm[s_ini ? 1 : 0];
    const int nwords  = (nz + 31) >> 5;  // <= FAST_BMW <= FAST_NT
    {
        const int cnt = (tid < nwords) ? __popc(B[tid]) : 0;
        const int inc = warp_incl_scan(cnt, lane);
        if (lane == 31) s_warp[wid] = inc;
        __syncthreads();
        int add = 0;
#pragma unroll
        for (int k = 0; k < FAST_NT / 32; ++k)
            if (k < wid) add += s_warp[k];
        if (tid < nwords) pre[tid] = (uint16_t)(add + inc - cnt);
        if (tid == FAST_NT - 1) pre[FAST_BMW] = (uint16_t)(add + inc);  // total
    }
    __syncthreads();
    const size_t slot = (size_t)img * P.cells_per_img + cell0 + kc;
    uint32_t* out     = cand + slot * P.cell_cap;
    for (int k = tid; k < n2; k += FAST_NT)
    {
        const int y = list2[k] >> 8, x = list2[k] & 255;
        const int i = (y - 3) * zw + (x - 3);
        const uint32_t word = B[i >> 5], bit = 1u << (i & 31);
        const int sc = score[y][x + a];
        score[y][x + a] = 0;  // list2 holds every entry phase B wrote: the tile is zero again for the next cell
        if (word & bit)
        {
            const int pos = pre[i >> 5] + __popc(word & (bit - 1));
            if (pos < P.cell_cap)
                out[pos] = (uint32_t)(x + c.off_x) | ((uint32_t)(y + c.off_y) << 12) | ((uint32_t)(sc - 1) << 24);
        }
    }
    if (tid == 0) cell_count[slot] = pre[FAST_BMW];
    c = cnext;
    __syncthreads();  // every read of this cell's tile, lists and bitmaps is done before they are reused
    }  // cells of this CTA
}

}  // namespace snk
#include "fast_group.cuh"
namespace snk
{
// --------------------------------------------------------------------------------------------------------------
// Quadtree distribution
// --------------------------------------------------------------------------------------------------------------
struct QNode
{
    float cx, cy, sx, sy;  // corner, size
    int from, to;
};
constexpr int QT_NT = 256;

__device__ __forceinline__ float kp_x(uint32_t p) { return (float)(p & 0xfffu); }
__device__ __forceinline__ float kp_y(uint32_t p) { return (float)((p >> 12) & 0xfffu); }

// dynamic shared memory: 2 node lists + leaf positions + sort keys + per-node child counts / positions = QT_NODE_BYTES
// per node (a leaf holds one keypoint: only its position in the keypoint array is kept). 76 bytes per node and 2 x 2048
// keypoints let six blocks share an SM at the BASELINE quotas.
constexpr int QT_NODE_BYTES = 2 * 24 + 4 + 4 + 16 + 4;
constexpr int QT_SMALL = 24;      // nodes up to this many keypoints are processed by a single thread
constexpr int QT_KP_SMEM = 2048;  // keypoints per (level, image) kept in shared memory (x2: array + partition buffer)
// the largest levels get their own launch with a larger capacity: camera noise of a few grey levels lifts level 0 of a
// 752x480 image above 2048 candidates, and the global-scratch path costs about 1.5x per candidate
constexpr int QT_KP_SMEM_BIG = 8192;
__global__ void __launch_bounds__(QT_NT) k_distribute(PyrDev P, const uint32_t* __restrict__ cand,
                                                      const int* __restrict__ cell_count, int* __restrict__ cell_off,
                                                      uint32_t* __restrict__ workA, uint32_t* __restrict__ workB,
                                                      uint32_t* __restrict__ sel, int* __restrict__ sel_count,
                                                      int node_cap, int kp_cap, int* __restrict__ err, int pass, int small_cap, int big_levels)
{
    extern __shared__ __align__(16) unsigned char smem[];
    QNode* listA    = reinterpret_cast<QNode*>(smem);
    QNode* listB    = listA + node_cap;
    uint32_t* leaves = reinterpret_cast<uint32_t*>(listB + node_cap);  // position of the leaf's keypoint
    uint32_t* keys  = leaves + node_cap;
    int* ccnt_big   = reinterpret_cast<int*>(keys + node_cap);           // child sizes of sorted node i (4 per node)
    short2* cpos    = reinterpret_cast<short2*>(ccnt_big + 4 * node_cap);  // (leaf position, next position) of node i (<= node_cap + 4 each)
    __shared__ int s_warp[QT_NT / 32], s_ctrl[10];

    const int level = blockIdx.y, img = blockIdx.x;  // all level-0 blocks (the longest) are dispatched first
    const LevelDev& L = P.lv[level];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const size_t wbase = (size_t)img * P.cand_per_img + L.cand_base;
    // keypoint array and its partition buffer: in shared memory when the level's candidates fit (the split loop walks
    // them node by node, a chain of dependent accesses), else in the global scratch (always large enough)
    uint32_t* A = reinterpret_cast<uint32_t*>(cpos + node_cap);
    uint32_t* B = A + kp_cap;
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
    // two launches share the (level, image) problems: pass 1 (shared memory for small_cap keypoints, six blocks per SM)
    // takes those with at most small_cap candidates, pass 2 (a larger array, fewer blocks per SM) the dense rest — on
    // ordinary images its blocks end here; pass 0 takes everything
    if ((pass == 1 && level < big_levels && n > small_cap) || (pass == 2 && n <= small_cap)) return;
    if (n > kp_cap)
    {
        A = workA + wbase;
        B = workB + wbase;
    }
    // one thread per cell (cells hold a handful of candidates; the per-cell loads of different threads overlap)
    for (int cidx = tid; cidx < L.n_cells; cidx += QT_NT)
    {
        const int cnt = min(cc[cidx], P.cell_cap), off = co[cidx];
        // cell lists are 16-byte aligned and cell_cap is a multiple of 4: four candidates per load, all loads of a cell
        // independent of each other
        const uint4* s = reinterpret_cast<const uint4*>(cand + ((size_t)img * P.cells_per_img + L.cell_base + cidx) * P.cell_cap);
        for (int k = 0; k < cnt; k += 4)
        {
            const uint4 v = __ldg(s + (k >> 2));
            A[off + k] = v.x;
            if (k + 1 < cnt) A[off + k + 1] = v.y;
            if (k + 2 < cnt) A[off + k + 2] = v.z;
            if (k + 3 < cnt) A[off + k + 3] = v.w;
        }
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
        // the whole block sorts (same permutation as the serial code, gnu_sort_cta.cuh); its scratch lives in the
        // child-count / position arrays, which are only written after the sort
        gnusort::cta_sort<uint32_t, gnusort::DescCount, QT_NT>(keys, m, reinterpret_cast<unsigned char*>(ccnt_big), tid);
        if (tid == 0)
        {
            // nodes are now ordered by size: [0, nbig) are handled by a warp each, the rest by a thread each
            int lo = 0, hi = m;
            while (lo < hi)
            {
                const int mid = (lo + hi) >> 1;
                if ((int)(keys[mid] >> gnusort::SHIFT) > QT_SMALL)
                    lo = mid + 1;
                else
                    hi = mid;
            }
            s_ctrl[8] = lo;
        }
        __syncthreads();
        const int nbig = s_ctrl[8];

        // child sizes of every inner node (the split itself is only applied to the nodes that get processed)
        for (int i = nbig + tid; i < m; i += QT_NT)
        {
            const QNode nd = inner[keys[i] & ((1u << gnusort::SHIFT) - 1)];
            const float mx = __fadd_rn(nd.cx, __fmul_rn(nd.sx, 0.5f)), my = __fadd_rn(nd.cy, __fmul_rn(nd.sy, 0.5f));
            int cq[4] = {0, 0, 0, 0};
            for (int k = nd.from; k < nd.to; ++k)
            {
                const uint32_t p = A[k];
                const int q      = ((kp_y(p) > my) ? 2 : 0) + ((kp_x(p) > mx) ? 1 : 0);
                cq[0] += (q == 0), cq[1] += (q == 1), cq[2] += (q == 2), cq[3] += (q == 3);
            }
            ccnt_big[4 * i + 0] = cq[0], ccnt_big[4 * i + 1] = cq[1], ccnt_big[4 * i + 2] = cq[2], ccnt_big[4 * i + 3] = cq[3];
        }
        for (int i = wid; i < nbig; i += QT_NT / 32)
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

        // bookkeeping of the split loop (:127-159): where each node's children land and where it stops. The serial
        // loop's running counts are prefix sums, its break is the first node that satisfies the finish rule: one warp
        // scans 32 nodes at a time (leaf count in the low, inner count in the high half of one word)
        if (wid == 0)
        {
            int nl = n_leaves, nn = 0, lpi = -1;
            for (int base = 0; base < m && lpi < 0; base += 32)
            {
                const int i = base + lane;
                int d = 0;
                if (i < m)
                {
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                    {
                        const int c = ccnt_big[4 * i + q];
                        d += (c == 1) + ((c > 1) << 16);
                    }
                }
                const int inc = warp_incl_scan(d, lane);
                const int il = inc & 0xffff, in = inc >> 16;
                if (i < m) cpos[i] = make_short2((short)min(nl + il - (d & 0xffff), 32767), (short)min(nn + in - (d >> 16), 32767));
                const bool hit    = (i < m) && (nl + il + nn + in + (m - i - 1) >= target);
                const unsigned hm = __ballot_sync(0xffffffffu, hit);
                const int f       = hm ? (__ffs(hm) - 1) : 31;  // the node the serial loop stops at, else the chunk's last
                if (hm) lpi = base + f;
                nl += __shfl_sync(0xffffffffu, il, f);
                nn += __shfl_sync(0xffffffffu, in, f);
            }
            if (lane == 0)
            {
                if (nl > node_cap || nn > node_cap) atomicExch(err, 2);
                s_ctrl[0] = nn;
                s_ctrl[1] = nl;
                s_ctrl[4] = (lpi >= 0);
                if (lpi >= 0) s_ctrl[3] = lpi;
                s_ctrl[7] = (lpi >= 0) ? lpi + 1 : m;
            }
        }
        __syncthreads();
        const int processed = s_ctrl[7];

        // splitAndSort (:15-70) of the processed nodes: stable 4-way partition, children appended in order
        for (int i = nbig + tid; i < processed; i += QT_NT)
        {
            // small node: the whole stable partition and the child bookkeeping by one thread
            const QNode nd = inner[keys[i] & ((1u << gnusort::SHIFT) - 1)];
            const float hx = __fmul_rn(nd.sx, 0.5f), hy = __fmul_rn(nd.sy, 0.5f);
            const float mx = __fadd_rn(nd.cx, hx), my = __fadd_rn(nd.cy, hy);
            const int cs[4] = {ccnt_big[4 * i], ccnt_big[4 * i + 1], ccnt_big[4 * i + 2], ccnt_big[4 * i + 3]};
            int o[4];
            o[0] = nd.from, o[1] = o[0] + cs[0], o[2] = o[1] + cs[1], o[3] = o[2] + cs[2];
            for (int k = nd.from; k < nd.to; ++k)
            {
                const uint32_t p = A[k];
                const int q      = ((kp_y(p) > my) ? 2 : 0) + ((kp_x(p) > mx) ? 1 : 0);
                const int dst    = (q == 0) ? o[0]++ : (q == 1) ? o[1]++ : (q == 2) ? o[2]++ : o[3]++;
                B[dst]           = p;
            }
            for (int k = nd.from; k < nd.to; ++k) A[k] = B[k];
            int2 pos = make_int2(cpos[i].x, cpos[i].y);
            int from = nd.from;
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
                    if (pos.x < node_cap) leaves[pos.x] = (uint32_t)ch.from;
                    ++pos.x;
                }
                else if (c > 1)
                {
                    if (pos.y < node_cap) next[pos.y] = ch;
                    ++pos.y;
                }
            }
        }
        for (int i = wid; i < min(nbig, processed); i += QT_NT / 32)
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
                int2 pos = make_int2(cpos[i].x, cpos[i].y);
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
                        if (pos.x < node_cap) leaves[pos.x] = (uint32_t)ch.from;
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
    // small nodes (all leaves, most others): one thread finds the first maximum serially
    for (int o = tid; o < total; o += QT_NT)
    {
        QNode nd;
        if (o < n_leaves)
        {
            if (o < out_cap) out[o] = A[leaves[o]];
            continue;
        }
        else if (o < n_leaves + n_next)
            nd = next[o - n_leaves];
        else
            nd = inner[keys[lpi + (o - n_leaves - n_next)] & ((1u << gnusort::SHIFT) - 1)];
        if (nd.to - nd.from > QT_SMALL) continue;
        int bk = nd.from;
        uint32_t br = A[bk] >> 24;
        for (int k = nd.from + 1; k < nd.to; ++k)
        {
            const uint32_t r = A[k] >> 24;
            if (br < r) br = r, bk = k;
        }
        if (o < out_cap) out[o] = A[bk];
    }
    for (int o = wid; o < total; o += QT_NT / 32)
    {
        QNode nd;
        if (o < n_leaves)
            continue;  // leaves hold one keypoint
        else if (o < n_leaves + n_next)
            nd = next[o - n_leaves];
        else
            nd = inner[keys[lpi + (o - n_leaves - n_next)] & ((1u << gnusort::SHIFT) - 1)];
        if (nd.to - nd.from <= QT_SMALL) continue;
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
//
// One warp per 128 x BL_ROWS strip, no shared memory. A lane owns 4 adjacent pixels (one aligned 32-bit word) and
// walks down the rows: it loads its word, gets the two neighbouring words by shuffle, forms the four 7-tap horizontal
// sums with byte permutes + dp4a (the taps fit in a byte), and keeps the last 7 rows of those sums in registers; the
// vertical pass runs on that register window. ~15 instructions per pixel; every source byte is loaded once per strip.
// Requires 4-byte aligned rows (internal buffers are; an unaligned caller buffer is copied first).
// --------------------------------------------------------------------------------------------------------------
constexpr int BL_TW = 128, BL_GROUPS = 8, BL_ROWS = 7 * BL_GROUPS - 6;  // 50 output rows per warp task (measured: 4 groups 0.69 ms, 6: 0.65, 8: 0.62, 9: 0.63, 12: 0.65)
struct TileDev
{
    short level, tx, ty, _pad;
};

__device__ __forceinline__ int reflect101(int p, int len)
{
    // valid for -len < p < 2*len-1 (halo of 3..4, levels are far larger)
    p = p < 0 ? -p : p;
    return p >= len ? 2 * (len - 1) - p : p;
}

// byte selector that rebuilds one word of the 12-byte tap window [x-4, x+8) of a lane at an image border.
// `first` is the pixel of the word's byte 0, `base` the pixel of byte 0 of the 8-byte register pair the selector reads
// from; pixels outside [0, W) are replaced by their reflect-101 partner. Out-of-pair sources are clamped: they can
// only be asked for by taps of outputs that are themselves outside the image and never stored.
__device__ __forceinline__ unsigned reflect_selector(int first, int base, int W)
{
    unsigned sel = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k)
    {
        const int src = reflect101(min(max(first + k, -(W - 1)), 2 * W - 2), W) - base;
        sel |= (unsigned)min(max(src, 0), 7) << (4 * k);
    }
    return sel;
}

__global__ void __launch_bounds__(256) k_blur(PyrDev P, const TileDev* __restrict__ tiles, int n_tiles)
{
    const int lane = threadIdx.x & 31;
    const int task = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (task >= n_tiles) return;
    const TileDev t   = tiles[task];
    const LevelDev& L = P.lv[t.level];
    const uint8_t* S  = L.img + (size_t)blockIdx.y * L.img_stride;
    uint8_t* D        = L.gauss + (size_t)blockIdx.y * L.gauss_stride;
    const int W = L.w, H = L.h, pitch = L.pitch;
    const int x0 = t.tx * BL_TW, y0 = t.ty * BL_ROWS;
    const int x  = x0 + 4 * lane;               // first of my 4 pixels
    const int words_in_row = pitch >> 2;        // addressable words of a row
    const int xw = x >> 2;
    const bool own_ok = xw < words_in_row;      // my word can be read
    const bool active = x < W;                  // I produce at least one output pixel
    // Borders without a slow path: the 12-byte window [wl wc wr] of a lane near the left / right image border is
    // rebuilt from the register pair that holds the reflected pixels, one byte permute per word.
    //   left  border (x == 0):          wl <- reflection of pixels -4..-1, read from (wc, wr)
    //   right border (x+4 > W):         wc, wr <- read from (wl, wc)   (my own word is cut by the border)
    //   right border (x+4 <= W < x+8):  wr <- read from (wc, wr)       (the next word is cut by the border)
    const bool fixl = (x == 0), fix_own = active && (x + 4 > W), fix_next = (x + 4 <= W) && (x + 8 > W);
    const unsigned sel_l = reflect_selector(-4, 0, W);
    const unsigned sel_c = fix_own ? reflect_selector(x, x - 4, W) : 0x7654u;
    const unsigned sel_r = fix_own ? reflect_selector(x + 4, x - 4, W) : (fix_next ? reflect_selector(x + 4, x, W) : 0x7654u);
    const unsigned KLO = 0x38302212u;  // horizontal taps 18,34,48,56 (little-endian bytes)
    const unsigned KHI = 0x00122230u;  // horizontal taps 48,34,18,0
    const unsigned V01 = 0x2212u, V23 = 0x3830u, V45 = 0x2230u;  // vertical taps as byte pairs (18,34) (48,56) (48,34)

    // vertical window: q[s][j] = h[r-1] | h[r] << 16 of the row r with r mod 7 == s (horizontal sums fit 16 bits), so
    // the 7-tap vertical sum is three dp2a on the pairs (r-6,r-5) (r-4,r-3) (r-2,r-1) plus 18*h[r]
    unsigned q[7][4], hprev[4] = {0u, 0u, 0u, 0u};
#pragma unroll 1
    for (int g = 0; g < BL_GROUPS; ++g)
    {
        // issue the loads of all 7 rows of the group before using any of them (memory-level parallelism)
        unsigned wcs[7], hal[7];
#pragma unroll
        for (int i = 0; i < 7; ++i)
        {
            const int gy = reflect101(min(y0 - 3 + 7 * g + i, H + 2), H);
            const uint32_t* row = reinterpret_cast<const uint32_t*>(S + (size_t)gy * pitch);
            wcs[i] = own_ok ? __ldg(row + xw) : 0u;
            hal[i] = 0u;
            if (lane == 0 && xw > 0) hal[i] = __ldg(row + xw - 1);
            if (lane == 31 && xw + 1 < words_in_row) hal[i] = __ldg(row + xw + 1);
        }
#pragma unroll
        for (int i = 0; i < 7; ++i)
        {
            const int r = 7 * g + i;  // 0 .. 7*BL_GROUPS-1, source row y0 - 3 + r
            unsigned wc = wcs[i];
            unsigned wl = __shfl_up_sync(0xffffffffu, wc, 1);
            unsigned wr = __shfl_down_sync(0xffffffffu, wc, 1);
            if (lane == 0) wl = hal[i];
            if (lane == 31) wr = hal[i];
            {
                const unsigned nl = __byte_perm(wc, wr, sel_l);
                const unsigned nc = __byte_perm(wl, wc, sel_c);
                const unsigned nr = fix_own ? __byte_perm(wl, wc, sel_r) : __byte_perm(wc, wr, sel_r);
                wl = fixl ? nl : wl;
                wc = nc;
                wr = nr;
            }
            // pixel j of my word: taps are bytes (j+1 .. j+7) of the 12-byte window [wl wc wr]
            unsigned hc[4];
            hc[0] = __dp4a(__byte_perm(wc, wr, 0x4321), KHI, __dp4a(__byte_perm(wl, wc, 0x4321), KLO, 0u));
            hc[1] = __dp4a(__byte_perm(wc, wr, 0x5432), KHI, __dp4a(__byte_perm(wl, wc, 0x5432), KLO, 0u));
            hc[2] = __dp4a(__byte_perm(wc, wr, 0x6543), KHI, __dp4a(__byte_perm(wl, wc, 0x6543), KLO, 0u));
            hc[3] = __dp4a(wr, KHI, __dp4a(wc, KLO, 0u));
            // output row y0 + r - 6 is complete once row r is known
            const int oy = y0 + r - 6;
            if ((g > 0 || i == 6) && oy < H && active)
            {
                unsigned packed = 0;
#pragma unroll
                for (int j = 0; j < 4; ++j)
                {
                    unsigned v = 18u * hc[j] + 32768u;
                    v          = __dp2a_lo(q[(i + 2) % 7][j], V01, v);  // rows r-6, r-5
                    v          = __dp2a_lo(q[(i + 4) % 7][j], V23, v);  // rows r-4, r-3
                    v          = __dp2a_lo(q[(i + 6) % 7][j], V45, v);  // rows r-2, r-1
                    packed |= (v >> 16) << (8 * j);
                }
                uint8_t* drow = D + (size_t)oy * L.gauss_pitch + x;
                if (x + 4 <= W)
                    *reinterpret_cast<uint32_t*>(drow) = packed;
                else
                    for (int j = 0; j < 4 && x + j < W; ++j) drow[j] = (uint8_t)(packed >> (8 * j));
            }
#pragma unroll
            for (int j = 0; j < 4; ++j)
            {
                q[i][j]  = __byte_perm(hprev[j], hc[j], 0x5410);  // h[r-1] | h[r] << 16
                hprev[j] = hc[j];
            }
        }
    }
}

// --------------------------------------------------------------------------------------------------------------
// Orientation + rBRIEF + output record. One warp per selected keypoint.
// --------------------------------------------------------------------------------------------------------------

constexpr int DS_WARPS = 8;
constexpr int DS_KP_PER_WARP = 8;  // keypoints per warp: the CTA prologue (pattern, level prefix) is paid once per 64
// Moment table of ORB::ComputeAngle for word-wise evaluation: entry [al][row][w] (al = (x - 15) & 3, row = v + 15,
// w = aligned word 0..8 of the 36-byte window that starts at the aligned column below x - 15) holds
//   .x  the four column offsets u of the word's bytes as signed bytes, 0 outside the disc  (m10 += dp4a(word, .x))
//   .y  1 per byte inside the disc                                          (m01 += v * dp4a(word, .y))
constexpr int DS_MOM_WORDS = 9, DS_MOM_TASKS = 31 * DS_MOM_WORDS;  // 279 word tasks per keypoint
constexpr int DS_MOM_TABLE = 4 * DS_MOM_TASKS;                     // uint2 entries

// sum of (unsigned byte of a) * (signed byte of b) + c
__device__ __forceinline__ int dp4a_u8s8(unsigned a, unsigned b, int c)
{
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

__global__ void __launch_bounds__(DS_WARPS * 32) k_describe(PyrDev P, const uint32_t* __restrict__ sel,
                                                            const int* __restrict__ sel_count,
                                                            snk_keypoint_f32* __restrict__ kps,
                                                            uint64_t* __restrict__ desc, int* __restrict__ counts,
                                                            const float2* __restrict__ pattern_t,
                                                            const uint2* __restrict__ moment_tab)
{
    __shared__ float2 pat[16][32];  // pat[j][lane] = point 16*lane + j as floats (conflict-free: lanes are consecutive)
    __shared__ int s_pref[SNK_MAX_LEVELS + 1];
    const int img = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    // pattern_t is the rBRIEF table already transposed to [j][lane] and converted to float (built once per handle)
    for (int i = tid; i < 512; i += DS_WARPS * 32) (&pat[0][0])[i] = __ldg(pattern_t + i);
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
    // the warp's keypoints: g(kk) = (blockIdx.x * DS_KP_PER_WARP + kk) * DS_WARPS + wid, kk < nk
    const int g0 = blockIdx.x * DS_KP_PER_WARP * DS_WARPS + wid;
    if (g0 >= total) return;
    const int nk = min(DS_KP_PER_WARP, (total - g0 + DS_WARPS - 1) / DS_WARPS);

    // ---- per keypoint (lane kk owns keypoint kk): level and packed position ----
    int my_level = 0;
    uint32_t my_p = 0;
    if (lane < nk)
    {
        const int g = g0 + lane * DS_WARPS;
        while (my_level + 1 < P.nlevels && g >= s_pref[my_level + 1]) ++my_level;
        my_p = sel[(size_t)img * P.sel_per_img + P.lv[my_level].sel_base + (g - s_pref[my_level])];
    }

    // ---- ORB::ComputeAngle: integer moments over the radius-15 disc of the un-blurred level ----
    // word-wise: 279 aligned words cover the 31 x 31 patch; lane takes tasks lane, lane + 32, ... (9 per lane)
    int my_m10 = 0, my_m01 = 0;
#pragma unroll 2
    for (int kk = 0; kk < nk; ++kk)
    {
        const int level   = __shfl_sync(0xffffffffu, my_level, kk);
        const uint32_t p  = __shfl_sync(0xffffffffu, my_p, kk);
        const LevelDev& L = P.lv[level];
        // kp.point += (minBorderX, minBorderY)   ORBExtractor.cpp:165-166
        const int x = (int)(p & 0xfffu) + 16, y = (int)((p >> 12) & 0xfffu) + 16;
        const int pitch = L.pitch;
        const int al    = (x - 15) & 3;
        // offsets inside one level image fit 32 bits (levels are at most 4000 x 4000)
        const uint8_t* base = L.img + (size_t)img * L.img_stride + ((y - 15) * pitch + (x - 15 - al));
        const uint2* tab    = moment_tab + al * DS_MOM_TASKS;
        int m10 = 0, m01 = 0;
#pragma unroll
        for (int it = 0; it < (DS_MOM_TASKS + 31) / 32; ++it)
        {
            const int t = lane + 32 * it;
            if (t < DS_MOM_TASKS)
            {
                const int r = (t * 7282) >> 16;  // t / 9 for t < 279
                const int w = t - 9 * r;
                const unsigned word = __ldg(reinterpret_cast<const unsigned*>(base + r * pitch) + w);
                const uint2 e       = __ldg(tab + t);
                m10 = dp4a_u8s8(word, e.x, m10);                         // u8 pixels x s8 column offsets
                m01 += (r - 15) * (int)__dp4a(word, e.y, 0u);            // v * sum of the row's pixels inside the disc
            }
        }
        m10 = warp_sum(m10);
        m01 = warp_sum(m01);
        if (lane == kk) my_m10 = m10, my_m01 = m01;
    }

    // ---- angle, sine and cosine of all keypoints of the warp at once (lane kk computes keypoint kk) ----
    // degrees(atan2f(m01, m10)): a * 180.f / pi_f, then wrap negatives by +360 (OrbDescriptors.cpp:52-53)
    float my_angle = snkmath::atan2f_fdlibm((float)my_m01, (float)my_m10);
    my_angle       = __fdiv_rn(__fmul_rn(my_angle, 180.0f), 3.14159265358979323846f);
    my_angle       = __fadd_rn((my_angle < 0.f) ? 360.f : 0.f, my_angle);
    const float my_rad = __fmul_rn(__fdiv_rn(my_angle, 180.0f), 3.14159265358979323846f);
    const float my_a = snkmath::cosf_glibc(my_rad), my_b = snkmath::sinf_glibc(my_rad);

    // ---- ORB::ComputeDescriptor on the blurred level; lane i produces byte i ----
#pragma unroll 1
    for (int kk = 0; kk < nk; ++kk)
    {
    const int g       = g0 + kk * DS_WARPS;
    const int level   = __shfl_sync(0xffffffffu, my_level, kk);
    const uint32_t p  = __shfl_sync(0xffffffffu, my_p, kk);
    const float angle = __shfl_sync(0xffffffffu, my_angle, kk);
    const float a     = __shfl_sync(0xffffffffu, my_a, kk);
    const float b     = __shfl_sync(0xffffffffu, my_b, kk);
    const LevelDev& L = P.lv[level];
    const int x = (int)(p & 0xfffu) + 16, y = (int)((p >> 12) & 0xfffu) + 16;
    const float resp = (float)(p >> 24);
    const int gpitch  = L.gauss_pitch;
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
            const float2 q = pat[2 * j + e][lane];
            const float fx = __fadd_rn(px, __fsub_rn(__fmul_rn(q.x, a), __fmul_rn(q.y, b)));
            const float fy = __fadd_rn(py, __fadd_rn(__fmul_rn(q.x, b), __fmul_rn(q.y, a)));
            // Saiga::iRound: (int)(v + (v >= 0 ? 0.5f : -0.5f)); the half takes the sign bit of v (v = -0.f gives
            // -0.5f instead of +0.5f, both truncate to 0)
            const int ix = (int)__fadd_rn(fx, __int_as_float(0x3f000000 | (__float_as_int(fx) & 0x80000000)));
            const int iy = (int)__fadd_rn(fy, __int_as_float(0x3f000000 | (__float_as_int(fy) & 0x80000000)));
            tv[e]        = gs[iy * gpitch + ix];
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
    }  // keypoints of this warp
}
}  // namespace snk

#include "blur_mma.cuh"
#include "blur_umma.cuh"
#include "describe_tma.cuh"
