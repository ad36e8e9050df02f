// cv::GaussianBlur(7 x 7, sigma 2, BORDER_REFLECT_101) of u8 images on the 5th-generation tensor cores: tcgen05.mma
// kind::i8, operands by TMA, accumulators in tensor memory (ORBExtractor.cpp:213; semantics as k_blur / k_blur_mma).
// The default blur since round 2 (SNK_BLUR_MMA=1 selects k_blur_mma, 0 k_blur); byte-equal to the oracle on every shape of
// tests/test_gpu_extract.py. 0.356 ms per 512 EuRoC images against 0.585 ms (mma.sync) and 0.62 ms (dp4a).
//
// Both passes of the fixed-point blur are banded matrix products with u8 weights (56 48 34 18, sum 256):
//   pass 1 (rows):     R[x][y]  = sum_k  Wc[x][k] * I[y][k]        Wc: band over image columns, reflect-101 folded in
//   pass 2 (columns):  O[x][y'] = sum_y  R[x][y] * Wr[y'][y]       Wr: band over image rows
//   out = (O + 32768) >> 16
// Pass 1 is computed TRANSPOSED (M = output columns on the TMEM lanes, N = image rows): the image tile is then the B
// operand exactly as it lies in memory (row = N index, columns = K, contiguous) and arrives by TMA with the swizzle the
// MMA wants; and the result has the image rows along the TMEM columns, i.e. along the K dimension of an A operand read
// from tensor memory — pass 2 takes it from there without a trip through shared memory. R is 16 bits wide, the operands
// are 8: the epilogue of pass 1 reads the accumulators (tcgen05.ld, two 16-bit values per register), splits them into low
// and high bytes (one PRMT each per four rows) and writes two u8 A operands back (tcgen05.st); pass 2 runs once on each,
// O = O_lo + 256 O_hi. The rounding constant rides along: the last input row of a tile lies outside every band, its high
// byte is set to 128 and its band weight to 1 (128 * 256 = 32768).
// Per tile (128 x 48 output pixels from 160 x 64 input pixels):
//   two TMA boxes (128 columns with the 128-byte swizzle, 32 with the 32-byte one) -> 5 MMAs (M 128, N 64, K 32) -> ld /
//   split / st -> 2 x 2 MMAs (M 128, N 48, K 32) -> ld, combine, bytes into a shared-memory tile -> TMA store (clipped at
//   the image's sides by the tensor map).
// TMEM: 128 columns per CTA, so that FOUR CTAs share an SM — the per-tile chain (MMA -> epilogue -> MMA -> epilogue) is
// latency, and tiles in flight are what hides it; the columns are reused inside a tile: [0,64) pass-1 accumulator, then
// [0,16) R_lo, [16,32) R_hi (u8 x 4), [32,80) O_lo, [80,128) O_hi.
// Persistent CTAs walk equal shares of the tile list (level by level, image by image, strip by strip, top down); warp 4's lane 0 issues the
// TMA loads (two tiles ahead) and the MMAs and waits only on mbarriers, warps 0-3 are the epilogue (a thread = an output
// column = a TMEM lane). The band matrices come from a table built once per pyramid geometry (k_umma_bands) and are
// copied in (cp.async) when the strip / the row-tile kind changes.
// Steps on the way (B200, 512 EuRoC images, all byte-equal): one CTA per strip with sequential phases 0.68 ms; warp-
// specialised 0.68; persistent over a tile list 0.70 (256 TMEM columns, two CTAs per SM: the chain of a tile is ~4400
// cycles whatever the schedule); 128 TMEM columns, four CTAs per SM 0.49; packed TMEM loads, one barrier per tile 0.47;
// 128 instead of 96 output columns per tile 0.39; tiles walked image by image (halo rows and columns from L2) 0.356.
#pragma once
#include <cuda.h>

#include <cstdint>

#include "blur_weights.cuh"

namespace snk
{
constexpr int UB_M        = 128;  // MMA M: TMEM lanes = output columns of the tile
constexpr int UB_OUT_COLS = 128;  // output columns per tile [x0, x0 + 128), x0 a multiple of 16 (TMA coordinates)
constexpr int UB_K        = 160;  // input columns per tile [x0 - 16, x0 + 144): the band needs [x0 - 3, x0 + 131); as two boxes,
constexpr int UB_K0       = 128;  //   128 columns with the 128-byte swizzle (four K slices of 32)
constexpr int UB_K1       = 32;   //   and 32 columns with the 32-byte swizzle (the fifth)
constexpr int UB_ROWS_IN  = 64;   // input rows per tile = N of pass 1 = K of pass 2: [y0 - 3, y0 + 61)
constexpr int UB_ROWS_OUT = 48;   // output rows per tile = N of pass 2 (multiple of 16, <= UB_ROWS_IN - 6 - 1: the last row carries the rounding constant)
constexpr int UB_EPI      = 128;  // 4 epilogue warps: warp w reads TMEM lanes 32 w .. + 31, a thread is one output column
constexpr int UB_NT       = UB_EPI + 32;  // + one warp whose lane 0 issues the TMA loads and the MMAs
constexpr int UB_CTAS_PER_SM = 4;
// 128 TMEM columns per CTA, four CTAs per SM. [0, 64): pass-1 accumulator D1; once the epilogue holds it in registers the
// columns are reused: [0, 16) R_lo and [16, 32) R_hi as u8 x 4 (A operands of pass 2), [32, 80) O_lo, [80, 128) O_hi
constexpr int UB_TMEM_COLS = 128;
constexpr int UB_COL_D1 = 0, UB_COL_ALO = 0, UB_COL_AHI = UB_ROWS_IN / 4, UB_COL_OLO = UB_ROWS_IN / 2, UB_COL_OHI = UB_ROWS_IN / 2 + UB_ROWS_OUT;
static_assert(UB_COL_OHI + UB_ROWS_OUT <= UB_TMEM_COLS && UB_ROWS_OUT + 7 <= UB_ROWS_IN && UB_ROWS_IN % 32 == 0 && UB_ROWS_OUT % 16 == 0, "tile shape");
// shared memory (operand tiles 1024-byte aligned for the swizzles)
constexpr int UB_IMG_BYTES = UB_ROWS_IN * (UB_K0 + UB_K1);      // an image tile: UB_ROWS_IN rows x 128 B, then UB_ROWS_IN rows x 32 B
constexpr int UB_WC_BYTES  = UB_M * (UB_K0 + UB_K1);            // column band: 128 rows (m) x 128 B, then 128 rows x 32 B
constexpr int UB_WR_BYTES  = UB_ROWS_OUT * 128;                 // row band: UB_ROWS_OUT rows (n) x 128 B (k, UB_ROWS_IN used)
constexpr int UB_OFF_IMG0 = 0;
constexpr int UB_OFF_IMG1 = UB_IMG_BYTES;
constexpr int UB_OFF_WC   = 2 * UB_IMG_BYTES;
constexpr int UB_OFF_WR   = UB_OFF_WC + UB_WC_BYTES;
constexpr int UB_OFF_OUT  = UB_OFF_WR + UB_WR_BYTES;            // output tile: UB_ROWS_OUT rows x UB_OUT_COLS B, dense
constexpr int UB_OFF_BAR  = UB_OFF_OUT + UB_ROWS_OUT * UB_OUT_COLS;  // six mbarriers, the TMEM address
constexpr int UB_SMEM     = UB_OFF_BAR + 64 + 1024;  // + alignment slack
static_assert(UB_OFF_IMG1 % 1024 == 0 && UB_OFF_WC % 1024 == 0 && UB_OFF_WR % 1024 == 0 && UB_OFF_OUT % 1024 == 0 &&
                  (UB_ROWS_IN * UB_K0) % 1024 == 0 && (UB_M * UB_K0) % 1024 == 0,
              "swizzle alignment");

struct UmmaBlurMaps
{
    CUtensorMap src[8];   // per level: (x, y, image) u8, box UB_K0 x UB_ROWS_IN x 1, SWIZZLE_128B
    CUtensorMap srcx[8];  // the same tensor, box UB_K1 x UB_ROWS_IN x 1, SWIZZLE_32B
    CUtensorMap dst[8];   // per level: box UB_OUT_COLS x UB_ROWS_OUT x 1, no swizzle
};
struct UmmaBlurTask  // one column strip of a level
{
    short level, x0, w, h;
    short n_row_tiles, strip, n_strips, wc_interior;  // wc_interior: no tap of the strip is folded (all such strips share one band)
    unsigned wc_off;  // byte offset of the strip's column band in the band table
    unsigned wr_off;  // byte offset of the level's row bands (one per row tile, UB_WR_BYTES apart)
};
// The tiles of a launch in the order the CTAs walk them: level by level, inside a level image by image, inside an image
// strip by strip, inside a strip from the top down. CTA c of G takes the tiles [c n / G, (c + 1) n / G): every CTA the same
// number of tiles, the pipeline never drains between strips or images, and a tile's halo — the rows it shares with the
// tile above, the columns it shares with the strip to the left — was read by the same CTA moments ago (L2 hits).
struct UmmaTileIter
{
    int task, img, row;
    UmmaBlurTask T;
    __device__ void seek(const UmmaBlurTask* __restrict__ tasks, int n_imgs, long long g)
    {
        task = 0;
        T    = tasks[0];
        for (;;)  // levels
        {
            const long long level_tiles = (long long)T.n_strips * T.n_row_tiles;
            if (g < n_imgs * level_tiles)
            {
                img = (int)(g / level_tiles);
                const int r = (int)(g - img * level_tiles);
                task += r / T.n_row_tiles;
                row = r % T.n_row_tiles;
                T   = tasks[task];
                return;
            }
            g -= n_imgs * level_tiles;
            task += T.n_strips;
            T = tasks[task];
        }
    }
    __device__ void next(const UmmaBlurTask* __restrict__ tasks, int n_tasks, int n_imgs)
    {
        if (++row < T.n_row_tiles) return;
        row = 0;
        if (T.strip + 1 < T.n_strips)
        {
            T = tasks[++task];
            return;
        }
        task -= T.strip;  // first strip of the level
        if (++img == n_imgs)
        {
            img = 0;
            task += T.n_strips;  // next level
        }
        if (task < n_tasks) T = tasks[task];
    }
};

// byte offset of (row r, byte b) in a K-major operand tile with 128-byte rows and the 128-byte swizzle
__device__ __forceinline__ unsigned ub_sw128(int r, int b)
{
    return (unsigned)((r >> 3) * 1024 + (r & 7) * 128 + ((((b >> 4) ^ (r & 7)) << 4) | (b & 15)));
}
// the same for 32-byte rows and the 32-byte swizzle
__device__ __forceinline__ unsigned ub_sw32(int r, int b) { return (unsigned)(r * 32 + ((((b >> 4) ^ ((r >> 2) & 1)) << 4) | (b & 15))); }
// shared-memory matrix descriptor: K-major, SWIZZLE_128B, 8-row groups 1024 bytes apart, descriptor version 1 (sm_100)
__device__ __forceinline__ unsigned long long ub_desc(unsigned smem_addr)
{
    return (unsigned long long)((smem_addr & 0x3ffffu) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// K-major, SWIZZLE_32B, 8-row groups 256 bytes apart
__device__ __forceinline__ unsigned long long ub_desc32(unsigned smem_addr)
{
    return (unsigned long long)((smem_addr & 0x3ffffu) >> 4) | (1ull << 16) | (16ull << 32) | (1ull << 46) | (6ull << 61);
}
// instruction descriptor: kind::i8, u8 x u8 -> s32, both operands K-major, M = 128
__device__ __forceinline__ unsigned ub_idesc(int n) { return (2u << 4) | ((unsigned)(n >> 3) << 17) | ((unsigned)(UB_M >> 4) << 24); }

__device__ __forceinline__ void ub_mma_ss(unsigned d_tmem, unsigned long long a_desc, unsigned long long b_desc, unsigned idesc, unsigned acc)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc)
        : "memory");
}
__device__ __forceinline__ void ub_mma_ts(unsigned d_tmem, unsigned a_tmem, unsigned long long b_desc, unsigned idesc, unsigned acc)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
        "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(acc)
        : "memory");
}
__device__ __forceinline__ void ub_commit(unsigned bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// bounded wait on an mbarrier phase: a copy or an MMA that never completes is reported, not waited for forever
__device__ __forceinline__ bool ub_wait(unsigned bar, unsigned parity)
{
    unsigned done = 0;
    for (int spin = 0; !done && spin < (1 << 12); ++spin)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, 0x186a0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(bar), "r"(parity)
                     : "memory");
    return done != 0;
}
// 16 consecutive TMEM columns of this thread's lane
__device__ __forceinline__ void ub_ld16(unsigned taddr, unsigned (&v)[16])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}
// the low 16 bits of 32 consecutive TMEM columns of this thread's lane, two columns per register
__device__ __forceinline__ void ub_ld16_pack(unsigned taddr, unsigned (&v)[16])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.pack::16b.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void ub_st4(unsigned taddr, unsigned a, unsigned b, unsigned c, unsigned d)
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%2,%3,%4};" ::"r"(taddr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// The band matrices of every (level, column strip) and (level, row tile) in the shared-memory layout of the MMA operands
// (128-byte rows, 128-byte swizzle), built once per pyramid geometry; k_blur_umma copies them in.
//   column band Wc[m][k]: weight of input column x0 - 16 + k in output column x0 + m (m < UB_OUT_COLS, else 0)
//   row band    Wr[n][k]: weight of input row y0 - 3 + k in output row y0 + n; Wr[n][UB_ROWS_IN - 1] = 1 carries the
//                         rounding constant (the high-byte operand holds 128 there: 128 * 256 = 32768)
struct UmmaBandJob
{
    unsigned off;         // byte offset in the table
    short origin, len;    // x0 (or y0) and the level's width (height)
    short is_rows, _pad;
};
__global__ void k_umma_bands(const UmmaBandJob* __restrict__ jobs, unsigned char* __restrict__ table)
{
    const UmmaBandJob J = jobs[blockIdx.x];
    const int r = threadIdx.x;  // one band row per thread, 128 threads
    unsigned char* base = table + J.off;
    if (J.is_rows)
    {
        if (r >= UB_ROWS_OUT) return;
        for (int kw = 0; kw < 32; ++kw)
        {
            unsigned w = band_word(J.origin - 3 + 4 * kw, J.origin + r, J.len, false);
            // k = UB_ROWS_IN - 1 lies outside every band (k <= n + 6): the constant's row
            if (kw == (UB_ROWS_IN - 1) / 4) w |= 1u << (8 * ((UB_ROWS_IN - 1) & 3));
            *reinterpret_cast<unsigned*>(base + ub_sw128(r, 4 * kw)) = w;
        }
    }
    else
    {
        for (int kw = 0; kw < UB_K / 4; ++kw)
        {
            const unsigned w = band_word(J.origin - 16 + 4 * kw, J.origin + r, J.len, false);
            unsigned char* at = kw < UB_K0 / 4 ? base + ub_sw128(r, 4 * kw) : base + UB_M * UB_K0 + ub_sw32(r, 4 * kw - UB_K0);
            *reinterpret_cast<unsigned*>(at) = w;
        }
    }
}

__device__ __forceinline__ void ub_arrive(unsigned bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }

__global__ void __launch_bounds__(UB_NT, UB_CTAS_PER_SM) k_blur_umma(const __grid_constant__ UmmaBlurMaps maps, const UmmaBlurTask* __restrict__ tasks,
                                                        int n_tasks, int n_imgs, int tiles_per_img,
                                                        const unsigned char* __restrict__ bands, int* __restrict__ err)
{
    extern __shared__ unsigned char ub_smem_raw[];
    const unsigned smem0 = ((unsigned)__cvta_generic_to_shared(ub_smem_raw) + 1023u) & ~1023u;  // the swizzles want 1024-byte alignment
    unsigned char* smem  = ub_smem_raw + (smem0 - (unsigned)__cvta_generic_to_shared(ub_smem_raw));
    // [0], [1]: image tile buffers (TMA); [2]: pass-1 MMAs done, [3]: pass-2 MMAs done (tcgen05.commit);
    // [4]: the A operands of pass 2 are written, [5]: the accumulators of pass 2 are read (UB_EPI arrivals each)
    unsigned long long* s_bar = reinterpret_cast<unsigned long long*>(smem + UB_OFF_BAR);
    unsigned& s_tmem          = *reinterpret_cast<unsigned*>(smem + UB_OFF_BAR + 48);
    const int tid = threadIdx.x, wid = tid >> 5, lane = tid & 31;
    const long long total = (long long)n_imgs * tiles_per_img;
    const long long g0 = total * blockIdx.x / gridDim.x, g1 = total * (blockIdx.x + 1) / gridDim.x;
    const int n_local = (int)(g1 - g0);
    if (n_local <= 0) return;
    const unsigned bar0  = (unsigned)__cvta_generic_to_shared(&s_bar[0]);
    const unsigned bar_p1 = bar0 + 16u, bar_p2 = bar0 + 24u, bar_e1 = bar0 + 32u, bar_e2 = bar0 + 40u;

    if (tid == 0)
    {
        for (int i = 0; i < 4; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8u * i));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar_e1), "r"(UB_EPI));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar_e2), "r"(UB_EPI));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (wid == UB_EPI / 32)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(&s_tmem)),
                     "n"(UB_TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    UmmaTileIter it;
    it.seek(tasks, n_imgs, g0);
    {   // the first strip's column band
        const uint4* src = reinterpret_cast<const uint4*>(bands + it.T.wc_off);
        uint4* dst       = reinterpret_cast<uint4*>(smem + UB_OFF_WC);
        for (int i = tid; i < UB_WC_BYTES / 16; i += UB_NT) dst[i] = __ldg(src + i);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the band was written by ordinary stores
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmem = s_tmem;

    if (wid == UB_EPI / 32)
    {
        // ================= the issuing warp: lane 0 feeds the TMA and the tensor core, everything it waits for is a barrier =========
        if (lane == 0)
        {
            int n_loaded = 0;  // `it` runs ahead of the MMAs: it is the next tile to load
            auto issue_load = [&]() {
                const int buf      = n_loaded & 1;
                const unsigned bar = bar0 + 8u * buf, dst = smem0 + (unsigned)(buf ? UB_OFF_IMG1 : UB_OFF_IMG0);
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((unsigned)UB_IMG_BYTES) : "memory");
                asm volatile(
                    "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
                    "l"(reinterpret_cast<unsigned long long>(&maps.src[it.T.level])), "r"(it.T.x0 - 16), "r"(it.row * UB_ROWS_OUT - 3),
                    "r"(it.img), "r"(bar)
                    : "memory");
                asm volatile(
                    "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
                        dst + (unsigned)(UB_ROWS_IN * UB_K0)),
                    "l"(reinterpret_cast<unsigned long long>(&maps.srcx[it.T.level])), "r"(it.T.x0 - 16 + UB_K0),
                    "r"(it.row * UB_ROWS_OUT - 3), "r"(it.img), "r"(bar)
                    : "memory");
                ++n_loaded;
                it.next(tasks, n_tasks, n_imgs);
            };
            const unsigned idesc1 = ub_idesc(UB_ROWS_IN), idesc2 = ub_idesc(UB_ROWS_OUT);
            const unsigned long long dwc = ub_desc(smem0 + UB_OFF_WC), dwcx = ub_desc32(smem0 + UB_OFF_WC + UB_M * UB_K0);
            const unsigned long long dwr = ub_desc(smem0 + UB_OFF_WR);
            issue_load();
            if (n_local > 1) issue_load();
            int failed = 0;
#pragma unroll 1
            for (int i = 0; i < n_local; ++i)
            {
                const int buf = i & 1;
                // pass 1: D1[m][n] = sum_k Wc[m][k] I[n][k], five K slices of 32 columns; its columns double as O_lo / O_hi of the
                // previous tile, which the epilogue must have read
                if (!ub_wait(bar0 + 8u * buf, (unsigned)((i >> 1) & 1))) { failed = 1; break; }
                if (i > 0 && !ub_wait(bar_e2, (unsigned)((i - 1) & 1))) { failed = 4; break; }
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const unsigned img_at = smem0 + (unsigned)(buf ? UB_OFF_IMG1 : UB_OFF_IMG0);
                const unsigned long long b = ub_desc(img_at);
#pragma unroll
                for (int k = 0; k < UB_K0 / 32; ++k) ub_mma_ss(tmem + UB_COL_D1, dwc + 2ull * k, b + 2ull * k, idesc1, k > 0);
                ub_mma_ss(tmem + UB_COL_D1, dwcx, ub_desc32(img_at + UB_ROWS_IN * UB_K0), idesc1, 1u);
                ub_commit(bar_p1);
                // pass 2 once the epilogue has written R_lo / R_hi (which also says pass 1 is complete: the tile buffer is free)
                if (!ub_wait(bar_e1, (unsigned)(i & 1))) { failed = 4; break; }
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                for (int k = 0; k < UB_ROWS_IN / 32; ++k) ub_mma_ts(tmem + UB_COL_OLO, tmem + UB_COL_ALO + 8 * k, dwr + 2ull * k, idesc2, k > 0);
#pragma unroll
                for (int k = 0; k < UB_ROWS_IN / 32; ++k) ub_mma_ts(tmem + UB_COL_OHI, tmem + UB_COL_AHI + 8 * k, dwr + 2ull * k, idesc2, k > 0);
                ub_commit(bar_p2);
                if (n_loaded < n_local) issue_load();  // after the MMAs: the epilogue is waiting for them
            }
            if (failed) atomicExch(err, 4 + failed);  // 5: a tile never arrived, 8: the epilogue never answered
        }
    }
    else
    {
        // ================= the epilogue warps =================
        const unsigned tlane = tmem + ((unsigned)(wid * 32) << 16);  // warp w reads lanes 32 w .. + 31
        int wr_level = -1, wr_row = -1;  // which row band lies in shared memory (interior tiles of a level share one)
        int failed   = 0;
#pragma unroll 1
        for (int i = 0; i < n_local; ++i)
        {
            const int y0 = it.row * UB_ROWS_OUT, H = it.T.h;
            const unsigned par = (unsigned)(i & 1);
            // the row band of this tile; away from the top and the bottom all tiles of a level share the band of the first
            // such tile. The previous tile's pass-2 MMAs, its only readers, have completed (bar_p2 was waited for)
            const bool interior = y0 >= 3 && y0 + UB_ROWS_OUT + 3 <= H;
            const bool have     = wr_level == it.T.level &&
                              (wr_row == it.row || (interior && wr_row * UB_ROWS_OUT >= 3 && wr_row * UB_ROWS_OUT + UB_ROWS_OUT + 3 <= H));
            if (!have)
            {
                const unsigned char* src = bands + it.T.wr_off + (size_t)it.row * UB_WR_BYTES;
                for (int j = tid; j < UB_WR_BYTES / 16; j += UB_EPI)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem0 + UB_OFF_WR + 16u * j), "l"(src + 16 * j) : "memory");
                wr_level = it.T.level, wr_row = it.row;
            }
            // this tile's output coordinates; then look at the next tile: a new strip needs its column band
            const int out_x = it.T.x0, out_img = it.img;
            const CUtensorMap* map_dst = &maps.dst[it.T.level];
            const int cur_task = it.task, cur_level = it.T.level;
            const bool cur_interior = it.T.wc_interior != 0;
            it.next(tasks, n_tasks, n_imgs);
            const bool new_strip = i + 1 < n_local && it.task != cur_task && !(cur_interior && it.T.wc_interior && it.T.level == cur_level);
            if (!ub_wait(bar_p1, par)) failed = 2;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (new_strip)  // pass 1 of this tile, the last reader of the old band, has completed
            {
                const unsigned char* src = bands + it.T.wc_off;
                for (int j = tid; j < UB_WC_BYTES / 16; j += UB_EPI)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem0 + UB_OFF_WC + 16u * j), "l"(src + 16 * j) : "memory");
            }
            // ---- split the 16-bit row sums into two u8 operands: image rows 4c .. 4c+3 -> bytes of TMEM column c. All of
            //      D1 goes into registers first: the operands are written over its columns ----
            {
                unsigned v[UB_ROWS_IN / 32][16];  // (row 2j, row 2j + 1) as 16-bit halves: the sums are below 2^16
#pragma unroll
                for (int c = 0; c < UB_ROWS_IN / 32; ++c) ub_ld16_pack(tlane + UB_COL_D1 + 32 * c, v[c]);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                // the last row is outside every band: its high byte feeds the rounding constant
                v[UB_ROWS_IN / 32 - 1][15] = (v[UB_ROWS_IN / 32 - 1][15] & 0xffffu) | 0x80000000u;
#pragma unroll
                for (int c = 0; c < UB_ROWS_IN / 32; ++c)
                {
                    unsigned lo[8], hi[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                    {
                        lo[j] = __byte_perm(v[c][2 * j], v[c][2 * j + 1], 0x6420);
                        hi[j] = __byte_perm(v[c][2 * j], v[c][2 * j + 1], 0x7531);
                    }
                    ub_st4(tlane + UB_COL_ALO + 8 * c, lo[0], lo[1], lo[2], lo[3]);
                    ub_st4(tlane + UB_COL_ALO + 8 * c + 4, lo[4], lo[5], lo[6], lo[7]);
                    ub_st4(tlane + UB_COL_AHI + 8 * c, hi[0], hi[1], hi[2], hi[3]);
                    ub_st4(tlane + UB_COL_AHI + 8 * c + 4, hi[4], hi[5], hi[6], hi[7]);
                }
            }
            if (!have || new_strip)
            {
                asm volatile("cp.async.wait_all;" ::: "memory");
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            }
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            // the previous tile's store must have read the output tile before this tile's bytes go there: thread 0 vouches for
            // it with its arrival — nobody gets past bar_p2 (pass 2 follows bar_e1) before
            if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            ub_arrive(bar_e1);
            if (!ub_wait(bar_p2, par)) failed = 3;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            // ---- the accumulators, 16 output rows at a time: combine (the rounding constant is in already) and lay the
            //      bytes out as image rows; this thread is column tid of the tile. After the last load the next tile's pass 1
            //      may take the columns over ----
            unsigned char* out = smem + UB_OFF_OUT + tid;
#pragma unroll
            for (int c = 0; c < UB_ROWS_OUT / 16; ++c)
            {
                unsigned ol[16], oh[16];
                ub_ld16(tlane + UB_COL_OLO + 16 * c, ol);
                ub_ld16(tlane + UB_COL_OHI + 16 * c, oh);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (c == UB_ROWS_OUT / 16 - 1)
                {
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    ub_arrive(bar_e2);
                }
#pragma unroll
                for (int j = 0; j < 16; ++j) out[(16 * c + j) * UB_OUT_COLS] = (unsigned char)((ol[j] + (oh[j] << 8)) >> 16);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            // one barrier per tile: the tile is complete; and nobody goes on alone after a barrier that never completed
            unsigned all_ok;
            asm volatile("{\n\t.reg .pred p, q;\n\tsetp.eq.u32 p, %1, 0;\n\tbarrier.cta.red.and.pred q, 1, %2, p;\n\tselp.u32 %0, 1, 0, q;\n\t}"
                         : "=r"(all_ok)
                         : "r"((unsigned)failed), "n"(UB_EPI)
                         : "memory");
            if (!all_ok)
            {
                if (failed && lane == 0) atomicExch(err, 4 + failed);  // 6 / 7: the MMAs of pass 1 / 2 never completed
                break;
            }
            if (tid == 0)
            {
                asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%1, %2, %3}], [%4];" ::"l"(
                                 reinterpret_cast<unsigned long long>(map_dst)),
                             "r"(out_x), "r"(y0), "r"(out_img), "r"(smem0 + UB_OFF_OUT)
                             : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        }
        if (tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (wid == UB_EPI / 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(UB_TMEM_COLS) : "memory");
}
}  // namespace snk
