// 7x7 Gaussian blur as two exact integer banded contractions on the tensor cores (included by orb_kernels.cuh).
//
// cv::GaussianBlur(level, 7x7, sigma 2, REFLECT_101) on 8-bit data is exact integer arithmetic: a row
// pass with the 8-bit kernel w = {18,34,48,56,48,34,18} (sum 256) giving 16-bit values, a column pass with the same
// kernel giving 24.16 fixed point, one rounding at the end: out = (sum_ij w_i w_j in[r+i-3][c+j-3] + 32768) >> 16.
// Both passes are multiplications with banded matrices, so they run as mma.sync.m16n8k32 (u8 x u8 -> s32):
//   row pass     D1[16 rows x 8 cols]  = In[16 rows x 32 cols] (A, row-major: 4 consecutive pixels per register, as they
//                                         lie in memory)  x  Bh[32 x 8] (B: the band, constants in registers)
//   column pass  Out[16 rows x 8 cols] = Bv[16 x 32 rows] (A: the band, constants)  x  D1[32 rows x 8 cols] (B), once for
//                                         the low and once for the high byte of the 16-bit D1: acc_lo + 256 acc_hi
// A K window of 32 holds the 7-tap band of 24 (row pass) / 16 (column pass) outputs: 32 multiply-adds per output and
// pass instead of 7 — the price of the tensor pipe, paid where it is idle; the ALU pipe, which bounds the dp4a kernel
// (k_blur: 28.7 thread-instructions per pixel), is left with the loads, the byte shuffles and the final rounding.
// Borders: REFLECT_101 folds the taps that fall outside back onto pixels inside, i.e. only the band matrices of the
// first / last strip and of the first / last row step differ (weights of folded taps add up); pixels outside are
// loaded as zero and carry zero weight.
// A warp owns a strip of 24 output columns and walks down a band of rows in steps of 16; the row-pass results of the 32
// rows a step needs live in a shared-memory ring (stored transposed, 16-bit), each step adds 16 new rows.
#pragma once

namespace snk
{
constexpr int BM_COLS  = 24;   // output columns per strip (three n-tiles of 8)
constexpr int BM_BAND  = 128;  // output rows per warp task
constexpr int BM_WARPS = 8;
constexpr int BM_CSTRIDE = 36;  // u16 entries per ring column (32 row slots + padding): 72 bytes, keeps the stores conflict-free

struct BlurTask
{
    short level, strip, band, _pad;
};

__device__ __forceinline__ void mma_u8(int (&d)[4], const unsigned (&a)[4], unsigned b0, unsigned b1)
{
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%12,%13};"
                 : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "r"(0), "r"(0), "r"(0), "r"(0));
}

// weight with which input position `in` (inside [0, len)) enters output position `out` under reflect-101
__device__ __forceinline__ unsigned band_weight(int in, int out, int len)
{
    if (in < 0 || in >= len) return 0u;
    unsigned s = 0;
#pragma unroll
    for (int d = -3; d <= 3; ++d)
    {
        const unsigned w = (d == 0) ? 56u : (d == 1 || d == -1) ? 48u : (d == 2 || d == -2) ? 34u : 18u;
        if (reflect101(out + d, len) == in) s += w;
    }
    return s;
}
// the same away from the borders: tap (in - out), no folding
__device__ __forceinline__ unsigned tap_weight(int d)
{
    const unsigned j = (unsigned)(d + 3);
    return j <= 6u ? (unsigned)((0x12223038302212ull >> (8 * j)) & 0xffu) : 0u;
}
// four consecutive input positions in0 .. in0+3 -> one packed register (byte i = weight of in0 + i).
// interior: no tap of this output is folded and every input it uses exists
__device__ __forceinline__ unsigned band_word(int in0, int out, int len, bool interior)
{
    if (interior)
        return tap_weight(in0 - out) | (tap_weight(in0 + 1 - out) << 8) | (tap_weight(in0 + 2 - out) << 16) | (tap_weight(in0 + 3 - out) << 24);
    return band_weight(in0, out, len) | (band_weight(in0 + 1, out, len) << 8) | (band_weight(in0 + 2, out, len) << 16) |
           (band_weight(in0 + 3, out, len) << 24);
}

__global__ void __launch_bounds__(BM_WARPS * 32) k_blur_mma(PyrDev P, const BlurTask* __restrict__ tasks, int n_tasks)
{
    __shared__ __align__(16) uint16_t ring_all[BM_WARPS][BM_COLS * BM_CSTRIDE];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int task = blockIdx.x * BM_WARPS + wid;
    if (task >= n_tasks) return;
    const int img    = blockIdx.y;
    const BlurTask T = tasks[task];
    const LevelDev& L = P.lv[T.level];
    const int W = L.w, H = L.h, pitch = L.pitch, gpitch = L.gauss_pitch;
    const uint8_t* src = L.img + (size_t)img * L.img_stride;
    uint8_t* dst       = L.gauss + (size_t)img * L.gauss_stride;
    uint16_t* ring     = ring_all[wid];
    const int g = lane >> 2, t = lane & 3;  // mma fragment coordinates: group id, thread in group

    const int c0   = T.strip * BM_COLS;   // first output column of the strip
    const int cwin = c0 - 4;              // first input column of the 32-wide window (4-byte aligned)
    const int rb   = T.band * BM_BAND, re = min(rb + BM_BAND, H);

    // ---- row-pass band: B fragments of the three n-tiles. b0: k = 4t..4t+3, b1: k = 16+4t..; n = g ----
    unsigned bh[3][2];
    const bool strip_interior = c0 >= 3 && c0 + BM_COLS + 3 <= W;  // warp-uniform
#pragma unroll
    for (int nt = 0; nt < 3; ++nt)
    {
        const int out = c0 + 8 * nt + g;
        bh[nt][0] = band_word(cwin + 4 * t, out, W, strip_interior);
        bh[nt][1] = band_word(cwin + 16 + 4 * t, out, W, strip_interior);
    }
    // column-pass band away from the top / bottom of the image: the same registers for every step
    unsigned av_in[4];
    av_in[0] = band_word(4 * t, 8 + g, 0, true);
    av_in[1] = band_word(4 * t, 16 + g, 0, true);
    av_in[2] = band_word(16 + 4 * t, 8 + g, 0, true);
    av_in[3] = band_word(16 + 4 * t, 16 + g, 0, true);

    // one row-pass tile: 16 rows starting at rt -> ring slots (rt - origin) & 31 ..; origin = rb - 8
    const int origin = rb - 8;
    auto row_pass = [&](int rt) {
        unsigned a[4];
        const int r_lo = rt + g, r_hi = rt + g + 8;
        const int cA = cwin + 4 * t, cB = cwin + 16 + 4 * t;
        const bool rl = r_lo >= 0 && r_lo < H, rh = r_hi >= 0 && r_hi < H;
        const bool ca = cA >= 0 && cA < pitch, cb = cB >= 0 && cB < pitch;
        a[0] = (rl && ca) ? __ldg(reinterpret_cast<const unsigned*>(src + (size_t)r_lo * pitch + cA)) : 0u;
        a[1] = (rh && ca) ? __ldg(reinterpret_cast<const unsigned*>(src + (size_t)r_hi * pitch + cA)) : 0u;
        a[2] = (rl && cb) ? __ldg(reinterpret_cast<const unsigned*>(src + (size_t)r_lo * pitch + cB)) : 0u;
        a[3] = (rh && cb) ? __ldg(reinterpret_cast<const unsigned*>(src + (size_t)r_hi * pitch + cB)) : 0u;
        const int s_lo = (rt - origin + g) & 31, s_hi = (rt - origin + g + 8) & 31;
#pragma unroll
        for (int nt = 0; nt < 3; ++nt)
        {
            int d[4];
            mma_u8(d, a, bh[nt][0], bh[nt][1]);
            // d0: (row g, col 2t), d1: (row g, col 2t+1), d2 / d3: row g+8 — stored transposed: ring[col][row slot]
            uint16_t* col = ring + (8 * nt + 2 * t) * BM_CSTRIDE;
            col[s_lo]              = (uint16_t)d[0];
            col[BM_CSTRIDE + s_lo] = (uint16_t)d[1];
            col[s_hi]              = (uint16_t)d[2];
            col[BM_CSTRIDE + s_hi] = (uint16_t)d[3];
        }
    };

    row_pass(rb - 8);  // rows [rb-8, rb+8): the upper half of the first window
    for (int r0 = rb; r0 < re; r0 += 16)
    {
        row_pass(r0 + 8);  // rows [r0+8, r0+24): the lower half of this step's window [r0-8, r0+24)
        __syncwarp();
        // ---- column-pass band: A fragment (rows of Out x rows of D1). a0: (out g, k 4t..), a1: (out g+8, k 4t..), a2 / a3: k+16
        const int rwin = r0 - 8;
        unsigned av[4] = {av_in[0], av_in[1], av_in[2], av_in[3]};
        if (r0 < 3 || r0 + 16 + 3 > H)  // warp-uniform: a tap of this step's outputs is folded
        {
            av[0] = band_word(rwin + 4 * t, r0 + g, H, false);
            av[1] = band_word(rwin + 4 * t, r0 + g + 8, H, false);
            av[2] = band_word(rwin + 16 + 4 * t, r0 + g, H, false);
            av[3] = band_word(rwin + 16 + 4 * t, r0 + g + 8, H, false);
        }
        const int s0 = (rwin - origin) & 31;  // ring slot of window row 0 (0 or 16)
#pragma unroll
        for (int nt = 0; nt < 3; ++nt)
        {
            // B fragment: column n = g of this n-tile, rows k = 4t..4t+3 (b0) and 16+4t.. (b1): 4 consecutive ring slots each
            const uint16_t* col = ring + (8 * nt + g) * BM_CSTRIDE;
            const uint2 x0 = *reinterpret_cast<const uint2*>(col + ((s0 + 4 * t) & 31));
            const uint2 x1 = *reinterpret_cast<const uint2*>(col + ((s0 + 16 + 4 * t) & 31));
            const unsigned lo0 = __byte_perm(x0.x, x0.y, 0x6420), hi0 = __byte_perm(x0.x, x0.y, 0x7531);
            const unsigned lo1 = __byte_perm(x1.x, x1.y, 0x6420), hi1 = __byte_perm(x1.x, x1.y, 0x7531);
            int dl[4], dh[4];
            mma_u8(dl, av, lo0, lo1);
            mma_u8(dh, av, hi0, hi1);
            const int col_out = c0 + 8 * nt + 2 * t;
            if (col_out < W)
            {
                const unsigned v0 = (unsigned)(dl[0] + (dh[0] << 8) + 32768) >> 16, v1 = (unsigned)(dl[1] + (dh[1] << 8) + 32768) >> 16;
                const unsigned v2 = (unsigned)(dl[2] + (dh[2] << 8) + 32768) >> 16, v3 = (unsigned)(dl[3] + (dh[3] << 8) + 32768) >> 16;
                const int ra = r0 + g, rbb = r0 + g + 8;
                if (col_out + 1 < W)
                {
                    if (ra < re) *reinterpret_cast<uint16_t*>(dst + (size_t)ra * gpitch + col_out) = (uint16_t)(v0 | (v1 << 8));
                    if (rbb < re) *reinterpret_cast<uint16_t*>(dst + (size_t)rbb * gpitch + col_out) = (uint16_t)(v2 | (v3 << 8));
                }
                else
                {
                    if (ra < re) dst[(size_t)ra * gpitch + col_out] = (uint8_t)v0;
                    if (rbb < re) dst[(size_t)rbb * gpitch + col_out] = (uint8_t)v2;
                }
            }
        }
        __syncwarp();  // the next step overwrites the half of the ring this step read first
    }
}
}  // namespace snk
