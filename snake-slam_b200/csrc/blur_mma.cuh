// 7x7 Gaussian blur as two exact banded contractions on the tensor cores (included by orb_kernels.cuh).
//
// cv::GaussianBlur(level, 7x7, sigma 2, REFLECT_101) on 8-bit data is exact integer arithmetic: a row
// pass with the 8-bit kernel w = {18,34,48,56,48,34,18} (sum 256) giving 16-bit values, a column pass with the same
// kernel giving 24.16 fixed point, one rounding at the end: out = (sum_ij w_i w_j in[r+i-3][c+j-3] + 32768) >> 16.
// Both passes are multiplications with banded matrices. They are chained in registers, transposed, so that the
// accumulator layout of the first is the operand layout of the second (no shared-memory round trip, no shuffles):
//   row pass     D1t[16 out cols x 8 rows] = Bh[16 out cols x 32 in cols] (A: the band, constants)
//                                            x Int[32 in cols x 8 rows]   (B: 4 consecutive pixels of a row per register,
//                                                                             exactly as they lie in memory)
//                mma.sync.m16n8k32 u8 x u8 -> s32. A thread ends up with D1t[col g / g+8][rows 2t, 2t+1]: two consecutive
//                ROWS of one column — which is what the A operand of m16n8k16 wants along K.
//   column pass  Outt[16 cols x 8 out rows] = D1t[16 cols x 16 rows] (A: two row-pass tiles, from registers)
//                                             x Bv[16 rows x 8 out rows] (B: the band, constants)
//                mma.sync.m16n8k16 f16 x f16 -> f32, once for the low and once for the high byte of the 16-bit D1t. Bytes
//                enter as the f16 numbers 1024 + b (bit pattern 0x6400 | b: no conversion instruction), weights are small
//                integers, every partial sum is an integer below 2^24: the f32 accumulation is exact. The accumulator
//                starts at 2^23 - 1024 * 256 (+ 32768 for the low byte), so the result's mantissa bits ARE the integer.
// Output rows [8j+4, 8j+12) use row tiles j and j+1 (rows 8j .. 8j+15): each row tile serves two steps without moving.
// A K window of 32 (16) holds the 7-tap band of 16 (8) outputs: 32 + 2 x 16 multiply-adds per pixel instead of 14 — the
// price of the tensor pipe, paid where it is idle; the ALU pipe, which bounds the dp4a kernel (k_blur: 28.7
// thread-instructions per pixel), is left with byte packing and the final shift.
// Borders: REFLECT_101 folds the taps that fall outside back onto pixels inside, i.e. only the band matrices of the
// first / last strip and of the first / last steps differ (weights of folded taps add up, their sum stays 256); pixels
// outside are loaded as zero and carry zero weight.
#pragma once
#include <cuda_fp16.h>

namespace snk
{
constexpr int BM_COLS  = 16;   // output columns per strip (M of both contractions)
constexpr int BM_BLOCKS = 16;  // 8-row output blocks per warp task (128 rows)
constexpr int BM_BAND  = 8 * BM_BLOCKS;
constexpr int BM_WARPS = 8;

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
__device__ __forceinline__ void mma_f16(float (&d)[4], unsigned a0, unsigned a1, unsigned a2, unsigned a3, unsigned b0, unsigned b1,
                                        float c)
{
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
                 : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "f"(c));
}

// weight with which input position `in` (inside [0, len)) enters output position `out` under reflect-101
__device__ __forceinline__ unsigned band_weight(int in, int out, int len)
{
    if (in < 0 || in >= len || out < 0 || out >= len) return 0u;
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
__device__ __forceinline__ unsigned weight_of(int in, int out, int len, bool interior)
{
    return interior ? tap_weight(in - out) : band_weight(in, out, len);
}
// four consecutive input positions in0 .. in0+3 as bytes of one register (u8 operand)
__device__ __forceinline__ unsigned band_word(int in0, int out, int len, bool interior)
{
    return weight_of(in0, out, len, interior) | (weight_of(in0 + 1, out, len, interior) << 8) |
           (weight_of(in0 + 2, out, len, interior) << 16) | (weight_of(in0 + 3, out, len, interior) << 24);
}
// two consecutive input positions as an f16 pair (small integers: exact)
__device__ __forceinline__ unsigned band_half2(int in0, int out, int len, bool interior)
{
    const __half2 h = __floats2half2_rn((float)weight_of(in0, out, len, interior), (float)weight_of(in0 + 1, out, len, interior));
    return *reinterpret_cast<const unsigned*>(&h);
}

__global__ void __launch_bounds__(BM_WARPS * 32) k_blur_mma(PyrDev P, const BlurTask* __restrict__ tasks, int n_tasks)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int task = blockIdx.x * BM_WARPS + wid;
    if (task >= n_tasks) return;
    const int img    = blockIdx.y;
    const BlurTask T = tasks[task];
    const LevelDev& L = P.lv[T.level];
    const int W = L.w, H = L.h, pitch = L.pitch, gpitch = L.gauss_pitch;
    const uint8_t* src = L.img + (size_t)img * L.img_stride;
    uint8_t* dst       = L.gauss + (size_t)img * L.gauss_stride;
    const int g = lane >> 2, t = lane & 3;  // mma fragment coordinates: group id, thread in group

    const int c0   = T.strip * BM_COLS;   // first output column of the strip
    const int cwin = c0 - 8;              // first input column of the 32-wide window (holds c0-3 .. c0+18; 8-byte aligned)
    // output blocks jb .. je-1; block j = rows [8j+4, 8j+12) from row tiles j and j+1; block -1 holds rows 0..3
    const int jb = T.band * BM_BLOCKS - 1, je = min(jb + BM_BLOCKS, (H + 3) / 8);

    // ---- row-pass band (A of the u8 mma): a0 (col g, k 4t..), a1 (col g+8, k 4t..), a2 / a3: k + 16 ----
    const bool strip_interior = c0 >= 3 && c0 + BM_COLS + 3 <= W;  // warp-uniform
    unsigned ah[4];
    ah[0] = band_word(cwin + 4 * t, c0 + g, W, strip_interior);
    ah[1] = band_word(cwin + 4 * t, c0 + g + 8, W, strip_interior);
    ah[2] = band_word(cwin + 16 + 4 * t, c0 + g, W, strip_interior);
    ah[3] = band_word(cwin + 16 + 4 * t, c0 + g + 8, W, strip_interior);
    // ---- column-pass band away from the top / bottom (B of the f16 mma): b0 (k 2t, 2t+1; out row g), b1: k + 8;
    //      window row k is image row 8j + k, output row g is image row 8j + 4 + g ----
    const unsigned bv_in0 = band_half2(2 * t, 4 + g, 0, true), bv_in1 = band_half2(2 * t + 8, 4 + g, 0, true);

    // one row-pass tile: rows 8j .. 8j+7 -> f16 operand pairs (low byte, high byte) for columns g and g+8
    const unsigned k64 = 0x64646464u;
    auto row_tile = [&](int j, unsigned& lo_a, unsigned& lo_b, unsigned& hi_a, unsigned& hi_b) {
        const int r = 8 * j + g, cA = cwin + 4 * t, cB = cwin + 16 + 4 * t;
        const bool rv = r >= 0 && r < H;
        const unsigned b0 = (rv && cA >= 0 && cA < pitch) ? __ldg(reinterpret_cast<const unsigned*>(src + (size_t)r * pitch + cA)) : 0u;
        const unsigned b1 = (rv && cB >= 0 && cB < pitch) ? __ldg(reinterpret_cast<const unsigned*>(src + (size_t)r * pitch + cB)) : 0u;
        int d[4];
        mma_u8(d, ah, b0, b1);
        // d0, d1: column g, rows 2t, 2t+1; d2, d3: column g+8. 16-bit values; bytes become f16 numbers 1024 + b
        const unsigned pa = __byte_perm((unsigned)d[0], (unsigned)d[1], 0x5410), pb = __byte_perm((unsigned)d[2], (unsigned)d[3], 0x5410);
        lo_a = __byte_perm(pa, k64, 0x4240), hi_a = __byte_perm(pa, k64, 0x4341);
        lo_b = __byte_perm(pb, k64, 0x4240), hi_b = __byte_perm(pb, k64, 0x4341);
    };

    unsigned la0, lb0, ha0, hb0, la1, lb1, ha1, hb1;  // tile j (upper half of the window) and tile j+1
    row_tile(jb, la0, lb0, ha0, hb0);
    for (int j = jb; j < je; ++j)
    {
        row_tile(j + 1, la1, lb1, ha1, hb1);
        unsigned bv0 = bv_in0, bv1 = bv_in1;
        const int rout = 8 * j + 4;  // first output row of the block
        if (rout < 3 || rout + 8 + 3 > H)  // warp-uniform: a tap of this block's outputs is folded (or the block leaves the image)
        {
            bv0 = band_half2(8 * j + 2 * t, rout + g, H, false);
            bv1 = band_half2(8 * j + 2 * t + 8, rout + g, H, false);
        }
        float fl[4], fh[4];
        mma_f16(fl, la0, lb0, la1, lb1, bv0, bv1, 8388608.0f - 262144.0f + 32768.0f);
        mma_f16(fh, ha0, hb0, ha1, hb1, bv0, bv1, 8388608.0f - 262144.0f);
        // d0: (col g, out row 2t), d1: (col g, out row 2t+1), d2 / d3: col g+8
#pragma unroll
        for (int i = 0; i < 4; ++i)
        {
            const unsigned v = (((unsigned)__float_as_int(fl[i]) & 0x7fffffu) + (((unsigned)__float_as_int(fh[i]) & 0x7fffffu) << 8)) >> 16;
            const int row = rout + 2 * t + (i & 1), col = c0 + g + 8 * (i >> 1);
            if (row >= 0 && row < H && col < W) dst[(size_t)row * gpitch + col] = (uint8_t)v;
        }
        la0 = la1, lb0 = lb1, ha0 = ha1, hb0 = hb1;
    }
}
}  // namespace snk
