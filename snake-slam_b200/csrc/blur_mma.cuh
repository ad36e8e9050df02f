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
// STATUS: the default blur since round 2. Byte-exact on every shape and input of tests/test_gpu_extract.py; 0.586 ms per
// 512 EuRoC images against 0.62 ms for the dp4a kernel k_blur (SNK_BLUR_MMA=0), 364 M against 513 M warp instructions.
// Earlier variants on the way (git history): strips with a shared-memory ring between the passes 0.99 ms,
// register chaining with direct byte stores 1.05 ms, CTA-staged tiles with a barrier per 8 rows 1.01 ms, this one at 92
// registers 0.89 ms (latency-bound at 25 % occupancy); what is left around the three tensor instructions per 16 x 8 block
// is ~80 instructions of operand packing, un-biasing, transposing and addressing. Tile shape sweep afterwards
// (tools/scratch/blur_variants.sh; strips per warp x CTAs per SM x one-tile-ahead loads): 2/4/ahead 0.581, 3/4 0.617,
// 3/4/ahead 0.629, 2/5/ahead 0.717, 3/5 0.806, 2/6/ahead 0.815 ms — nothing beyond noise over 4/4 (0.586); more CTAs per
// SM spill.
// Borders: REFLECT_101 folds the taps that fall outside back onto pixels inside, i.e. only the band matrices of the
// first / last strip and of the first / last steps differ (weights of folded taps add up, their sum stays 256); pixels
// outside are loaded as zero and carry zero weight.
#pragma once
#include <cuda_fp16.h>

#include "blur_weights.cuh"

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

// two consecutive input positions as an f16 pair (small integers: exact)
__device__ __forceinline__ unsigned band_half2(int in0, int out, int len, bool interior)
{
    const __half2 h = __floats2half2_rn((float)weight_of(in0, out, len, interior), (float)weight_of(in0 + 1, out, len, interior));
    return *reinterpret_cast<const unsigned*>(&h);
}

// A warp works alone on BM_STRIPS adjacent strips (64 output columns) and walks down a band of rows, one 8-row block
// per step: no shared-memory staging, no barrier. Loads: lane (g, t) reads words t, t+4, .. of row g of the tile — every
// word is the second operand register of one strip and the first of the next. Stores: a 4 x 4 byte transpose over four
// lanes (two shuffles) turns the accumulator layout into whole words of four consecutive columns.
constexpr int BM_STRIPS = 4;
constexpr int BM_TILE_W = BM_STRIPS * BM_COLS;  // 64 output columns per warp

__global__ void __launch_bounds__(BM_WARPS * 32, 4) k_blur_mma(PyrDev P, const BlurTask* __restrict__ tasks, int n_tasks)
{
    __shared__ uint4 s_ah[BM_WARPS][BM_STRIPS][32];  // row-pass band registers per strip (they differ at the image's sides)
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

    const int ct0 = T.strip * BM_TILE_W;  // first output column of the warp's tile; strip s covers [ct0 + 16 s, +16)
    // output blocks jb .. je-1; block j = rows [8j+4, 8j+12) from row tiles j and j+1; block -1 holds rows 0..3
    const int jb = T.band * BM_BLOCKS - 1, je = min(jb + BM_BLOCKS, (H + 3) / 8);

    // ---- row-pass band (A of the u8 mma) per strip: a0 (col g, k 4t..), a1 (col g+8, k 4t..), a2 / a3: k + 16;
    //      the strip's 32-wide input window starts at c0 - 8 (holds c0-3 .. c0+18) ----
    {
        // away from the image's sides every strip has the same band: in - out = (4t + i - 8) - g
        const uint4 a_in = make_uint4(band_word(4 * t - 8, g, 0, true), band_word(4 * t - 8, g + 8, 0, true),
                                      band_word(4 * t + 8, g, 0, true), band_word(4 * t + 8, g + 8, 0, true));
#pragma unroll
        for (int s = 0; s < BM_STRIPS; ++s)
        {
            const int c0 = ct0 + s * BM_COLS, cwin = c0 - 8;
            uint4 a = a_in;
            if (!(c0 >= 3 && c0 + BM_COLS + 3 <= W))  // warp-uniform: a tap of this strip is folded at the left / right side
                a = make_uint4(band_word(cwin + 4 * t, c0 + g, W, false), band_word(cwin + 4 * t, c0 + g + 8, W, false),
                               band_word(cwin + 16 + 4 * t, c0 + g, W, false), band_word(cwin + 16 + 4 * t, c0 + g + 8, W, false));
            s_ah[wid][s][lane] = a;
        }
    }
    __syncwarp();
    // ---- column-pass band away from the top / bottom (B of the f16 mma): b0 (k 2t, 2t+1; out row g), b1: k + 8;
    //      window row k is image row 8j + k, output row g is image row 8j + 4 + g ----
    const unsigned bv_in0 = band_half2(2 * t, 4 + g, 0, true), bv_in1 = band_half2(2 * t + 8, 4 + g, 0, true);

    // loads: word k of the lane = columns ct0 - 8 + 4t + 16k .. +3 of row g of the tile
    const int lcol = ct0 - 8 + 4 * t;
    unsigned lmask = 0;  // words of the lane that lie inside the row
#pragma unroll
    for (int k = 0; k <= BM_STRIPS; ++k)
        if (lcol + 16 * k >= 0 && lcol + 16 * k < pitch) lmask |= 1u << k;
    const uint8_t* lsrc = src + lcol;
    // stores after the transpose: this lane ends up with row 2t + (gi & 1), columns 4 q + 8 (gi >> 1) .. +3 of each strip
    const int gi = g & 3, q = g >> 2;
    const unsigned selA = (gi & 1) ? 0x3715u : 0x6240u, selB = (gi & 2) ? 0x3276u : 0x5410u;
    const int srow = 2 * t + (gi & 1), scol = ct0 + 4 * q + 8 * (gi >> 1);
    uint8_t* sdst = dst + scol;
    unsigned sfull = 0, spart = 0;  // per strip: the lane's word lies fully / partly inside the image
#pragma unroll
    for (int s = 0; s < BM_STRIPS; ++s)
    {
        const int col = scol + s * BM_COLS;
        if (col + 4 <= W) sfull |= 1u << s;
        else if (col < W) spart |= 1u << s;
    }
    const unsigned k64 = 0x64646464u;

    unsigned st[BM_STRIPS][4];  // per strip: f16 operand pairs of the upper row tile (lo / hi byte; columns g / g+8)
    auto row_ops = [&](const unsigned (&w)[BM_STRIPS + 1], int s, unsigned (&o)[4]) {
        const uint4 a4 = s_ah[wid][s][lane];
        const unsigned a[4] = {a4.x, a4.y, a4.z, a4.w};
        int d[4];
        mma_u8(d, a, w[s], w[s + 1]);
        // d0, d1: column g, rows 2t, 2t+1; d2, d3: column g+8. 16-bit values; bytes become f16 numbers 1024 + b
        const unsigned pa = __byte_perm((unsigned)d[0], (unsigned)d[1], 0x5410), pb = __byte_perm((unsigned)d[2], (unsigned)d[3], 0x5410);
        o[0] = __byte_perm(pa, k64, 0x4240), o[1] = __byte_perm(pb, k64, 0x4240);  // low bytes: columns g, g+8
        o[2] = __byte_perm(pa, k64, 0x4341), o[3] = __byte_perm(pb, k64, 0x4341);  // high bytes
    };
    auto load_tile = [&](int j, unsigned (&w)[BM_STRIPS + 1]) {
        const int r = 8 * j + g;
        const unsigned m = (r >= 0 && r < H) ? lmask : 0u;
        const uint8_t* p = lsrc + (size_t)(m ? r : 0) * pitch;
#pragma unroll
        for (int k = 0; k <= BM_STRIPS; ++k) w[k] = (m >> k & 1u) ? __ldg(reinterpret_cast<const unsigned*>(p + 16 * k)) : 0u;
    };

    {
        unsigned w[BM_STRIPS + 1];
        load_tile(jb, w);
#pragma unroll
        for (int s = 0; s < BM_STRIPS; ++s) row_ops(w, s, st[s]);
    }
    for (int j = jb; j < je; ++j)
    {
        unsigned w[BM_STRIPS + 1];
        load_tile(j + 1, w);
        unsigned bv0 = bv_in0, bv1 = bv_in1;
        const int rout = 8 * j + 4;  // first output row of the block
        if (rout < 3 || rout + 8 + 3 > H)  // warp-uniform: a tap of this block's outputs is folded (or the block leaves the image)
        {
            bv0 = band_half2(8 * j + 2 * t, rout + g, H, false);
            bv1 = band_half2(8 * j + 2 * t + 8, rout + g, H, false);
        }
        const int row = rout + srow;
        const bool row_ok = row >= 0 && row < H;
        uint8_t* orow = sdst + (size_t)(row_ok ? row : 0) * gpitch;
#pragma unroll
        for (int s = 0; s < BM_STRIPS; ++s)
        {
            unsigned nx[4];
            row_ops(w, s, nx);
            float fl[4], fh[4];
            mma_f16(fl, st[s][0], st[s][1], nx[0], nx[1], bv0, bv1, 8388608.0f - 262144.0f + 32768.0f);
            mma_f16(fh, st[s][2], st[s][3], nx[2], nx[3], bv0, bv1, 8388608.0f - 262144.0f);
            // result = (lo + 256 hi + 32768) >> 16 = byte 2 of lo_bits + (hi_bits << 8) (the 0x4B exponent bytes cancel out of
            // byte 2: they sit in byte 3 of lo_bits and leave the word from hi_bits << 8)
            unsigned sum[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) sum[i] = (unsigned)__float_as_int(fl[i]) + ((unsigned)__float_as_int(fh[i]) << 8);
            const unsigned x = __byte_perm(__byte_perm(sum[0], sum[1], 0x0062), __byte_perm(sum[2], sum[3], 0x0062), 0x5410);
            // x = [ (row 2t, col g), (row 2t+1, col g), (row 2t, col g+8), (row 2t+1, col g+8) ] -> 4 x 4 transpose over lanes gi = 0..3
            const unsigned y = __byte_perm(x, __shfl_xor_sync(0xffffffffu, x, 4), selA);
            const unsigned z = __byte_perm(y, __shfl_xor_sync(0xffffffffu, y, 8), selB);
            if (row_ok)
            {
                uint8_t* o = orow + s * BM_COLS;
                if (sfull >> s & 1u)
                    *reinterpret_cast<unsigned*>(o) = z;
                else if (spart >> s & 1u)
                    for (int b = 0; scol + s * BM_COLS + b < W; ++b) o[b] = (uint8_t)(z >> (8 * b));
            }
            st[s][0] = nx[0], st[s][1] = nx[1], st[s][2] = nx[2], st[s][3] = nx[3];
        }
    }
}
}  // namespace snk
