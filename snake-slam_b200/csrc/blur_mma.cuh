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
// STATUS: an experiment that lost. Byte-exact on every shape and input of tests/test_gpu_extract.py, but 1.01 ms per 512
// images against 0.62 ms for the dp4a kernel k_blur (B200, profiles/r2m_*): the three tensor instructions per 16 x 8
// block are free, the 165 other instructions around them (staging, operand packing, un-biasing, byte stores, one
// barrier per 8 rows) are not; k_blur needs 115 per 128 pixels. Selected with SNK_BLUR_MMA=1; k_blur stays the default.
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

// tap weight: w[d + 3] for -3 <= d <= 3, else 0
__device__ __forceinline__ unsigned tap_weight(int d)
{
    const unsigned j = (unsigned)(d + 3);
    return j <= 6u ? (unsigned)((0x12223038302212ull >> (8 * j)) & 0xffu) : 0u;
}
// weight with which input position `in` enters output position `out` under reflect-101: tap position p = out + d lands
// on `in` directly (p == in), through the left mirror (p == -in, in > 0) or through the right one (p == 2 (len-1) - in,
// in < len-1); the three cases are disjoint for len > 6. interior: no folding possible, one lookup.
__device__ __forceinline__ unsigned weight_of(int in, int out, int len, bool interior)
{
    if (interior) return tap_weight(in - out);
    if (in < 0 || in >= len || out < 0 || out >= len) return 0u;
    unsigned w = tap_weight(in - out);
    if (in > 0) w += tap_weight(-in - out);
    if (in < len - 1) w += tap_weight(2 * (len - 1) - in - out);
    return w;
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

// One CTA = BM_WARPS adjacent strips (128 output columns) walking down a band of rows, one 8-row block per step. The
// CTA stages each 8-row input tile in shared memory with coalesced 16-byte cp.async copies (zero-filled outside the
// image) two tiles ahead, the warps read their operand words from it without bank conflicts, write their 16 x 8 output
// bytes into a shared tile, and the CTA stores that tile one step later as whole 128-byte rows.
constexpr int BM_TILE_W   = BM_WARPS * BM_COLS;  // 128 output columns per CTA
constexpr int BM_IN_CHUNKS = 10;                 // 16-byte chunks per staged row: columns ct0-16 .. ct0+143
constexpr int BM_IN_WORDS = 44;                  // row stride in words (176 bytes): operand reads hit 32 different banks
constexpr int BM_OUT_BYTES = 4 * (BM_TILE_W / 4 + 1);  // 132: the byte stores of a warp hit 8 different banks

__global__ void __launch_bounds__(BM_WARPS * 32) k_blur_mma(PyrDev P, const BlurTask* __restrict__ tasks, int n_tasks)
{
    __shared__ __align__(16) unsigned s_in[2][8][BM_IN_WORDS];
    __shared__ __align__(16) uint8_t s_out[2][8][BM_OUT_BYTES];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int img    = blockIdx.y;
    const BlurTask T = tasks[blockIdx.x];
    const LevelDev& L = P.lv[T.level];
    const int W = L.w, H = L.h, pitch = L.pitch, gpitch = L.gauss_pitch;
    const uint8_t* src = L.img + (size_t)img * L.img_stride;
    uint8_t* dst       = L.gauss + (size_t)img * L.gauss_stride;
    const int g = lane >> 2, t = lane & 3;  // mma fragment coordinates: group id, thread in group

    const int ct0  = T.strip * BM_TILE_W;      // first output column of the CTA tile
    const int c0   = ct0 + wid * BM_COLS;      // first output column of this warp's strip
    const int cwin = c0 - 8;                   // first input column of the warp's 32-wide window (holds c0-3 .. c0+18)
    // output blocks jb .. je-1; block j = rows [8j+4, 8j+12) from row tiles j and j+1; block -1 holds rows 0..3
    const int jb = T.band * BM_BLOCKS - 1, je = min(jb + BM_BLOCKS, (H + 3) / 8);

    // ---- staging: thread `tid` < 80 copies chunk tid % 10 of row tid / 10 of every tile (16 bytes, zero-filled outside) ----
    const int st_r = tid / BM_IN_CHUNKS, st_c = tid - st_r * BM_IN_CHUNKS;
    const int st_col = ct0 - 16 + 16 * st_c;
    const bool st_on = tid < 8 * BM_IN_CHUNKS, st_col_ok = st_col >= 0 && st_col < pitch;
    const unsigned st_dst = (unsigned)__cvta_generic_to_shared(&s_in[0][st_r][4 * st_c]);
    const uint8_t* st_src = src + st_col;
    auto stage = [&](int j) {
        if (st_on)
        {
            const int row  = 8 * j + st_r;
            const bool ok  = st_col_ok && row >= 0 && row < H;
            const uint8_t* ga = ok ? st_src + (size_t)row * pitch : src;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(st_dst + (unsigned)((j & 1) * (8 * BM_IN_WORDS * 4))), "l"(ga),
                         "r"(ok ? 16 : 0)
                         : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

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

    // one row-pass tile from shared memory -> f16 operand pairs (low byte, high byte) for columns g and g+8
    const unsigned k64 = 0x64646464u;
    const unsigned* in_w = &s_in[0][g][2 + 4 * wid + t];  // the warp's window starts 8 bytes into the staged row; bank 12g + 4 wid + t + 2
    auto row_tile = [&](int j, unsigned& lo_a, unsigned& lo_b, unsigned& hi_a, unsigned& hi_b) {
        const unsigned* rowp = in_w + (j & 1) * (8 * BM_IN_WORDS);
        int d[4];
        mma_u8(d, ah, rowp[0], rowp[4]);
        // d0, d1: column g, rows 2t, 2t+1; d2, d3: column g+8. 16-bit values; bytes become f16 numbers 1024 + b
        const unsigned pa = __byte_perm((unsigned)d[0], (unsigned)d[1], 0x5410), pb = __byte_perm((unsigned)d[2], (unsigned)d[3], 0x5410);
        lo_a = __byte_perm(pa, k64, 0x4240), hi_a = __byte_perm(pa, k64, 0x4341);
        lo_b = __byte_perm(pb, k64, 0x4240), hi_b = __byte_perm(pb, k64, 0x4341);
    };
    // the CTA's output tile of block j (rows 8j+4 .. 8j+11): whole 128-byte rows, one word per thread
    const int fl_col = ct0 + 4 * lane;
    const unsigned* fl_src = reinterpret_cast<const unsigned*>(&s_out[0][wid][4 * lane]);
    uint8_t* fl_dst = dst + fl_col;
    auto flush = [&](int j) {
        const int row = 8 * j + 4 + wid;
        if (row >= 0 && row < H && fl_col < W)
        {
            const unsigned v = fl_src[(j & 1) * (8 * BM_OUT_BYTES / 4)];
            uint8_t* o = fl_dst + (size_t)row * gpitch;
            if (fl_col + 4 <= W)
                *reinterpret_cast<unsigned*>(o) = v;
            else
                for (int b = 0; fl_col + b < W; ++b) o[b] = (uint8_t)(v >> (8 * b));
        }
    };
    uint8_t* so = &s_out[0][2 * t][wid * BM_COLS + g];  // this thread's four output bytes: rows 2t, 2t+1, columns g, g+8

    unsigned la0, lb0, ha0, hb0, la1, lb1, ha1, hb1;  // tile j (upper half of the window) and tile j+1
    stage(jb);
    stage(jb + 1);
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncthreads();
    row_tile(jb, la0, lb0, ha0, hb0);
    for (int j = jb; j < je; ++j)
    {
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();  // tile j+1 is in shared memory; every warp is done with tile j's buffer and with block j-1's output
        if (j > jb) flush(j - 1);
        stage(j + 2);     // into the buffer tile j occupied
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
        // d0: (col g, out row 2t), d1: (col g, out row 2t+1), d2 / d3: col g+8 -> bytes of the CTA's output tile
        uint8_t* sob = so + (j & 1) * (8 * BM_OUT_BYTES);
#pragma unroll
        for (int i = 0; i < 4; ++i)
        {
            const unsigned v = (((unsigned)__float_as_int(fl[i]) & 0x7fffffu) + (((unsigned)__float_as_int(fh[i]) & 0x7fffffu) << 8)) >> 16;
            sob[(i & 1) * BM_OUT_BYTES + 8 * (i >> 1)] = (uint8_t)v;
        }
        la0 = la1, lb0 = lb1, ha0 = ha1, hb0 = hb1;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    if (je > jb) flush(je - 1);
}
}  // namespace snk
