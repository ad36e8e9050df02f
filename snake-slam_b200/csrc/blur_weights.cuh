// Band weights of the 7-tap Gaussian row / column pass with cv's BORDER_REFLECT_101 folded in (shared by the tensor-core
// blur kernels, blur_mma.cuh and blur_umma.cuh).
#pragma once

namespace snk
{
// tap weight: w[d + 3] for -3 <= d <= 3, else 0
__device__ __forceinline__ unsigned tap_weight(int d)
{
    const unsigned j = (unsigned)(d < 0 ? -d : d);  // the kernel is symmetric: {56, 48, 34, 18} by |d|
    return j <= 3u ? ((0x12223038u >> (8 * j)) & 0xffu) : 0u;
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
}  // namespace snk
