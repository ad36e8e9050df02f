// Colour conversion in front of the extractor.
//
// cv::cvtColor(src, dst, cv::COLOR_RGBA2GRAY) on 8-bit data as Input::computeGrayscaleImage calls it
// (Snake/Preprocess/Input.cpp:394-413). OpenCV's 8-bit path is fixed point with 15 fractional bits:
//   gray = (R*9798 + G*19235 + B*3735 + (1<<14)) >> 15          (alpha ignored)
// — checked exhaustively (all 2^24 colours) against cv2 4.13 by tests/golden/make_golden.py.
#pragma once
#include <cstdint>

namespace snk
{
__device__ __forceinline__ unsigned gray_of_rgba(unsigned px)
{
    // one dp4a: bytes (R,G,B,A) x (38,75,15,0) would be the 7-bit form; the 15-bit coefficients need 16-bit lanes
    const unsigned r = px & 255u, g = (px >> 8) & 255u, b = (px >> 16) & 255u;
    return (r * 9798u + g * 19235u + b * 3735u + (1u << 14)) >> 15;
}

// One thread converts four pixels: one 16-byte load, one 4-byte store. HBM-bound (5 bytes per pixel).
// src rows must be 16-byte aligned (pitch % 16 == 0), dst rows 4-byte aligned.
static __global__ void __launch_bounds__(128)
k_rgba_to_gray(const uint8_t* __restrict__ src, size_t src_stride, size_t src_pitch, uint8_t* __restrict__ dst,
               size_t dst_stride, int dst_pitch, int w, int h)
{
    const int q = blockIdx.x * 128 + threadIdx.x;  // group of 4 pixels
    const int y = blockIdx.y;
    if (4 * q >= w) return;
    const uint8_t* s = src + (size_t)blockIdx.z * src_stride + (size_t)y * src_pitch;
    uint8_t* d       = dst + (size_t)blockIdx.z * dst_stride + (size_t)y * dst_pitch;
    unsigned out;
    if (4 * q + 4 <= w)
    {
        const uint4 v = __ldcs(reinterpret_cast<const uint4*>(s) + q);  // streamed: never read again
        out = gray_of_rgba(v.x) | (gray_of_rgba(v.y) << 8) | (gray_of_rgba(v.z) << 16) | (gray_of_rgba(v.w) << 24);
    }
    else
    {
        out = 0;
        for (int j = 0; 4 * q + j < w; ++j)
            out |= gray_of_rgba(reinterpret_cast<const unsigned*>(s)[4 * q + j]) << (8 * j);
    }
    if (4 * q + 4 <= dst_pitch)
        reinterpret_cast<unsigned*>(d)[q] = out;
    else
        for (int j = 0; 4 * q + j < w; ++j) d[4 * q + j] = (uint8_t)(out >> (8 * j));
}
}  // namespace snk
