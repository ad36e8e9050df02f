// TEST INFRASTRUCTURE ONLY — CPU oracle, never on the product path.
//
// cvlite: plain-integer restatements of the three OpenCV primitives the reference's
// CPU ORB extractor calls (OpenCV itself is an un-vendored dependency of the reference,
// version unpinned; pinned here to opencv-python-headless 4.13.0 which is what
// tests/golden was generated with — see tests/golden/make_golden.py).
//
//   resize_linear_u8   <- cv::resize(..., INTER_LINEAR)      ORBExtractor.cpp:282
//   gaussian7_u8       <- cv::GaussianBlur(7x7, s=2, R101)   ORBExtractor.cpp:213
//   fast9_16_nms       <- cv::FAST(roi, th, true)            ORBExtractor.cpp:133,138
//
// (paths relative to External/saiga/src/saiga/vision/features/ in the reference)
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace cvlite
{
struct FastKp
{
    int x, y;   // ROI-relative pixel
    int score;  // cv::KeyPoint::response
};

// cv::cvtColor(src, dst, COLOR_RGBA2GRAY), 8-bit (Input.cpp:407): 15-bit fixed point, alpha ignored
void rgba_to_gray_u8(const uint8_t* src, int h, int w, size_t spitch, uint8_t* dst, size_t dpitch);
void resize_linear_u8(const uint8_t* src, int sh, int sw, size_t spitch, uint8_t* dst, int dh, int dw, size_t dpitch);

// coefficient tables exactly as cv::resize builds them (also used by tests to cross-check the device tables)
void resize_tables(int ssize, int dsize, std::vector<int>& ofs, std::vector<int16_t>& coef /* 2 per dst */, bool clamp_coef);

void gaussian7_u8(const uint8_t* src, int h, int w, size_t spitch, uint8_t* dst, size_t dpitch);

// FAST-9/16 corner score of one pixel: max over the 16 arcs of 9 contiguous ring pixels of
// min(center - ring) and min(ring - center); corner at threshold t  <=>  arc > t; cv response = arc - 1.
int fast_arc(const uint8_t* p, size_t pitch);

void fast9_16_nms(const uint8_t* roi, int h, int w, size_t pitch, int threshold, std::vector<FastKp>& out);
}  // namespace cvlite
