// TEST INFRASTRUCTURE ONLY — CPU oracle, never on the product path. See cvlite.h.
#include "cvlite.h"

#include <algorithm>
#include <cmath>
#include <cstring>

namespace cvlite
{
// ---------------------------------------------------------------------------------------------
// cv::resize INTER_LINEAR, 8UC1. Fixed point: 11-bit coefficients, horizontal pass to int32,
// vertical pass ((b0*(H0>>4))>>16) + ((b1*(H1>>4))>>16) + 2) >> 2.
// Horizontal taps at the right/left edge get their fractional part zeroed (clamp_coef = true);
// vertical taps keep the fraction and only clamp the row index (clamp_coef = false).
// ---------------------------------------------------------------------------------------------
void resize_tables(int ssize, int dsize, std::vector<int>& ofs, std::vector<int16_t>& coef, bool clamp_coef)
{
    ofs.resize(dsize);
    coef.resize(2 * (size_t)dsize);
    const double inv_scale = (double)dsize / ssize;
    const double scale     = 1.0 / inv_scale;
    for (int d = 0; d < dsize; ++d)
    {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s   = (int)std::floor(f);
        f -= s;
        if (clamp_coef)
        {
            if (s < 0)
            {
                f = 0;
                s = 0;
            }
            if (s >= ssize - 1)
            {
                f = 0;
                s = ssize - 1;
            }
        }
        ofs[d]          = s;
        coef[2 * d + 0] = (int16_t)std::lrintf((1.f - f) * 2048.f);
        coef[2 * d + 1] = (int16_t)std::lrintf(f * 2048.f);
    }
}

void resize_linear_u8(const uint8_t* src, int sh, int sw, size_t spitch, uint8_t* dst, int dh, int dw, size_t dpitch)
{
    std::vector<int> xofs, yofs;
    std::vector<int16_t> ialpha, ibeta;
    resize_tables(sw, dw, xofs, ialpha, true);
    resize_tables(sh, dh, yofs, ibeta, false);

    std::vector<int> rows[2];
    rows[0].resize(dw);
    rows[1].resize(dw);
    int cached[2] = {-1, -1};  // which source row each buffer holds

    auto hresize = [&](int sy, std::vector<int>& out) {
        const uint8_t* S = src + (size_t)sy * spitch;
        for (int x = 0; x < dw; ++x)
        {
            int sx  = xofs[x];
            int sx1 = std::min(sx + 1, sw - 1);
            out[x]  = S[sx] * ialpha[2 * x] + S[sx1] * ialpha[2 * x + 1];
        }
    };

    for (int y = 0; y < dh; ++y)
    {
        int sy0 = std::min(std::max(yofs[y], 0), sh - 1);
        int sy1 = std::min(std::max(yofs[y] + 1, 0), sh - 1);
        // tiny 2-row cache (rows advance monotonically)
        int slot0 = -1, slot1 = -1;
        for (int k = 0; k < 2; ++k)
        {
            if (cached[k] == sy0) slot0 = k;
        }
        if (slot0 < 0)
        {
            slot0 = (cached[0] == sy1) ? 1 : 0;
            hresize(sy0, rows[slot0]);
            cached[slot0] = sy0;
        }
        for (int k = 0; k < 2; ++k)
        {
            if (cached[k] == sy1) slot1 = k;
        }
        if (slot1 < 0)
        {
            slot1 = 1 - slot0;
            hresize(sy1, rows[slot1]);
            cached[slot1] = sy1;
        }
        const int* S0 = rows[slot0].data();
        const int* S1 = rows[slot1].data();
        const int b0 = ibeta[2 * y], b1 = ibeta[2 * y + 1];
        uint8_t* D = dst + (size_t)y * dpitch;
        for (int x = 0; x < dw; ++x)
        {
            int v = (((b0 * (S0[x] >> 4)) >> 16) + ((b1 * (S1[x] >> 4)) >> 16) + 2) >> 2;
            D[x]  = (uint8_t)std::min(std::max(v, 0), 255);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// cv::GaussianBlur(Size(7,7), 2, 2, BORDER_REFLECT_101), 8UC1: separable 8.8 fixed-point kernel
// {18,34,48,56,48,34,18}/256, 16.16 after the second pass, round-to-nearest.
// ---------------------------------------------------------------------------------------------
static inline int reflect101(int p, int len)
{
    if (len == 1) return 0;
    while (p < 0 || p >= len)
    {
        if (p < 0)
            p = -p;
        else
            p = 2 * (len - 1) - p;
    }
    return p;
}

void gaussian7_u8(const uint8_t* src, int h, int w, size_t spitch, uint8_t* dst, size_t dpitch)
{
    static const int K[7] = {18, 34, 48, 56, 48, 34, 18};
    std::vector<uint16_t> tmp((size_t)h * w);
    std::vector<int> xi(w + 6);
    for (int x = -3; x < w + 3; ++x) xi[x + 3] = reflect101(x, w);
    for (int y = 0; y < h; ++y)
    {
        const uint8_t* S = src + (size_t)y * spitch;
        uint16_t* T      = tmp.data() + (size_t)y * w;
        for (int x = 0; x < w; ++x)
        {
            int a = 0;
            for (int i = 0; i < 7; ++i) a += K[i] * S[xi[x + i]];
            T[x] = (uint16_t)a;
        }
    }
    for (int y = 0; y < h; ++y)
    {
        const uint16_t* R[7];
        for (int i = 0; i < 7; ++i) R[i] = tmp.data() + (size_t)reflect101(y + i - 3, h) * w;
        uint8_t* D = dst + (size_t)y * dpitch;
        for (int x = 0; x < w; ++x)
        {
            uint32_t a = 0;
            for (int i = 0; i < 7; ++i) a += (uint32_t)K[i] * R[i][x];
            D[x] = (uint8_t)((a + 32768u) >> 16);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// cv::FAST TYPE_9_16 with non-max suppression.
// ---------------------------------------------------------------------------------------------
static const int RING_DX[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int RING_DY[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

int fast_arc(const uint8_t* p, size_t pitch)
{
    int d[25];
    const int v = p[0];
    for (int k = 0; k < 16; ++k) d[k] = v - p[(ptrdiff_t)RING_DY[k] * (ptrdiff_t)pitch + RING_DX[k]];
    for (int k = 16; k < 25; ++k) d[k] = d[k - 16];
    int best = -256;
    for (int s = 0; s < 16; ++s)
    {
        int mn = d[s], mx = d[s];
        for (int j = 1; j < 9; ++j)
        {
            mn = std::min(mn, d[s + j]);
            mx = std::max(mx, d[s + j]);
        }
        best = std::max(best, std::max(mn, -mx));
    }
    return best;
}

void fast9_16_nms(const uint8_t* roi, int h, int w, size_t pitch, int threshold, std::vector<FastKp>& out)
{
    out.clear();
    if (h < 7 || w < 7) return;
    // score plane, zero where not a corner (also the 3-px frame), exactly what cv's rolling buffers hold
    std::vector<uint8_t> score((size_t)h * w, 0);
    for (int y = 3; y < h - 3; ++y)
    {
        const uint8_t* row = roi + (size_t)y * pitch;
        uint8_t* srow      = score.data() + (size_t)y * w;
        for (int x = 3; x < w - 3; ++x)
        {
            const uint8_t* p = row + x;
            const int v      = p[0];
            // any arc of 9 contains one pixel of every opposite pair: cheap rejection
            const int lo = v - threshold, hi = v + threshold;
            int a = p[(ptrdiff_t)3 * (ptrdiff_t)pitch], b = p[-(ptrdiff_t)3 * (ptrdiff_t)pitch];
            if (a >= lo && a <= hi && b >= lo && b <= hi) continue;
            a = p[3];
            b = p[-3];
            if (a >= lo && a <= hi && b >= lo && b <= hi) continue;
            const int arc = fast_arc(p, pitch);
            if (arc > threshold) srow[x] = (uint8_t)(arc - 1);
        }
    }
    for (int y = 3; y < h - 3; ++y)
    {
        const uint8_t* s0 = score.data() + (size_t)(y - 1) * w;
        const uint8_t* s1 = s0 + w;
        const uint8_t* s2 = s1 + w;
        for (int x = 3; x < w - 3; ++x)
        {
            const int s = s1[x];
            if (!s) continue;
            // a corner's score is arc-1 >= threshold; cv requires it to beat all 8 neighbours strictly
            if (s > s0[x - 1] && s > s0[x] && s > s0[x + 1] && s > s1[x - 1] && s > s1[x + 1] && s > s2[x - 1] &&
                s > s2[x] && s > s2[x + 1])
            {
                out.push_back({x, y, s});
            }
        }
    }
}
// cv::cvtColor RGBA -> GRAY for CV_8U (OpenCV imgproc color_rgb: RGB2Gray<uchar>, gray_shift = 15,
// RY15 = 9798, GY15 = 19235, BY15 = 3735, CV_DESCALE adds 1 << 14); identical to cv2 4.13 for all 2^24 colours
void rgba_to_gray_u8(const uint8_t* src, int h, int w, size_t spitch, uint8_t* dst, size_t dpitch)
{
    for (int y = 0; y < h; ++y)
    {
        const uint8_t* s = src + (size_t)y * spitch;
        uint8_t* d       = dst + (size_t)y * dpitch;
        for (int x = 0; x < w; ++x)
            d[x] = (uint8_t)((s[4 * x] * 9798 + s[4 * x + 1] * 19235 + s[4 * x + 2] * 3735 + (1 << 14)) >> 15);
    }
}
}  // namespace cvlite
