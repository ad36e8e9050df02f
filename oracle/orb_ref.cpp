// TEST INFRASTRUCTURE ONLY — CPU oracle, never on the product path. See orb_ref.h.
#include "orb_ref.h"

#include "cvlite.h"

#include <algorithm>
#include <cmath>
#include <cstring>

#define SNK_PATTERN_QUAL static const
#include "../include/snakeorb_pattern.h"

namespace orbref
{
// ---------------------------------------------------------------------------------------------
// ScalePyramid  (vision/util/ScalePyramid.cpp:16-55)
// ---------------------------------------------------------------------------------------------
ScalePyramid::ScalePyramid(int levels, double sf, int total_features)
    : num_levels(levels), total_num_features(total_features), scale_factor(sf)
{
    log_scale_factor = std::log(scale_factor);
    scale.assign(levels, 1.0);
    inv_scale.assign(levels, 1.0);
    squared_scale.assign(levels, 1.0);
    inv_squared_scale.assign(levels, 1.0);
    num_features.assign(levels, 0);

    const double factor = 1.0 / scale_factor;
    double per_level    = total_num_features * (1 - factor) / (1 - std::pow(factor, (double)num_levels));
    int sum             = 0;
    for (int i = 0; i < num_levels - 1; ++i)
    {
        num_features[i] = iRound((float)per_level);  // iRound takes a float (imath.h:73)
        sum += num_features[i];
        per_level *= factor;
    }
    for (int i = 1; i < num_levels; ++i)
    {
        scale[i]             = scale_factor * scale[i - 1];
        inv_scale[i]         = 1 / scale[i];
        squared_scale[i]     = scale[i] * scale[i];
        inv_squared_scale[i] = 1.f / squared_scale[i];
    }
    num_features[num_levels - 1] = std::max(total_num_features - sum, 0);
}

// ScalePyramid.h:26-41
void ScalePyramid::EstimateMinMaxDistance(double dist, int level, double& mn, double& mx) const
{
    double max_distance = dist * scale[level];
    double min_distance = max_distance / scale[num_levels - 1];
    min_distance /= scale_factor;
    max_distance *= scale_factor;
    mn = min_distance;
    mx = max_distance;
}

// ScalePyramid.h:44-57
double ScalePyramid::PredictScaleLevel(double ref_dist, int ref_level, double dist) const
{
    double max_distance    = ref_dist * scale[ref_level];
    double predicted_scale = max_distance / dist;
    return std::log(predicted_scale) / log_scale_factor;
}

// ScalePyramid.h:78
double ScalePyramid::ScaleForContiniousLevel(double level) const
{
    return std::pow(scale_factor, level);
}

// ---------------------------------------------------------------------------------------------
// Orientation / descriptor
// ---------------------------------------------------------------------------------------------
std::vector<int> AngleUmax()
{
    // OrbPattern.cpp:16-36: row half-widths of the radius-15 disc, made symmetric about the diagonal
    const int R = 15;
    std::vector<int> u(R + 1);
    int vmax         = iFloor(R * std::sqrt(2.f) / 2 + 1);
    int vmin         = iCeil(R * std::sqrt(2.f) / 2);
    const double hp2 = R * R;
    for (int v = 0; v <= vmax; ++v) u[v] = iRound((float)std::sqrt(hp2 - v * v));
    for (int v = R, v0 = 0; v >= vmin; --v)
    {
        while (u[v0] == u[v0 + 1]) ++v0;
        u[v] = v0;
        ++v0;
    }
    return u;
}

float ComputeAngle(const uint8_t* img, size_t pitch, float px, float py, const int* u_max)
{
    // OrbDescriptors.cpp:26-54. Integer first-order moments over the disc on the un-blurred level.
    int m_01 = 0, m_10 = 0;
    const uint8_t* c = img + (size_t)iRound(py) * pitch + iRound(px);
    for (int u = -15; u <= 15; ++u) m_10 += u * c[u];
    const int step = (int)pitch;
    for (int v = 1; v <= 15; ++v)
    {
        int v_sum = 0;
        int d     = u_max[v];
        for (int u = -d; u <= d; ++u)
        {
            int lo = c[u + v * step], hi = c[u - v * step];
            v_sum += (lo - hi);
            m_10 += u * (lo + hi);
        }
        m_01 += v * v_sum;
    }
    // Saiga::degrees<float>: a * 180.f / pi_f   (core/math/EigenGLMInterfaceFloat.h:69-74)
    float angle = atan2f((float)m_01, (float)m_10) * 180.0f / 3.14159265358979323846f;
    return (angle < 0) * 360 + angle;
}

Descriptor ComputeDescriptor(const uint8_t* img, size_t pitch, float px, float py, float angle_degrees)
{
    // OrbDescriptors.cpp:58-125. Saiga::radians<float>: a / 180.f * pi_f
    Descriptor out;
    uint8_t* bytes = (uint8_t*)out.data();
    float angle    = angle_degrees / 180.0f * 3.14159265358979323846f;
    float a = cosf(angle), b = sinf(angle);
    auto sample = [&](int k) -> int {
        // volatile-free but order-pinned: products rounded to float, then sub/add (no FMA: oracle is
        // built with -ffp-contract=off)
        float ux = (float)kOrbPattern[2 * k], uy = (float)kOrbPattern[2 * k + 1];
        float fx = px + (ux * a - uy * b);
        float fy = py + (ux * b + uy * a);
        return img[(size_t)iRound(fy) * pitch + iRound(fx)];
    };
    for (int i = 0; i < 32; ++i)
    {
        int val = 0;
        for (int j = 0; j < 8; ++j)
        {
            int t0 = sample(16 * i + 2 * j);
            int t1 = sample(16 * i + 2 * j + 1);
            val |= (t0 < t1) << j;
        }
        bytes[i] = (uint8_t)val;
    }
    return out;
}

// ---------------------------------------------------------------------------------------------
// Quadtree distribution  (FeatureDistribution.cpp:15-189)
// ---------------------------------------------------------------------------------------------
namespace
{
struct Node
{
    float cx, cy;  // corner
    float sx, sy;  // size
    int from, to;
    int count() const { return to - from; }
};

// FeatureDistribution.cpp:15-70: stable 4-way partition of the node's range by (x > mid) + 2*(y > mid)
void split(const Node& n, std::vector<KeyPointF>& kps, std::array<std::vector<KeyPointF>, 4>& bins, Node out[4])
{
    const float hx = n.sx * 0.5f, hy = n.sy * 0.5f;
    const float mx = n.cx + hx, my = n.cy + hy;
    for (auto& b : bins) b.clear();
    for (int i = n.from; i < n.to; ++i)
    {
        const KeyPointF& kp = kps[i];
        int q               = ((kp.y > my) << 1) + (kp.x > mx);
        bins[q].push_back(kp);
    }
    int pos = n.from;
    for (int q = 0; q < 4; ++q)
    {
        out[q].from = pos;
        for (auto& kp : bins[q]) kps[pos++] = kp;
        out[q].to = pos;
        out[q].sx = hx;
        out[q].sy = hy;
    }
    out[0].cx = n.cx, out[0].cy = n.cy;
    out[1].cx = mx, out[1].cy = n.cy;
    out[2].cx = n.cx, out[2].cy = my;
    out[3].cx = mx, out[3].cy = my;
}
}  // namespace

std::vector<KeyPointF> Distribute(std::vector<KeyPointF>& kps, float min_x, float min_y, float max_x, float max_y,
                                  int target_n)
{
    if ((int)kps.size() <= target_n) return kps;  // :83-86

    std::vector<Node> leaves, inner, next;
    std::array<std::vector<KeyPointF>, 4> bins;

    // root square (:96-103); the caller passes bounds that are offset w.r.t. the keypoint frame — kept as is
    const float ccx = (min_x + max_x) * 0.5f, ccy = (min_y + max_y) * 0.5f;
    const float wx = max_x - min_x, wy = max_y - min_y;
    const float side = std::max(wx, wy);
    next.push_back({ccx - side * 0.5f, ccy - side * 0.5f, side, side, 0, (int)kps.size()});

    int last_node_count      = 0;
    int last_processed_inner = -1;
    bool finished            = false;
    while (!finished && !next.empty())
    {
        inner.swap(next);
        next.clear();
        int node_count = (int)(leaves.size() + inner.size());
        if (node_count == last_node_count) break;
        last_node_count = node_count;

        // :124 — std::sort is not stable; libstdc++'s introsort permutation of ties is part of the result
        std::sort(inner.begin(), inner.end(), [](const Node& a, const Node& b) { return a.count() > b.count(); });

        for (int i = 0; i < (int)inner.size(); ++i)
        {
            Node ch[4];
            split(inner[i], kps, bins, ch);
            for (int q = 0; q < 4; ++q)
            {
                int c = ch[q].count();
                if (c == 0) continue;
                if (c == 1)
                    leaves.push_back(ch[q]);
                else
                    next.push_back(ch[q]);
            }
            int remaining = (int)inner.size() - i - 1;
            if ((int)leaves.size() + (int)next.size() + remaining >= target_n)
            {
                last_processed_inner = i;  // :153 — the node just split is emitted again below
                finished             = true;
                break;
            }
        }
    }

    std::vector<KeyPointF> result;
    result.reserve(leaves.size() + inner.size());
    auto best = [&](const Node& n) {
        int b = n.from;  // std::max_element keeps the first maximum
        for (int i = n.from + 1; i < n.to; ++i)
            if (kps[b].response < kps[i].response) b = i;
        result.push_back(kps[b]);
    };
    for (auto& n : leaves) best(n);
    for (auto& n : next) best(n);
    for (size_t i = (size_t)last_processed_inner; i < inner.size(); ++i) best(inner[i]);  // :183 size_t(-1) => none
    return result;
}

// ---------------------------------------------------------------------------------------------
// Extractor driver  (ORBExtractor.cpp:82-289)
// ---------------------------------------------------------------------------------------------
CellGrid MakeCellGrid(int rows, int cols)
{
    // ORBExtractor.cpp:99-112
    CellGrid g;
    g.minB             = 19 - 3;
    g.maxBX            = cols - 19 + 3;
    g.maxBY            = rows - 19 + 3;
    const float width  = (float)(g.maxBX - g.minB);
    const float height = (float)(g.maxBY - g.minB);
    g.nCols            = (int)(width / 30.f);
    g.nRows            = (int)(height / 30.f);
    g.wCell            = (int)std::ceil(width / g.nCols);
    g.hCell            = (int)std::ceil(height / g.nRows);
    return g;
}

OrbExtractor::OrbExtractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int threads)
    : pyramid(nlevels, scaleFactor, nfeatures),
      num_levels(nlevels),
      th_fast(iniThFAST),
      th_fast_min(minThFAST),
      num_threads(threads)
{
    levels.resize(nlevels);
    u_max = AngleUmax();
}

void OrbExtractor::AllocatePyramid(int rows, int cols)
{
    // ORBExtractor.cpp:236-259
    if (!levels.front().image.empty() && levels.front().rows == rows && levels.front().cols == cols) return;
    for (int l = 0; l < num_levels; ++l)
    {
        float s        = (float)pyramid.inv_scale[l];
        levels[l].rows = iRound(rows * s);
        levels[l].cols = iRound(cols * s);
        levels[l].image.assign((size_t)levels[l].rows * levels[l].cols, 0);
        levels[l].gauss.assign((size_t)levels[l].rows * levels[l].cols, 0);
    }
}

void OrbExtractor::ComputePyramid(const uint8_t* img, int rows, int cols, size_t pitch)
{
    // ORBExtractor.cpp:261-287. Level 0 = copy; level l = resize(level l-1). The reflect-101 aprons the
    // reference writes are never read by later stages, so they are not materialised.
    AllocatePyramid(rows, cols);
    for (int y = 0; y < rows; ++y) std::memcpy(levels[0].image.data() + (size_t)y * cols, img + y * pitch, cols);
    for (int l = 1; l < num_levels; ++l)
    {
        auto& p = levels[l - 1];
        auto& c = levels[l];
        cvlite::resize_linear_u8(p.image.data(), p.rows, p.cols, p.cols, c.image.data(), c.rows, c.cols, c.cols);
    }
}

void OrbExtractor::DetectKeypoints()
{
    // ORBExtractor.cpp:90-175
#pragma omp parallel for num_threads(num_threads) schedule(dynamic)
    for (int level = 0; level < num_levels; ++level)
    {
        auto& L = levels[level];
        L.raw.clear();
        const CellGrid g = MakeCellGrid(L.rows, L.cols);
        std::vector<cvlite::FastKp> cell;
        for (int i = 0; i < g.nRows; ++i)
        {
            const int iniY = g.minB + i * g.hCell;
            int maxY       = iniY + g.hCell + 6;
            if (iniY >= g.maxBY - 3) continue;
            if (maxY > g.maxBY) maxY = g.maxBY;
            for (int j = 0; j < g.nCols; ++j)
            {
                const int iniX = g.minB + j * g.wCell;
                int maxX       = iniX + g.wCell + 6;
                if (iniX >= g.maxBX - 6) continue;
                if (maxX > g.maxBX) maxX = g.maxBX;
                const uint8_t* roi = L.image.data() + (size_t)iniY * L.cols + iniX;
                cvlite::fast9_16_nms(roi, maxY - iniY, maxX - iniX, L.cols, th_fast, cell);
                if (cell.empty()) cvlite::fast9_16_nms(roi, maxY - iniY, maxX - iniX, L.cols, th_fast_min, cell);
                for (auto& c : cell)
                {
                    // cv::KeyPoint defaults: size 7, angle -1, octave 0
                    KeyPointF kp{(float)c.x, (float)c.y, 7.f, -1.f, (float)c.score, 0};
                    kp.x += j * g.wCell;
                    kp.y += i * g.hCell;
                    L.raw.push_back(kp);
                }
            }
        }
        std::vector<KeyPointF> work = L.raw;  // Distribute permutes its input; keep the raw list for tests
        L.kept = Distribute(work, (float)g.minB, (float)g.minB, (float)g.maxBX, (float)g.maxBY,
                            pyramid.num_features[level]);
        const int scaledPatchSize = (int)(31 * pyramid.scale[level]);
        for (auto& kp : L.kept)
        {
            kp.x += g.minB;
            kp.y += g.minB;
            kp.octave = level;
            kp.size   = (float)scaledPatchSize;
            kp.angle  = ComputeAngle(L.image.data(), L.cols, kp.x, kp.y, u_max.data());
        }
    }
}

void OrbExtractor::Detect(const uint8_t* img, int rows, int cols, size_t pitch, std::vector<KeyPointF>& kps,
                          std::vector<Descriptor>& desc)
{
    // ORBExtractor.cpp:178-233
    if (!img || rows <= 0 || cols <= 0) return;
    desc.clear();
    ComputePyramid(img, rows, cols, pitch);
    DetectKeypoints();
    int n = 0;
    for (auto& L : levels)
    {
        L.offset = n;
        n += (int)L.kept.size();
    }
    desc.resize(n);
    kps.resize(n);
#pragma omp parallel for num_threads(num_threads) schedule(dynamic)
    for (int level = 0; level < num_levels; ++level)
    {
        auto& L = levels[level];
        if (L.kept.empty()) continue;
        cvlite::gaussian7_u8(L.image.data(), L.rows, L.cols, L.cols, L.gauss.data(), L.cols);
        for (size_t i = 0; i < L.kept.size(); ++i)
            desc[L.offset + i] = ComputeDescriptor(L.gauss.data(), L.cols, L.kept[i].x, L.kept[i].y, L.kept[i].angle);
        std::vector<KeyPointF> out = L.kept;
        if (level != 0)
        {
            float s = (float)pyramid.scale[level];
            for (auto& kp : out)
            {
                kp.x *= s;
                kp.y *= s;
            }
        }
        for (size_t i = 0; i < out.size(); ++i) kps[L.offset + i] = out[i];
    }
}
}  // namespace orbref
