// TEST INFRASTRUCTURE ONLY — CPU oracle, never on the product path.
//
// Restatement of the reference's CPU ORB extractor (Saiga::ORBExtractor and the classes it
// uses) on plain structs, for parity tests and for the timed CPU baseline of bench.py.
// Every function cites the reference lines it follows; paths are relative to
// /root/reference/External/saiga/src/saiga/ .
//
// Pinning: the OpenCV primitives (cvlite.*) are pinned against cv2 4.13.0 golden vectors
// (tests/golden/); ScalePyramid / Distribute / ComputeAngle / ComputeDescriptor are pinned
// against the reference's own translation units compiled unmodified into oracle/_ref
// (tests/test_oracle_vs_ref.py) and against golden vectors produced by them.
#pragma once
#include <array>
#include <cstddef>
#include <cstdint>
#include <vector>

namespace orbref
{
// == Saiga::KeyPoint<float> (vision/features/Features.h:21-67), 24 bytes
struct KeyPointF
{
    float x, y;
    float size;
    float angle;
    float response;
    int octave;
};
static_assert(sizeof(KeyPointF) == 24, "layout");

using Descriptor = std::array<uint64_t, 4>;  // == Saiga::DescriptorORB (Features.h:70)

// core/math/imath.h:63-76
inline int iFloor(double v) { return ((int)v) - (((int)v) > v); }
inline int iCeil(float v) { return ((int)v) + (((int)v) < v); }
inline int iRound(float v) { return (int)(v + (v >= 0 ? 0.5f : -0.5f)); }

// vision/util/ScalePyramid.cpp:16-55, ScalePyramid.h:26-84
struct ScalePyramid
{
    int num_levels         = 1;
    int total_num_features = 0;
    double scale_factor    = 1;
    double log_scale_factor = 0;
    std::vector<double> scale, inv_scale, squared_scale, inv_squared_scale;
    std::vector<int> num_features;

    ScalePyramid() = default;
    ScalePyramid(int levels, double scale_factor, int total_features);

    void EstimateMinMaxDistance(double dist, int level, double& mn, double& mx) const;
    double PredictScaleLevel(double ref_dist, int ref_level, double dist) const;
    static bool PredictionConsistent(double predicted, int level)
    {
        double e = predicted - level;
        return e * e < (1.2 * 1.2);
    }
    double ScaleForContiniousLevel(double level) const;
};

// vision/features/OrbPattern.cpp:16-36
std::vector<int> AngleUmax();

// vision/features/OrbDescriptors.cpp:26-54
float ComputeAngle(const uint8_t* img, size_t pitch, float px, float py, const int* u_max);
// vision/features/OrbDescriptors.cpp:58-125
Descriptor ComputeDescriptor(const uint8_t* img, size_t pitch, float px, float py, float angle_degrees);

// vision/features/FeatureDistribution.cpp:72-189 (+ splitAndSort :15-70). Permutes `kps` in place.
std::vector<KeyPointF> Distribute(std::vector<KeyPointF>& kps, float min_x, float min_y, float max_x, float max_y,
                                  int target_n);

// cell geometry of ORBExtractor::DetectKeypoints, vision/features/ORBExtractor.cpp:99-127
struct CellGrid
{
    int nCols, nRows, wCell, hCell, minB, maxBX, maxBY;
};
CellGrid MakeCellGrid(int rows, int cols);

// vision/features/ORBExtractor.cpp:82-289
class OrbExtractor
{
   public:
    OrbExtractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int threads = 1);

    void Detect(const uint8_t* img, int rows, int cols, size_t pitch, std::vector<KeyPointF>& kps,
                std::vector<Descriptor>& desc);

    struct Level
    {
        int rows = 0, cols = 0;
        std::vector<uint8_t> image;  // dense rows*cols (the reference's 19-px apron is never read)
        std::vector<uint8_t> gauss;
        std::vector<KeyPointF> raw;   // FAST candidates before distribution (cell-row-major order)
        std::vector<KeyPointF> kept;  // after Distribute + (16,16) shift + angle
        int offset = 0;
    };
    std::vector<Level> levels;
    ScalePyramid pyramid;

   private:
    void AllocatePyramid(int rows, int cols);
    void ComputePyramid(const uint8_t* img, int rows, int cols, size_t pitch);
    void DetectKeypoints();
    int num_levels, th_fast, th_fast_min, num_threads;
    std::vector<int> u_max;
};
}  // namespace orbref
