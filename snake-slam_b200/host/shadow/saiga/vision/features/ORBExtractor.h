// Shadow of saiga/vision/features/ORBExtractor.h: put `snake-slam_b200/host/shadow` in front of Saiga's own include
// directory and every translation unit that includes "saiga/vision/features/ORBExtractor.h" — in Snake-SLAM that is
// Snake/Preprocess/FeatureDetector.cpp:12 — gets a `Saiga::ORBExtractor` backed by libsnakeorb.so, with the public
// interface of the original (External/saiga/src/saiga/vision/features/ORBExtractor.h:44-63): the same constructor
// arguments, Detect, GetImage(level) and getPyramid(). FeatureDetector.cpp compiles against it unchanged
// (tests/test_adapter_reference_build.py does exactly that where /root/reference exists).
#pragma once
#ifndef SNAKEORB_WITH_SAIGA
#    define SNAKEORB_WITH_SAIGA
#endif
#include "saiga/core/image/templatedImage.h"

#include "../../../../snakeorb_saiga_adapter.h"

namespace Saiga
{
class ORBExtractor
{
   public:
    using KeypointType = Saiga::KeyPoint<float>;

    ORBExtractor(int nfeatures, float scaleFactor, int num_levels, int th_fast, int th_fast_min, int num_threads)
        : impl(nfeatures, scaleFactor, num_levels, th_fast, th_fast_min, num_threads), pyramid(num_levels, scaleFactor, nfeatures),
          mirror(num_levels)
    {
    }

    void Detect(Saiga::ImageView<unsigned char> inputImage, std::vector<KeypointType>& keypoints,
                std::vector<Saiga::DescriptorORB>& outputDescriptors)
    {
        impl.Detect(inputImage, keypoints, outputDescriptors);
        for (auto& m : mirror) m.valid = false;
    }

    // The scaled image of `level` after Detect; the view is invalidated by the next Detect, as in the original. The
    // pyramid lives on the device: the level is copied into a host mirror the first time it is asked for.
    ImageView<unsigned char> GetImage(int level)
    {
        auto& m = mirror[level];
        if (!m.valid)
        {
            int rows = 0, cols = 0;
            snkadapter::check(snk_orb_get_level(impl.handle(), 0, level, 0, nullptr, &rows, &cols, nullptr), "snk_orb_get_level");
            m.image.create(rows, cols);
            snkadapter::check(snk_orb_read_level(impl.handle(), 0, level, 0, m.image.data8(), m.image.pitchBytes), "snk_orb_read_level");
            m.valid = true;
        }
        return m.image.getImageView();
    }

    ScalePyramid getPyramid() { return pyramid; }

   protected:
    struct Mirror
    {
        Saiga::TemplatedImage<unsigned char> image;
        bool valid = false;
    };
    snkadapter::ORBExtractorB200 impl;
    Saiga::ScalePyramid pyramid;
    std::vector<Mirror> mirror;
};
}  // namespace Saiga
