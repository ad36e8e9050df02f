// Shadow of saiga/cuda/imageProcessing/ORBExtractorGPU.h (External/saiga/src/saiga/cuda/imageProcessing/
// ORBExtractorGPU.h:38-49): `Saiga::ORBExtractorGPU` with the original's constructor and
// `int Detect(ImageView<uchar>, vector<KeyPoint<float>>&, vector<DescriptorORB>&)` (returns the number of keypoints),
// backed by libsnakeorb.so. Used by Snake/Preprocess/FeatureDetector.cpp:31-33,119,149 when SNAKE_CUDA is defined and
// settings.fd_gpu is set. The original asserts pitchBytes % 4 == 0 and drops cols % 4 columns
// (ORBExtractorGPU.cu:110-113); this one takes any pitch and keeps every column (it follows the CPU extractor).
#pragma once
#include "../../vision/features/ORBExtractor.h"

namespace Saiga
{
class ORBExtractorGPU
{
   public:
    ORBExtractorGPU(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST)
        : impl(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
    {
    }
    int Detect(Saiga::ImageView<unsigned char> inputImage, std::vector<Saiga::KeyPoint<float>>& keypoints,
               std::vector<Saiga::DescriptorORB>& descriptors)
    {
        return impl.Detect(inputImage, keypoints, descriptors);
    }

   private:
    snkadapter::ORBExtractorB200 impl;
};
}  // namespace Saiga
