// TEST INFRASTRUCTURE ONLY. /root/reference/Snake/Preprocess/Preprocess.cpp includes <opencv2/imgproc.hpp>; OpenCV's C++
// headers are absent from this image and nothing in that translation unit uses a cv:: name, so an empty header lets the
// reference file compile unmodified (oracle/Makefile, _ref/libsnake_ref.so).
#pragma once
