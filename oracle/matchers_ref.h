// TEST INFRASTRUCTURE ONLY — CPU oracle, never on the product path. See matchers_ref.cpp.
#pragma once
#include "../include/snakeorb.h"

namespace orbref
{
int hamming(const uint64_t* a, const uint64_t* b);
void match_knn2(const uint64_t* d1, int n, const uint64_t* d2, int m, snk_knn2* out, int threads);
int filter_matches(const snk_knn2* knn2, int n, int threshold, float ratio, int32_t* matches_ij);
int stereo_match(const snk_stereo_params* p, const snk_keypoint_f64* left, const uint64_t* dl, int nl,
                 const snk_keypoint_f64* right, const uint64_t* dr, int nr, float* right_points, float* depth,
                 long long* pair_ops);
int stereo_from_rgbd(const snk_rgbd_params* p, const snk_keypoint_f64* kp, int n, const float* depth_img, int rows,
                     int cols, size_t pitch, float* right_points, float* depth);
void undistort_keypoints(const snk_rectify* r, const snk_keypoint_f64* kps, int n, snk_keypoint_f64* kp_un,
                         double* normalized);
void feature_grid(const snk_grid_view* b, const snk_keypoint_f64* kps, int n, int32_t* permutation, int32_t* cells);
int search_projection_fine(const snk_frame_view* f, snk_point_fine* pts, int n, const snk_camera* K,
                           const snk_scale_pyramid* sp, float th, float ratio, int32_t* kp_to_point_out,
                           int32_t* best_idx, int32_t* best_dist, uint8_t* visible, long long* pair_ops);
int search_projection_coarse(const snk_frame_view* f, const snk_point_coarse* pts, int n, const snk_camera* K,
                             const snk_scale_pyramid* sp, float th, int feature_error, int motion,
                             int32_t* kp_to_point_out, long long* pair_ops);
int search_projection_frame_to_kf(const snk_frame_view* f, const snk_point_kf* pts, int n, const snk_camera* K,
                                  float th, int feature_error, int32_t* kp_to_point_out, long long* pair_ops);
int match_bow(const uint64_t* d1, int n1, const uint64_t* d2, int n2, const snk_featvec_view* fv1,
              const snk_featvec_view* fv2, const uint8_t* valid1, const uint8_t* valid2, int th, float ratio,
              int32_t* match12, long long* pair_ops);
}  // namespace orbref
