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
double epipolar_distance_squared(const double* p1, const double* p2, const double* F);
int fuse_search(const snk_frame_view* kf, const uint8_t* point_mask, const snk_point_fusion* pts, int n,
                const snk_camera* K, const snk_scale_pyramid* sp, float th, float obs_factor, int feature_th,
                int32_t* cand_kp, int32_t* cand_id, long long* pair_ops);
void fuse_search_loop(const snk_frame_view* kf, const uint8_t* skip, const snk_point_fusion* pts, int n,
                      const snk_camera* K, float th, int32_t* best_idx, int32_t* best_dist);
int search_triangulation_bow(const snk_frame_view* kf1, const double* np1, const snk_featvec_view* fv1,
                             const snk_frame_view* kf2, const double* np2, const snk_featvec_view* fv2, const double* E,
                             const snk_camera* K, float epipolarDistance, int featureDistance, int32_t* pairs,
                             long long* pair_ops);
int search_triangulation_project(const snk_frame_view* kf1, const double* np1, const snk_frame_view* kf2,
                                 const double* np2, const double* grid, int grid_rows, int grid_cols, const double* E12,
                                 const snk_camera* K, float epipolarDistance, int featureDistance, int32_t* pairs,
                                 long long* pair_ops);
int match_prediction(const uint64_t* desc_left, int n_left, double* last_positions, const snk_frame_view* right,
                     float radius, int featureError, int32_t* matches);
int features_in_area(const snk_frame_view* f, double px, double py, float r, int32_t* out);
void grid_bounds(int w, int h, const double* K, const double* dist, int resolution, snk_grid_view* out);
}  // namespace orbref
