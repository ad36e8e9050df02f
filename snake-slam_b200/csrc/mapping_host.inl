// Host wrappers of the local-mapping / loop-closing searches (included at the end of matchers.cu).
namespace
{
enum : unsigned
{
    UP_KP    = 1u,
    UP_DESC  = 2u,
    UP_RIGHT = 4u,
    UP_K2P   = 8u,
    UP_GRID  = 16u,
    UP_DEPTH = 32u
};
struct LiteFrame
{
    FrameDev f{};
    const float* d_depth = nullptr;
};
size_t lite_bytes(const snk_frame_view* v, unsigned what)
{
    const size_t n = std::max(v->n, 1), cells = (size_t)std::max(v->grid.rows * v->grid.cols, 1);
    size_t b = 0;
    if (what & UP_KP) b += padded<snk_keypoint_f64>(n);
    if (what & UP_DESC) b += padded<uint64_t>(4 * n);
    if (what & UP_RIGHT) b += padded<float>(n);
    if (what & UP_K2P) b += padded<int>(n);
    if (what & UP_GRID) b += padded<int>(2 * cells);
    if (what & UP_DEPTH) b += padded<float>(n);
    return b;
}
int check_lite(const snk_frame_view* v, unsigned what)
{
    SNK_REQUIRE(v && v->n >= 0, "bad frame view");
    if (v->n)
    {
        SNK_REQUIRE(!(what & UP_KP) || v->kp_un, "frame view without undistorted keypoints");
        SNK_REQUIRE(!(what & UP_DESC) || v->desc, "frame view without descriptors");
        SNK_REQUIRE(!(what & UP_RIGHT) || v->right_points, "frame view without right_points");
        SNK_REQUIRE(!(what & UP_K2P) || v->kp_to_point, "frame view without kp_to_point");
        SNK_REQUIRE(!(what & UP_DEPTH) || v->depth, "frame view without depth");
    }
    SNK_REQUIRE(!(what & UP_GRID) || (v->grid.rows >= 1 && v->grid.cols >= 1 && v->grid.cells), "bad feature grid");
    return SNK_OK;
}
int upload_lite(snk_ctx* c, Carver& cv, const snk_frame_view* v, unsigned what, LiteFrame& u)
{
    const size_t n = std::max(v->n, 1), cells = (size_t)std::max(v->grid.rows * v->grid.cols, 1);
    cudaStream_t st = c->stream;
    FrameDev& f = u.f;
    f.n         = v->n;
    if (what & UP_KP)
    {
        auto* p = cv.take<snk_keypoint_f64>(n);
        if (v->n) SNK_CUDA(cudaMemcpyAsync(p, v->kp_un, sizeof(snk_keypoint_f64) * (size_t)v->n, cudaMemcpyHostToDevice, st));
        f.kp_un = p;
    }
    if (what & UP_DESC)
    {
        auto* p = cv.take<uint64_t>(4 * n);
        if (v->n) SNK_CUDA(cudaMemcpyAsync(p, v->desc, 32 * (size_t)v->n, cudaMemcpyHostToDevice, st));
        f.desc = p;
    }
    if (what & UP_RIGHT)
    {
        auto* p = cv.take<float>(n);
        if (v->n) SNK_CUDA(cudaMemcpyAsync(p, v->right_points, 4 * (size_t)v->n, cudaMemcpyHostToDevice, st));
        f.right_points = p;
    }
    if (what & UP_K2P)
    {
        auto* p = cv.take<int>(n);
        if (v->n) SNK_CUDA(cudaMemcpyAsync(p, v->kp_to_point, 4 * (size_t)v->n, cudaMemcpyHostToDevice, st));
        f.kp_to_point = p;
    }
    if (what & UP_DEPTH)
    {
        auto* p = cv.take<float>(n);
        if (v->n) SNK_CUDA(cudaMemcpyAsync(p, v->depth, 4 * (size_t)v->n, cudaMemcpyHostToDevice, st));
        u.d_depth = p;
    }
    if (what & UP_GRID)
    {
        auto* p = cv.take<int>(2 * cells);
        SNK_CUDA(cudaMemcpyAsync(p, v->grid.cells, 8 * (size_t)v->grid.rows * v->grid.cols, cudaMemcpyHostToDevice, st));
        f.cells = p;
    }
    for (int i = 0; i < 4; ++i) f.q[i] = v->pose_q[i];
    for (int i = 0; i < 3; ++i) f.t[i] = v->pose_t[i];
    f.bmin_x = v->grid.bmin_x, f.bmin_y = v->grid.bmin_y, f.bmax_x = v->grid.bmax_x, f.bmax_y = v->grid.bmax_y;
    f.inv_x = v->grid.cell_size_inv_x, f.inv_y = v->grid.cell_size_inv_y;
    f.rows = v->grid.rows, f.cols = v->grid.cols;
    return SNK_OK;
}
// pinned staging for small per-point results would need a growable pinned buffer; the result arrays are pageable
// host memory of the caller, so a plain synchronous-at-the-end copy is used
int download_ints(snk_ctx* c, std::vector<int>& host, const int* dev, size_t n)
{
    host.resize(n);
    if (n) SNK_CUDA(cudaMemcpyAsync(host.data(), dev, 4 * n, cudaMemcpyDeviceToHost, c->stream));
    SNK_CUDA(cudaStreamSynchronize(c->stream));
    return SNK_OK;
}
Mat3Dev mat3(const double* m)
{
    Mat3Dev r;
    for (int i = 0; i < 9; ++i) r.m[i] = m[i];
    return r;
}
}  // namespace
extern "C" {

int snk_fuse_search(snk_ctx* c, const snk_frame_view* kf, const uint8_t* point_mask, const snk_point_fusion* points,
                    int n_points, const snk_camera* cam, const snk_scale_pyramid* pyr, float th, float obs_factor,
                    int feature_th, int32_t* cand_kp, int32_t* cand_id, int* n_candidates)
{
    SNK_REQUIRE(c && cam && pyr && n_candidates && n_points >= 0, "bad argument");
    const unsigned what = UP_KP | UP_DESC | UP_RIGHT | UP_GRID | UP_DEPTH;
    int rc = check_lite(kf, what);
    if (rc) return rc;
    SNK_REQUIRE(pyr->nlevels >= 1 && pyr->nlevels <= 16, "bad pyramid");
    *n_candidates = 0;
    if (n_points == 0) return SNK_OK;
    SNK_REQUIRE(points && cand_kp && cand_id, "null point arrays");
    for (int i = 0; i < n_points; ++i)
        SNK_REQUIRE(points[i].ref_level >= 0 && points[i].ref_level < pyr->nlevels, "point %d: reference level %d", i, points[i].ref_level);
    SNK_CUDA(cudaSetDevice(c->device));
    rc = ctx_reserve(c, lite_bytes(kf, what) + padded<snk_point_fusion>(n_points) + padded<uint8_t>(n_points) + padded<int>(n_points));
    if (rc) return rc;
    Carver cv{c->d_buf};
    LiteFrame u;
    if ((rc = upload_lite(c, cv, kf, what, u))) return rc;
    auto* dp = cv.take<snk_point_fusion>(n_points);
    auto* dm = cv.take<uint8_t>(n_points);
    int* db  = cv.take<int>(n_points);
    cudaStream_t st = c->stream;
    SNK_CUDA(cudaMemcpyAsync(dp, points, sizeof(snk_point_fusion) * (size_t)n_points, cudaMemcpyHostToDevice, st));
    if (point_mask) SNK_CUDA(cudaMemcpyAsync(dm, point_mask, n_points, cudaMemcpyHostToDevice, st));
    k_fuse<<<(n_points + 127) / 128, 128, 0, st>>>(u.f, u.d_depth, point_mask ? dm : nullptr, dp, n_points, *cam, *pyr, th,
                                                   obs_factor, feature_th, std::log(pyr->scale_factor), db);
    SNK_CUDA(cudaGetLastError());
    std::vector<int> best;
    if ((rc = download_ints(c, best, db, n_points))) return rc;
    int m = 0;
    for (int i = 0; i < n_points; ++i)  // fuseCandidates.emplace_back(bestIdx, lmp.id) in point order (:475-479)
        if (best[i] >= 0) cand_kp[m] = best[i], cand_id[m] = points[i].id, ++m;
    *n_candidates = m;
    return SNK_OK;
}

int snk_fuse_search_loop(snk_ctx* c, const snk_frame_view* kf, const uint8_t* skip, const snk_point_fusion* points,
                         int n_points, const snk_camera* cam, float th, int32_t* best_idx, int32_t* best_dist)
{
    SNK_REQUIRE(c && cam && n_points >= 0, "bad argument");
    const unsigned what = UP_KP | UP_DESC | UP_GRID;
    int rc = check_lite(kf, what);
    if (rc) return rc;
    if (n_points == 0) return SNK_OK;
    SNK_REQUIRE(points && best_idx && best_dist, "null point arrays");
    SNK_CUDA(cudaSetDevice(c->device));
    rc = ctx_reserve(c, lite_bytes(kf, what) + padded<snk_point_fusion>(n_points) + padded<uint8_t>(n_points) + 2 * padded<int>(n_points));
    if (rc) return rc;
    Carver cv{c->d_buf};
    LiteFrame u;
    if ((rc = upload_lite(c, cv, kf, what, u))) return rc;
    auto* dp = cv.take<snk_point_fusion>(n_points);
    auto* dm = cv.take<uint8_t>(n_points);
    int* db  = cv.take<int>(n_points);
    int* dd  = cv.take<int>(n_points);
    cudaStream_t st = c->stream;
    SNK_CUDA(cudaMemcpyAsync(dp, points, sizeof(snk_point_fusion) * (size_t)n_points, cudaMemcpyHostToDevice, st));
    if (skip) SNK_CUDA(cudaMemcpyAsync(dm, skip, n_points, cudaMemcpyHostToDevice, st));
    k_fuse_loop<<<(n_points + 127) / 128, 128, 0, st>>>(u.f, skip ? dm : nullptr, dp, n_points, *cam, th, db, dd);
    SNK_CUDA(cudaGetLastError());
    SNK_CUDA(cudaMemcpyAsync(best_idx, db, 4 * (size_t)n_points, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(best_dist, dd, 4 * (size_t)n_points, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaStreamSynchronize(st));
    return SNK_OK;
}

int snk_search_triangulation_bow(snk_ctx* c, const snk_frame_view* kf1, const double* np1, const snk_featvec_view* fv1,
                                 const snk_frame_view* kf2, const double* np2, const snk_featvec_view* fv2, const double* E,
                                 const snk_camera* cam, float epipolar_distance, int feature_distance, int32_t* pairs,
                                 int* n_pairs)
{
    SNK_REQUIRE(c && fv1 && fv2 && E && cam && n_pairs, "bad argument");
    const unsigned what = UP_DESC | UP_K2P;
    int rc = check_lite(kf1, what);
    if (rc) return rc;
    if ((rc = check_lite(kf2, what))) return rc;
    *n_pairs = 0;
    if (kf1->n == 0 || kf2->n == 0 || fv1->n_nodes <= 0 || fv2->n_nodes <= 0) return SNK_OK;
    SNK_REQUIRE(np1 && np2 && pairs, "null arrays");
    SNK_REQUIRE(fv1->node_ids && fv1->offsets && fv1->indices && fv2->node_ids && fv2->offsets && fv2->indices, "null feature vector");
    const int slots1 = fv1->offsets[fv1->n_nodes], slots2 = fv2->offsets[fv2->n_nodes];
    SNK_REQUIRE(slots1 >= 0 && slots1 <= kf1->n && slots2 >= 0 && slots2 <= kf2->n, "feature vector larger than its frame");
    if (slots1 == 0 || slots2 == 0) return SNK_OK;
    SNK_CUDA(cudaSetDevice(c->device));
    rc = ctx_reserve(c, lite_bytes(kf1, what) + lite_bytes(kf2, what) + padded<double>(2 * (size_t)kf1->n) +
                            padded<double>(2 * (size_t)kf2->n) + 2 * padded<int>(fv1->n_nodes + 1) + 2 * padded<int>(fv2->n_nodes + 1) +
                            padded<int>(slots1) + padded<int>(slots2) + padded<int>(slots1));
    if (rc) return rc;
    Carver cv{c->d_buf};
    LiteFrame a, b;
    if ((rc = upload_lite(c, cv, kf1, what, a))) return rc;
    if ((rc = upload_lite(c, cv, kf2, what, b))) return rc;
    cudaStream_t st = c->stream;
    auto up = [&](const void* src, size_t bytes) -> void* {
        void* d = cv.take<unsigned char>(bytes);
        cudaMemcpyAsync(d, src, bytes, cudaMemcpyHostToDevice, st);
        return d;
    };
    const double* dn1 = (const double*)up(np1, 16 * (size_t)kf1->n);
    const double* dn2 = (const double*)up(np2, 16 * (size_t)kf2->n);
    FeatVecDev g1{fv1->n_nodes, (const int*)up(fv1->node_ids, 4 * (size_t)fv1->n_nodes),
                  (const int*)up(fv1->offsets, 4 * (size_t)(fv1->n_nodes + 1)), (const int*)up(fv1->indices, 4 * (size_t)slots1)};
    FeatVecDev g2{fv2->n_nodes, (const int*)up(fv2->node_ids, 4 * (size_t)fv2->n_nodes),
                  (const int*)up(fv2->offsets, 4 * (size_t)(fv2->n_nodes + 1)), (const int*)up(fv2->indices, 4 * (size_t)slots2)};
    int* db = cv.take<int>(slots1);
    const double th_chi1 = epipolar_distance * 2 / cam->fx;  // float product, double quotient (:19)
    k_tri_bow<<<(slots1 + 127) / 128, 128, 0, st>>>(g1, g2, slots1, a.f.kp_to_point, b.f.kp_to_point, a.f.desc, b.f.desc, dn1, dn2,
                                                    mat3(E), th_chi1 * th_chi1, feature_distance, db);
    SNK_CUDA(cudaGetLastError());
    std::vector<int> best;
    if ((rc = download_ints(c, best, db, slots1))) return rc;
    int m = 0;
    for (int s = 0; s < slots1; ++s)  // vMatchedPairs.emplace_back(idx1, bestIdx2) in the order the merge visits idx1
        if (best[s] >= 0) pairs[2 * m] = fv1->indices[s], pairs[2 * m + 1] = best[s], ++m;
    *n_pairs = m;
    return SNK_OK;
}

int snk_search_triangulation_project(snk_ctx* c, const snk_frame_view* kf1, const double* np1, const snk_frame_view* kf2,
                                     const double* np2, const double* depth_grid, int grid_rows, int grid_cols,
                                     const double* E12, const snk_camera* cam, float epipolar_distance, int feature_distance,
                                     int32_t* pairs, int* n_pairs)
{
    SNK_REQUIRE(c && E12 && cam && n_pairs && depth_grid, "bad argument");
    const unsigned what1 = UP_KP | UP_DESC | UP_K2P, what2 = UP_KP | UP_DESC | UP_K2P | UP_GRID;
    int rc = check_lite(kf1, what1);
    if (rc) return rc;
    if ((rc = check_lite(kf2, what2))) return rc;
    SNK_REQUIRE(kf1->grid.rows >= 1 && kf1->grid.cols >= 1, "kf1 needs the feature grid bounds");
    // SAIGA_ASSERT(grid.rows() >= Rows / 4 && grid.cols() >= Cols / 4) (:177-178); the lookup reads (Rows-1)/4, (Cols-1)/4
    SNK_REQUIRE(grid_rows > (kf1->grid.rows - 1) / 4 && grid_cols > (kf1->grid.cols - 1) / 4, "depth grid %dx%d too small", grid_rows, grid_cols);
    *n_pairs = 0;
    if (kf1->n == 0) return SNK_OK;
    SNK_REQUIRE(np1 && (np2 || kf2->n == 0) && pairs, "null arrays");
    SNK_CUDA(cudaSetDevice(c->device));
    rc = ctx_reserve(c, lite_bytes(kf1, what1) + lite_bytes(kf2, what2) + padded<double>(2 * (size_t)kf1->n) +
                            padded<double>(2 * (size_t)std::max(kf2->n, 1)) + padded<double>((size_t)grid_rows * grid_cols) + padded<int>(kf1->n));
    if (rc) return rc;
    Carver cv{c->d_buf};
    LiteFrame a, b;
    if ((rc = upload_lite(c, cv, kf1, what1, a))) return rc;
    if ((rc = upload_lite(c, cv, kf2, what2, b))) return rc;
    cudaStream_t st = c->stream;
    double* dn1 = cv.take<double>(2 * (size_t)kf1->n);
    double* dn2 = cv.take<double>(2 * (size_t)std::max(kf2->n, 1));
    double* dg  = cv.take<double>((size_t)grid_rows * grid_cols);
    int* db     = cv.take<int>(kf1->n);
    SNK_CUDA(cudaMemcpyAsync(dn1, np1, 16 * (size_t)kf1->n, cudaMemcpyHostToDevice, st));
    if (kf2->n) SNK_CUDA(cudaMemcpyAsync(dn2, np2, 16 * (size_t)kf2->n, cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaMemcpyAsync(dg, depth_grid, 8 * (size_t)grid_rows * grid_cols, cudaMemcpyHostToDevice, st));
    const double th_chi1 = epipolar_distance / cam->fx;
    k_tri_project<<<(kf1->n + 127) / 128, 128, 0, st>>>(a.f, b.f, dn1, dn2, dg, grid_cols, mat3(E12), *cam, th_chi1 * th_chi1,
                                                        feature_distance, db);
    SNK_CUDA(cudaGetLastError());
    std::vector<int> best;
    if ((rc = download_ints(c, best, db, kf1->n))) return rc;
    int m = 0;
    for (int i = 0; i < kf1->n; ++i)
        if (best[i] >= 0) pairs[2 * m] = i, pairs[2 * m + 1] = best[i], ++m;
    *n_pairs = m;
    return SNK_OK;
}

int snk_match_prediction(snk_ctx* c, const uint64_t* desc_left, int n_left, double* last_positions, const snk_frame_view* right,
                         float radius, int feature_error, int32_t* matches, int* n_matches)
{
    SNK_REQUIRE(c && n_matches && n_left >= 0, "bad argument");
    const unsigned what = UP_KP | UP_DESC | UP_GRID;
    int rc = check_lite(right, what);
    if (rc) return rc;
    *n_matches = 0;
    if (n_left == 0) return SNK_OK;
    SNK_REQUIRE(desc_left && last_positions && matches, "null arrays");
    SNK_CUDA(cudaSetDevice(c->device));
    rc = ctx_reserve(c, lite_bytes(right, what) + padded<uint64_t>(4 * (size_t)n_left) + padded<double>(2 * (size_t)n_left) + padded<int>(n_left));
    if (rc) return rc;
    Carver cv{c->d_buf};
    LiteFrame r;
    if ((rc = upload_lite(c, cv, right, what, r))) return rc;
    cudaStream_t st = c->stream;
    auto* dd   = cv.take<uint64_t>(4 * (size_t)n_left);
    double* dp = cv.take<double>(2 * (size_t)n_left);
    int* dm    = cv.take<int>(n_left);
    SNK_CUDA(cudaMemcpyAsync(dd, desc_left, 32 * (size_t)n_left, cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaMemcpyAsync(dp, last_positions, 16 * (size_t)n_left, cudaMemcpyHostToDevice, st));
    k_match_prediction<<<(n_left + 127) / 128, 128, 0, st>>>(r.f, dd, n_left, dp, radius, feature_error, dm);
    SNK_CUDA(cudaGetLastError());
    SNK_CUDA(cudaMemcpyAsync(last_positions, dp, 16 * (size_t)n_left, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(matches, dm, 4 * (size_t)n_left, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaStreamSynchronize(st));
    int m = 0;
    for (int i = 0; i < n_left; ++i) m += matches[i] >= 0;
    *n_matches = m;
    return SNK_OK;
}
}  // extern "C"
