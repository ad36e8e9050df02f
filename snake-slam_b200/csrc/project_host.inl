// Host wrappers of the projection searches and BoW matching (included at the end of matchers.cu).
namespace
{
struct FrameUpload
{
    FrameDev f{};
    int* d_k2p_out = nullptr;
    int* d_winner  = nullptr;
    size_t bytes(const snk_frame_view* v) const
    {
        const size_t n = std::max(v->n, 1), cells = (size_t)std::max(v->grid.rows * v->grid.cols, 1);
        return padded<snk_keypoint_f64>(n) + padded<uint64_t>(4 * n) + padded<float>(n) + 3 * padded<int>(n) +
               padded<int>(2 * cells);
    }
};

int upload_frame(snk_ctx* c, Carver& cv, const snk_frame_view* v, FrameUpload& u)
{
    const size_t n = std::max(v->n, 1), cells = (size_t)std::max(v->grid.rows * v->grid.cols, 1);
    auto* kp  = cv.take<snk_keypoint_f64>(n);
    auto* ds  = cv.take<uint64_t>(4 * n);
    auto* rp  = cv.take<float>(n);
    auto* k2p = cv.take<int>(n);
    u.d_k2p_out = cv.take<int>(n);
    u.d_winner  = cv.take<int>(n);
    auto* cl  = cv.take<int>(2 * cells);
    cudaStream_t st = c->stream;
    if (v->n)
    {
        SNK_CUDA(cudaMemcpyAsync(kp, v->kp_un, sizeof(snk_keypoint_f64) * (size_t)v->n, cudaMemcpyHostToDevice, st));
        SNK_CUDA(cudaMemcpyAsync(ds, v->desc, 32 * (size_t)v->n, cudaMemcpyHostToDevice, st));
        SNK_CUDA(cudaMemcpyAsync(rp, v->right_points, 4 * (size_t)v->n, cudaMemcpyHostToDevice, st));
        SNK_CUDA(cudaMemcpyAsync(k2p, v->kp_to_point, 4 * (size_t)v->n, cudaMemcpyHostToDevice, st));
        SNK_CUDA(cudaMemcpyAsync(u.d_k2p_out, v->kp_to_point, 4 * (size_t)v->n, cudaMemcpyHostToDevice, st));
        SNK_CUDA(cudaMemsetAsync(u.d_winner, 0x7f, 4 * (size_t)v->n, st));  // 0x7f7f7f7f: replaced below
    }
    SNK_CUDA(cudaMemcpyAsync(cl, v->grid.cells, 8 * (size_t)v->grid.rows * v->grid.cols, cudaMemcpyHostToDevice, st));
    FrameDev& f = u.f;
    f.n = v->n, f.kp_un = kp, f.desc = ds, f.right_points = rp, f.kp_to_point = k2p;
    for (int i = 0; i < 4; ++i) f.q[i] = v->pose_q[i];
    for (int i = 0; i < 3; ++i) f.t[i] = v->pose_t[i];
    f.bmin_x = v->grid.bmin_x, f.bmin_y = v->grid.bmin_y, f.bmax_x = v->grid.bmax_x, f.bmax_y = v->grid.bmax_y;
    f.inv_x = v->grid.cell_size_inv_x, f.inv_y = v->grid.cell_size_inv_y;
    f.rows = v->grid.rows, f.cols = v->grid.cols, f.cells = cl;
    return SNK_OK;
}

__global__ void k_gather_valid(const snk_point_fine* p, int n, int* out)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = p[i].valid;
}

__global__ void k_fill_int(int* p, int n, int v)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

int check_frame(const snk_frame_view* f)
{
    SNK_REQUIRE(f && f->n >= 0, "bad frame view");
    SNK_REQUIRE(f->grid.rows >= 1 && f->grid.cols >= 1 && f->grid.cells, "bad feature grid");
    SNK_REQUIRE(f->n == 0 || (f->kp_un && f->desc && f->right_points && f->kp_to_point), "null frame arrays");
    return SNK_OK;
}
}  // namespace

extern "C" {

int snk_search_projection_fine(snk_ctx* c, const snk_frame_view* frame, snk_point_fine* points, int n_points,
                               const snk_camera* cam, const snk_scale_pyramid* pyr, float th, float ratio,
                               int32_t* kp_to_point_out, int32_t* best_idx, int32_t* best_dist, uint8_t* visible,
                               int* n_matches)
{
    SNK_REQUIRE(c && cam && pyr && n_matches && n_points >= 0, "bad argument");
    int rc = check_frame(frame);
    if (rc) return rc;
    SNK_REQUIRE(n_points == 0 || (points && best_idx && best_dist && visible), "null point arrays");
    SNK_REQUIRE(frame->n == 0 || kp_to_point_out, "null output");
    SNK_REQUIRE(pyr->nlevels >= 1 && pyr->nlevels <= 16, "bad pyramid");
    for (int i = 0; i < n_points; ++i)
        SNK_REQUIRE(points[i].ref_level >= 0 && points[i].ref_level < pyr->nlevels, "ref_level out of range");
    SNK_CUDA(cudaSetDevice(c->device));
    FrameUpload u;
    const size_t np = std::max(n_points, 1);
    rc = ctx_reserve(c, u.bytes(frame) + padded<snk_point_fine>(np) + 2 * padded<int>(np) + padded<uint8_t>(np) + padded<int>(4));
    if (rc) return rc;
    Carver cv{c->d_buf};
    rc = upload_frame(c, cv, frame, u);
    if (rc) return rc;
    auto* dp = cv.take<snk_point_fine>(np);
    int* dbi = cv.take<int>(np);
    int* dbd = cv.take<int>(np);
    auto* dv = cv.take<uint8_t>(np);
    int* dn  = cv.take<int>(4);
    cudaStream_t st = c->stream;
    SNK_CUDA(cudaMemsetAsync(dn, 0, 16, st));
    if (frame->n) k_fill_int<<<(frame->n + 255) / 256, 256, 0, st>>>(u.d_winner, frame->n, 0x7fffffff);
    if (n_points)
    {
        SNK_CUDA(cudaMemcpyAsync(dp, points, sizeof(snk_point_fine) * (size_t)n_points, cudaMemcpyHostToDevice, st));
        k_proj_fine<<<(n_points + 127) / 128, 128, 0, st>>>(u.f, dp, n_points, *cam, *pyr, th, ratio,
                                                           std::log(pyr->scale_factor), dbi, dbd, dv, u.d_winner);
    }
    if (frame->n)
        k_assign<snk_point_fine><<<(frame->n + 255) / 256, 256, 0, st>>>(u.f.kp_to_point, u.d_winner, dp, nullptr, frame->n,
                                                                        u.d_k2p_out, nullptr, dn);
    if (n_points)
    {
        SNK_CUDA(cudaMemcpyAsync(points, dp, sizeof(snk_point_fine) * (size_t)n_points, cudaMemcpyDeviceToHost, st));
        SNK_CUDA(cudaMemcpyAsync(best_idx, dbi, 4 * (size_t)n_points, cudaMemcpyDeviceToHost, st));
        SNK_CUDA(cudaMemcpyAsync(best_dist, dbd, 4 * (size_t)n_points, cudaMemcpyDeviceToHost, st));
        SNK_CUDA(cudaMemcpyAsync(visible, dv, (size_t)n_points, cudaMemcpyDeviceToHost, st));
    }
    if (frame->n) SNK_CUDA(cudaMemcpyAsync(kp_to_point_out, u.d_k2p_out, 4 * (size_t)frame->n, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(c->h_pin + 4, dn, 4, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaStreamSynchronize(st));
    *n_matches = c->h_pin[4];
    return SNK_OK;
}

int snk_search_projection_coarse(snk_ctx* c, const snk_frame_view* frame, const snk_point_coarse* points, int n_points,
                                 const snk_camera* cam, const snk_scale_pyramid* pyr, float th, int feature_error,
                                 int motion, int32_t* kp_to_point_out, int* n_matches)
{
    SNK_REQUIRE(c && cam && pyr && n_matches && n_points >= 0, "bad argument");
    int rc = check_frame(frame);
    if (rc) return rc;
    SNK_REQUIRE(n_points == 0 || points, "null point array");
    SNK_REQUIRE(frame->n == 0 || kp_to_point_out, "null output");
    SNK_REQUIRE(pyr->nlevels >= 1 && pyr->nlevels <= 16, "bad pyramid");
    for (int i = 0; i < n_points; ++i)
        SNK_REQUIRE(points[i].octave >= 0 && points[i].octave < pyr->nlevels, "octave out of range");
    SNK_CUDA(cudaSetDevice(c->device));
    FrameUpload u;
    const size_t np = std::max(n_points, 1);
    rc = ctx_reserve(c, u.bytes(frame) + padded<snk_point_coarse>(np) + 2 * padded<int>(np) + padded<int>(64));
    if (rc) return rc;
    Carver cv{c->d_buf};
    rc = upload_frame(c, cv, frame, u);
    if (rc) return rc;
    auto* dp = cv.take<snk_point_coarse>(np);
    int* dbi = cv.take<int>(np);
    int* dbb = cv.take<int>(np);
    int* dn  = cv.take<int>(64);  // [0] matches, [32..61] histogram
    cudaStream_t st = c->stream;
    SNK_CUDA(cudaMemsetAsync(dn, 0, 256, st));
    if (frame->n) k_fill_int<<<(frame->n + 255) / 256, 256, 0, st>>>(u.d_winner, frame->n, 0x7fffffff);
    if (n_points)
    {
        SNK_CUDA(cudaMemcpyAsync(dp, points, sizeof(snk_point_coarse) * (size_t)n_points, cudaMemcpyHostToDevice, st));
        k_proj_coarse<<<(n_points + 127) / 128, 128, 0, st>>>(u.f, dp, n_points, *cam, *pyr, th, feature_error, motion, dbi,
                                                             dbb, u.d_winner);
    }
    if (frame->n)
    {
        k_assign<snk_point_coarse><<<(frame->n + 255) / 256, 256, 0, st>>>(u.f.kp_to_point, u.d_winner, dp, dbb, frame->n,
                                                                          u.d_k2p_out, dn + 32, dn);
        k_rot_cull<<<1, 256, 0, st>>>(u.f.kp_to_point, u.d_winner, dbb, frame->n, dn + 32, u.d_k2p_out, dn);
        SNK_CUDA(cudaMemcpyAsync(kp_to_point_out, u.d_k2p_out, 4 * (size_t)frame->n, cudaMemcpyDeviceToHost, st));
    }
    SNK_CUDA(cudaMemcpyAsync(c->h_pin + 4, dn, 4, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaStreamSynchronize(st));
    *n_matches = c->h_pin[4];
    return SNK_OK;
}

int snk_search_projection_frame_to_kf(snk_ctx* c, const snk_frame_view* frame, const snk_point_kf* points, int n_points,
                                      const snk_camera* cam, float th, int feature_error, int32_t* kp_to_point_out,
                                      int* n_matches)
{
    SNK_REQUIRE(c && cam && n_matches && n_points >= 0, "bad argument");
    int rc = check_frame(frame);
    if (rc) return rc;
    SNK_REQUIRE(n_points == 0 || points, "null point array");
    SNK_REQUIRE(frame->n == 0 || kp_to_point_out, "null output");
    SNK_CUDA(cudaSetDevice(c->device));
    FrameUpload u;
    const size_t np = std::max(n_points, 1);
    rc = ctx_reserve(c, u.bytes(frame) + padded<snk_point_kf>(np) + padded<int>(4));
    if (rc) return rc;
    Carver cv{c->d_buf};
    rc = upload_frame(c, cv, frame, u);
    if (rc) return rc;
    auto* dp = cv.take<snk_point_kf>(np);
    int* dn  = cv.take<int>(4);
    cudaStream_t st = c->stream;
    SNK_CUDA(cudaMemsetAsync(dn, 0, 16, st));
    if (n_points)
    {
        SNK_CUDA(cudaMemcpyAsync(dp, points, sizeof(snk_point_kf) * (size_t)n_points, cudaMemcpyHostToDevice, st));
        k_proj_frame_to_kf<<<1, 32, 0, st>>>(u.f, dp, n_points, *cam, th, feature_error, u.d_k2p_out, dn);
    }
    if (frame->n) SNK_CUDA(cudaMemcpyAsync(kp_to_point_out, u.d_k2p_out, 4 * (size_t)frame->n, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(c->h_pin + 4, dn, 4, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaStreamSynchronize(st));
    *n_matches = c->h_pin[4];
    return SNK_OK;
}

int snk_feature_grid(snk_ctx* c, const snk_grid_view* b, const snk_keypoint_f64* kp_un, int n, int32_t* permutation,
                     int32_t* cells_out)
{
    SNK_REQUIRE(c && b && cells_out && n >= 0 && (n == 0 || (kp_un && permutation)), "bad argument");
    SNK_REQUIRE(b->rows >= 1 && b->cols >= 1 && (long long)b->rows * b->cols < (1 << 18), "bad grid size");
    SNK_REQUIRE(n <= 8192, "snk_feature_grid handles at most 8192 keypoints");
    const int ncell = b->rows * b->cols;
    SNK_CUDA(cudaSetDevice(c->device));
    const size_t nn = std::max(n, 1);
    int rc = ctx_reserve(c, padded<snk_keypoint_f64>(nn) + padded<int>(nn) + padded<int>(2 * (size_t)ncell));
    if (rc) return rc;
    Carver cv{c->d_buf};
    auto* dk = cv.take<snk_keypoint_f64>(nn);
    int* dp  = cv.take<int>(nn);
    int* dc  = cv.take<int>(2 * (size_t)ncell);
    cudaStream_t st = c->stream;
    if (n) SNK_CUDA(cudaMemcpyAsync(dk, kp_un, sizeof(snk_keypoint_f64) * (size_t)n, cudaMemcpyHostToDevice, st));
    FrameDev f{};
    f.n = n, f.kp_un = dk;
    f.bmin_x = b->bmin_x, f.bmin_y = b->bmin_y, f.bmax_x = b->bmax_x, f.bmax_y = b->bmax_y;
    f.inv_x = b->cell_size_inv_x, f.inv_y = b->cell_size_inv_y, f.rows = b->rows, f.cols = b->cols;
    int np2 = 1;
    while (np2 < n) np2 <<= 1;
    k_feature_grid<<<1, 256, (size_t)np2 * 4, st>>>(f, n, np2, dp, dc);
    if (n) SNK_CUDA(cudaMemcpyAsync(permutation, dp, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(cells_out, dc, 8 * (size_t)ncell, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaStreamSynchronize(st));
    return SNK_OK;
}

int snk_match_bow(snk_ctx* c, const uint64_t* d1, int n1, const uint64_t* d2, int n2, const snk_featvec_view* fv1,
                  const snk_featvec_view* fv2, const uint8_t* valid1, const uint8_t* valid2, int th, float ratio,
                  int32_t* match12, int* n_matches)
{
    SNK_REQUIRE(c && fv1 && fv2 && n_matches && n1 >= 0 && n2 >= 0, "bad argument");
    SNK_REQUIRE(n1 == 0 || (d1 && valid1 && match12), "null keyframe-1 arrays");
    SNK_REQUIRE(n2 == 0 || (d2 && valid2), "null keyframe-2 arrays");
    *n_matches = 0;
    for (int i = 0; i < n1; ++i) match12[i] = -1;
    if (n1 == 0 || n2 == 0 || fv1->n_nodes == 0 || fv2->n_nodes == 0) return SNK_OK;
    const int e1 = fv1->offsets[fv1->n_nodes], e2 = fv2->offsets[fv2->n_nodes];
    SNK_REQUIRE(e2 < (1 << 20), "feature vector too long");
    for (int i = 0; i < e1; ++i) SNK_REQUIRE(fv1->indices[i] >= 0 && fv1->indices[i] < n1, "fv1 index out of range");
    for (int i = 0; i < e2; ++i) SNK_REQUIRE(fv2->indices[i] >= 0 && fv2->indices[i] < n2, "fv2 index out of range");
    SNK_CUDA(cudaSetDevice(c->device));
    int rc = ctx_reserve(c, padded<uint64_t>(4 * (size_t)n1) + padded<uint64_t>(4 * (size_t)n2) + padded<uint8_t>(n1) +
                                2 * padded<uint8_t>(n2) + padded<int>(n1) + 2 * padded<int>(fv1->n_nodes + 1) +
                                2 * padded<int>(fv2->n_nodes + 1) + padded<int>(e1 + 1) + padded<int>(e2 + 1) + padded<int>(4));
    if (rc) return rc;
    Carver cv{c->d_buf};
    auto* a   = cv.take<uint64_t>(4 * (size_t)n1);
    auto* b   = cv.take<uint64_t>(4 * (size_t)n2);
    auto* v1  = cv.take<uint8_t>(n1);
    auto* v2  = cv.take<uint8_t>(n2);
    auto* m2  = cv.take<uint8_t>(n2);
    int* m12  = cv.take<int>(n1);
    int* ids1 = cv.take<int>(fv1->n_nodes + 1);
    int* of1  = cv.take<int>(fv1->n_nodes + 1);
    int* ids2 = cv.take<int>(fv2->n_nodes + 1);
    int* of2  = cv.take<int>(fv2->n_nodes + 1);
    int* in1  = cv.take<int>(e1 + 1);
    int* in2  = cv.take<int>(e2 + 1);
    int* dn   = cv.take<int>(4);
    cudaStream_t st = c->stream;
    SNK_CUDA(cudaMemcpyAsync(a, d1, 32 * (size_t)n1, cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaMemcpyAsync(b, d2, 32 * (size_t)n2, cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaMemcpyAsync(v1, valid1, n1, cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaMemcpyAsync(v2, valid2, n2, cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaMemsetAsync(m2, 0, n2, st));
    SNK_CUDA(cudaMemsetAsync(m12, 0xff, 4 * (size_t)n1, st));
    SNK_CUDA(cudaMemsetAsync(dn, 0, 16, st));
    SNK_CUDA(cudaMemcpyAsync(ids1, fv1->node_ids, 4 * (size_t)fv1->n_nodes, cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaMemcpyAsync(of1, fv1->offsets, 4 * (size_t)(fv1->n_nodes + 1), cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaMemcpyAsync(ids2, fv2->node_ids, 4 * (size_t)fv2->n_nodes, cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaMemcpyAsync(of2, fv2->offsets, 4 * (size_t)(fv2->n_nodes + 1), cudaMemcpyHostToDevice, st));
    if (e1) SNK_CUDA(cudaMemcpyAsync(in1, fv1->indices, 4 * (size_t)e1, cudaMemcpyHostToDevice, st));
    if (e2) SNK_CUDA(cudaMemcpyAsync(in2, fv2->indices, 4 * (size_t)e2, cudaMemcpyHostToDevice, st));
    FeatVecDev f1{fv1->n_nodes, ids1, of1, in1}, f2{fv2->n_nodes, ids2, of2, in2};
    k_match_bow<<<(fv1->n_nodes + 3) / 4, 128, 0, st>>>(a, b, f1, f2, v1, v2, th, ratio, m12, m2, dn);
    SNK_CUDA(cudaMemcpyAsync(match12, m12, 4 * (size_t)n1, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(c->h_pin + 4, dn, 4, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaStreamSynchronize(st));
    *n_matches = c->h_pin[4];
    return SNK_OK;
}

// One keyframe against n_cand candidates in one upload, one launch and one download.
int snk_match_bow_batch(snk_ctx* c, const uint64_t* d1, int n1, const snk_featvec_view* fv1, const uint8_t* valid1,
                        const snk_bow_candidate* cands, int n_cand, int th, float ratio, int32_t* match12, int32_t* n_matches)
{
    SNK_REQUIRE(c && fv1 && n1 >= 0 && n_cand >= 0, "bad argument");
    SNK_REQUIRE(n_cand == 0 || (cands && n_matches), "null candidate arrays");
    SNK_REQUIRE(n1 == 0 || n_cand == 0 || (d1 && valid1 && match12), "null keyframe-1 arrays");
    SNK_REQUIRE(n_cand <= 65535, "at most 65535 candidates per call");
    for (int k = 0; k < n_cand; ++k) n_matches[k] = 0;
    for (size_t i = 0; i < (size_t)n_cand * n1; ++i) match12[i] = -1;
    if (n1 == 0 || n_cand == 0 || fv1->n_nodes == 0) return SNK_OK;
    const int e1 = fv1->offsets[fv1->n_nodes];
    for (int i = 0; i < e1; ++i) SNK_REQUIRE(fv1->indices[i] >= 0 && fv1->indices[i] < n1, "fv1 index out of range");
    // layout of the blob: keyframe 1 (descriptors, valid, node ids, offsets, indices), the candidate table, then each
    // candidate's arrays; every array 16-byte aligned
    size_t off = 0, m2 = 0;
    auto put   = [&](size_t bytes) {
        const size_t at = off;
        off             = align_up(off + bytes, (size_t)16);
        return at;
    };
    const size_t o_d1 = put(32 * (size_t)n1), o_v1 = put(n1), o_ids1 = put(4 * (size_t)fv1->n_nodes),
                 o_of1 = put(4 * (size_t)(fv1->n_nodes + 1)), o_in1 = put(4 * (size_t)e1), o_tab = put(sizeof(BowCandDev) * (size_t)n_cand);
    std::vector<BowCandDev> tab(n_cand);
    for (int k = 0; k < n_cand; ++k)
    {
        const snk_bow_candidate& q = cands[k];
        SNK_REQUIRE(q.n >= 0 && q.fv && (q.n == 0 || (q.desc && q.valid)), "candidate %d: null arrays", k);
        const int nn = q.n == 0 ? 0 : q.fv->n_nodes, e2 = nn ? q.fv->offsets[nn] : 0;
        SNK_REQUIRE(e2 < (1 << 20), "candidate %d: feature vector too long", k);
        for (int i = 0; i < e2; ++i) SNK_REQUIRE(q.fv->indices[i] >= 0 && q.fv->indices[i] < q.n, "candidate %d: fv index out of range", k);
        BowCandDev& t = tab[k];
        t.n_nodes = nn, t.n2 = q.n, t.m2_off = (int)m2;
        t.desc_off  = (int)put(32 * (size_t)q.n);
        t.valid_off = (int)put(q.n);
        t.ids_off   = (int)put(4 * (size_t)nn);
        t.offs_off  = (int)put(4 * (size_t)(nn + 1));
        t.inds_off  = (int)put(4 * (size_t)e2);
        m2 += align_up((size_t)q.n, (size_t)16);
        SNK_REQUIRE(off < ((size_t)1 << 31), "batch too large (2 GiB of candidate data)");
    }
    SNK_CUDA(cudaSetDevice(c->device));
    if (off > c->h_stage_cap)
    {
        SNK_CUDA(cudaStreamSynchronize(c->stream));
        if (c->h_stage) SNK_CUDA(cudaFreeHost(c->h_stage));
        c->h_stage = nullptr, c->h_stage_cap = 0;
        SNK_CUDA(cudaMallocHost(&c->h_stage, off + off / 4));
        c->h_stage_cap = off + off / 4;
    }
    unsigned char* h = c->h_stage;
    auto cp          = [&](size_t at, const void* src, size_t bytes) {
        if (bytes) std::memcpy(h + at, src, bytes);
    };
    cp(o_d1, d1, 32 * (size_t)n1), cp(o_v1, valid1, n1), cp(o_ids1, fv1->node_ids, 4 * (size_t)fv1->n_nodes);
    cp(o_of1, fv1->offsets, 4 * (size_t)(fv1->n_nodes + 1)), cp(o_in1, fv1->indices, 4 * (size_t)e1);
    cp(o_tab, tab.data(), sizeof(BowCandDev) * (size_t)n_cand);
    for (int k = 0; k < n_cand; ++k)
    {
        const snk_bow_candidate& q = cands[k];
        const BowCandDev& t        = tab[k];
        cp(t.desc_off, q.desc, 32 * (size_t)q.n), cp(t.valid_off, q.valid, q.n);
        if (t.n_nodes)
        {
            cp(t.ids_off, q.fv->node_ids, 4 * (size_t)t.n_nodes), cp(t.offs_off, q.fv->offsets, 4 * (size_t)(t.n_nodes + 1));
            cp(t.inds_off, q.fv->indices, 4 * (size_t)q.fv->offsets[t.n_nodes]);
        }
    }
    int rc = ctx_reserve(c, padded<unsigned char>(off) + padded<uint8_t>(std::max(m2, (size_t)1)) + padded<int>((size_t)n_cand * n1) +
                                padded<int>(n_cand));
    if (rc) return rc;
    Carver cv{c->d_buf};
    auto* blob = cv.take<unsigned char>(off);
    auto* dm2  = cv.take<uint8_t>(std::max(m2, (size_t)1));
    int* m12   = cv.take<int>((size_t)n_cand * n1);
    int* dn    = cv.take<int>(n_cand);
    cudaStream_t st = c->stream;
    SNK_CUDA(cudaMemcpyAsync(blob, h, off, cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaMemsetAsync(dm2, 0, std::max(m2, (size_t)1), st));
    SNK_CUDA(cudaMemsetAsync(m12, 0xff, 4 * (size_t)n_cand * n1, st));
    SNK_CUDA(cudaMemsetAsync(dn, 0, 4 * (size_t)n_cand, st));
    const FeatVecDev f1{fv1->n_nodes, reinterpret_cast<const int*>(blob + o_ids1), reinterpret_cast<const int*>(blob + o_of1),
                        reinterpret_cast<const int*>(blob + o_in1)};
    k_match_bow_batch<<<dim3((fv1->n_nodes + 3) / 4, n_cand), 128, 0, st>>>(blob, reinterpret_cast<const BowCandDev*>(blob + o_tab),
                                                                             reinterpret_cast<const uint64_t*>(blob + o_d1), f1, blob + o_v1, n1,
                                                                             th, ratio, m12, dm2, dn);
    SNK_CUDA(cudaGetLastError());
    SNK_CUDA(cudaMemcpyAsync(match12, m12, 4 * (size_t)n_cand * n1, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(n_matches, dn, 4 * (size_t)n_cand, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaStreamSynchronize(st));
    return SNK_OK;
}


// ---- device-resident frame and local map (no reference counterpart: there both live in host memory, Snake/Map/Frame.h,
// Snake/Map/LocalMap.h:83-186) -----------------------------------------------------------------------------------------
struct snk_frame_dev
{
    snk_ctx* c = nullptr;
    FrameDev f{};
    int cap_n = 0, cap_cells = 0;
    unsigned char* buf = nullptr;
    int *k2p = nullptr, *k2p_out = nullptr, *winner = nullptr;
};
struct snk_localmap
{
    snk_ctx* c        = nullptr;
    snk_point_fine* d = nullptr;
    int *bi = nullptr, *bd = nullptr;
    uint8_t* vis = nullptr;
    int capacity = 0, n = 0;
};

int snk_frame_dev_create(snk_ctx* c, const snk_frame_view* v, snk_frame_dev** out)
{
    SNK_REQUIRE(c && out, "null argument");
    int rc = check_frame(v);
    if (rc) return rc;
    SNK_CUDA(cudaSetDevice(c->device));
    auto* fd = new snk_frame_dev;
    fd->c    = c;
    const size_t n = std::max(v->n, 1), cells = (size_t)v->grid.rows * v->grid.cols;
    const size_t bytes = padded<snk_keypoint_f64>(n) + padded<uint64_t>(4 * n) + padded<float>(n) + 3 * padded<int>(n) + padded<int>(2 * cells);
    if (cudaMalloc(&fd->buf, bytes) != cudaSuccess)
    {
        set_error("snk_frame_dev_create: cudaMalloc(%zu) failed", bytes);
        delete fd;
        return SNK_ERR_CUDA;
    }
    Carver cv{fd->buf};
    auto* kp = cv.take<snk_keypoint_f64>(n);
    auto* ds = cv.take<uint64_t>(4 * n);
    auto* rp = cv.take<float>(n);
    fd->k2p     = cv.take<int>(n);
    fd->k2p_out = cv.take<int>(n);
    fd->winner  = cv.take<int>(n);
    auto* cl    = cv.take<int>(2 * cells);
    cudaStream_t st = c->stream;
    if (v->n)
    {
        SNK_CUDA(cudaMemcpyAsync(kp, v->kp_un, sizeof(snk_keypoint_f64) * (size_t)v->n, cudaMemcpyHostToDevice, st));
        SNK_CUDA(cudaMemcpyAsync(ds, v->desc, 32 * (size_t)v->n, cudaMemcpyHostToDevice, st));
        SNK_CUDA(cudaMemcpyAsync(rp, v->right_points, 4 * (size_t)v->n, cudaMemcpyHostToDevice, st));
        SNK_CUDA(cudaMemcpyAsync(fd->k2p, v->kp_to_point, 4 * (size_t)v->n, cudaMemcpyHostToDevice, st));
    }
    SNK_CUDA(cudaMemcpyAsync(cl, v->grid.cells, 8 * cells, cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaStreamSynchronize(st));  // the caller's arrays may go away after this call
    FrameDev& f = fd->f;
    f.n = v->n, f.kp_un = kp, f.desc = ds, f.right_points = rp, f.kp_to_point = fd->k2p;
    for (int i = 0; i < 4; ++i) f.q[i] = v->pose_q[i];
    for (int i = 0; i < 3; ++i) f.t[i] = v->pose_t[i];
    f.bmin_x = v->grid.bmin_x, f.bmin_y = v->grid.bmin_y, f.bmax_x = v->grid.bmax_x, f.bmax_y = v->grid.bmax_y;
    f.inv_x = v->grid.cell_size_inv_x, f.inv_y = v->grid.cell_size_inv_y;
    f.rows = v->grid.rows, f.cols = v->grid.cols, f.cells = cl;
    *out = fd;
    return SNK_OK;
}

void snk_frame_dev_destroy(snk_frame_dev* f)
{
    if (!f) return;
    cudaSetDevice(f->c->device);
    cudaStreamSynchronize(f->c->stream);
    if (f->buf) cudaFree(f->buf);
    delete f;
}

int snk_frame_dev_set_pose(snk_frame_dev* f, const double* q_xyzw, const double* t)
{
    SNK_REQUIRE(f && q_xyzw && t, "null argument");
    for (int i = 0; i < 4; ++i) f->f.q[i] = q_xyzw[i];
    for (int i = 0; i < 3; ++i) f->f.t[i] = t[i];
    return SNK_OK;
}

int snk_frame_dev_set_assignment(snk_frame_dev* f, const int32_t* kp_to_point)
{
    SNK_REQUIRE(f && (f->f.n == 0 || kp_to_point), "null argument");
    if (f->f.n == 0) return SNK_OK;
    SNK_CUDA(cudaSetDevice(f->c->device));
    SNK_CUDA(cudaMemcpyAsync(f->k2p, kp_to_point, 4 * (size_t)f->f.n, cudaMemcpyHostToDevice, f->c->stream));
    SNK_CUDA(cudaStreamSynchronize(f->c->stream));
    return SNK_OK;
}

int snk_localmap_create(snk_ctx* c, int capacity, snk_localmap** out)
{
    SNK_REQUIRE(c && out && capacity >= 1, "bad argument");
    SNK_CUDA(cudaSetDevice(c->device));
    auto* lm     = new snk_localmap;
    lm->c        = c;
    lm->capacity = capacity;
    if (cudaMalloc(&lm->d, sizeof(snk_point_fine) * (size_t)capacity) != cudaSuccess || cudaMalloc(&lm->bi, 4 * (size_t)capacity) != cudaSuccess ||
        cudaMalloc(&lm->bd, 4 * (size_t)capacity) != cudaSuccess || cudaMalloc(&lm->vis, (size_t)capacity) != cudaSuccess)
    {
        set_error("snk_localmap_create: cudaMalloc failed for %d points", capacity);
        cudaFree(lm->d), cudaFree(lm->bi), cudaFree(lm->bd), cudaFree(lm->vis);
        delete lm;
        return SNK_ERR_CUDA;
    }
    *out = lm;
    return SNK_OK;
}

void snk_localmap_destroy(snk_localmap* lm)
{
    if (!lm) return;
    cudaSetDevice(lm->c->device);
    cudaStreamSynchronize(lm->c->stream);
    cudaFree(lm->d), cudaFree(lm->bi), cudaFree(lm->bd), cudaFree(lm->vis);
    delete lm;
}

int snk_localmap_upload(snk_localmap* lm, const snk_point_fine* points, int first, int n, int new_size)
{
    SNK_REQUIRE(lm && first >= 0 && n >= 0 && (n == 0 || points), "bad argument");
    SNK_REQUIRE(first + n <= lm->capacity && new_size <= lm->capacity && new_size >= 0, "local map capacity %d exceeded", lm->capacity);
    for (int i = 0; i < n; ++i) SNK_REQUIRE(points[i].ref_level >= 0 && points[i].ref_level < 16, "ref_level out of range");
    SNK_CUDA(cudaSetDevice(lm->c->device));
    if (n)
    {
        SNK_CUDA(cudaMemcpyAsync(lm->d + first, points, sizeof(snk_point_fine) * (size_t)n, cudaMemcpyHostToDevice, lm->c->stream));
        SNK_CUDA(cudaStreamSynchronize(lm->c->stream));
    }
    lm->n = new_size;
    return SNK_OK;
}

int snk_search_projection_fine_resident(snk_ctx* c, snk_frame_dev* fd, snk_localmap* lm, const snk_camera* cam,
                                        const snk_scale_pyramid* pyr, float th, float ratio, int keep_assignment,
                                        int32_t* kp_to_point_out, int32_t* point_valid_out, uint8_t* visible_out, int* n_matches)
{
    SNK_REQUIRE(c && fd && lm && cam && pyr && n_matches, "null argument");
    SNK_REQUIRE(fd->c == c && lm->c == c, "frame and local map must belong to this context");
    SNK_REQUIRE(pyr->nlevels >= 1 && pyr->nlevels <= 16, "bad pyramid");
    SNK_CUDA(cudaSetDevice(c->device));
    int rc = ctx_reserve(c, padded<int>(std::max(lm->n, 1)) + padded<int>(4));
    if (rc) return rc;
    Carver cv{c->d_buf};
    int* dvalid = cv.take<int>(std::max(lm->n, 1));
    int* dn     = cv.take<int>(4);
    cudaStream_t st = c->stream;
    const int nk = fd->f.n, np = lm->n;
    SNK_CUDA(cudaMemsetAsync(dn, 0, 16, st));
    if (nk) k_fill_int<<<(nk + 255) / 256, 256, 0, st>>>(fd->winner, nk, 0x7fffffff);
    if (np)
        k_proj_fine<<<(np + 127) / 128, 128, 0, st>>>(fd->f, lm->d, np, *cam, *pyr, th, ratio, std::log(pyr->scale_factor), lm->bi, lm->bd,
                                                     lm->vis, fd->winner);
    if (nk)
        k_assign<snk_point_fine><<<(nk + 255) / 256, 256, 0, st>>>(fd->k2p, fd->winner, lm->d, nullptr, nk, fd->k2p_out, nullptr, dn);
    if (np && point_valid_out)
    {
        k_gather_valid<<<(np + 255) / 256, 256, 0, st>>>(lm->d, np, dvalid);
        SNK_CUDA(cudaMemcpyAsync(point_valid_out, dvalid, 4 * (size_t)np, cudaMemcpyDeviceToHost, st));
    }
    if (np && visible_out) SNK_CUDA(cudaMemcpyAsync(visible_out, lm->vis, (size_t)np, cudaMemcpyDeviceToHost, st));
    if (nk && kp_to_point_out) SNK_CUDA(cudaMemcpyAsync(kp_to_point_out, fd->k2p_out, 4 * (size_t)nk, cudaMemcpyDeviceToHost, st));
    if (nk && keep_assignment) SNK_CUDA(cudaMemcpyAsync(fd->k2p, fd->k2p_out, 4 * (size_t)nk, cudaMemcpyDeviceToDevice, st));
    SNK_CUDA(cudaMemcpyAsync(c->h_pin + 4, dn, 4, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaStreamSynchronize(st));
    *n_matches = c->h_pin[4];
    return SNK_OK;
}

}  // extern "C"
