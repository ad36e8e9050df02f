// Host side of the vocabulary handle and of snk_bow_transform (included at the end of matchers.cu).
struct snk_voc
{
    int device = 0;
    int k = 0, L = 0, n_nodes = 0, n_words = 0;
    int *d_child_begin = nullptr, *d_child_id = nullptr, *d_word_id = nullptr;
    uint4* d_child_desc = nullptr;
    float* d_weight     = nullptr;
    VocDev dev() const { return VocDev{k, L, n_nodes, n_words, d_child_begin, d_child_id, d_child_desc, d_weight, d_word_id}; }
};

namespace
{
int voc_build(int device, int k, int L, int n, const int32_t* parent, const double* weight, const int32_t* word_id,
              const uint64_t* desc, snk_voc** out)
{
    SNK_REQUIRE(out && parent && weight && word_id && desc, "null argument");
    SNK_REQUIRE(n >= 2 && k >= 1 && L >= 1, "a vocabulary needs a root and at least one word");
    int rc = select_device(device);
    if (rc) return rc;
    // loadRaw (MiniBow2.h:1033-1041): node ids are file positions; every node but the root is appended to its
    // parent's child list in file order
    std::vector<int> begin(n + 1, 0), child_id(n - 1);
    for (int i = 1; i < n; ++i)
    {
        SNK_REQUIRE(parent[i] >= 0 && parent[i] < n && parent[i] != i, "node %d has parent %d", i, parent[i]);
        ++begin[parent[i] + 1];
    }
    for (int i = 0; i < n; ++i) begin[i + 1] += begin[i];
    SNK_REQUIRE(begin[1] > 0, "the root has no children");
    std::vector<int> fill(begin.begin(), begin.end() - 1);
    std::vector<uint64_t> cdesc(4 * (size_t)(n - 1));
    std::vector<float> w(n);
    int n_words = 0;
    for (int i = 1; i < n; ++i)
    {
        const int slot = fill[parent[i]]++;
        child_id[slot] = i;
        memcpy(&cdesc[4 * (size_t)slot], desc + 4 * (size_t)i, 32);
    }
    for (int i = 0; i < n; ++i)
    {
        w[i] = (float)weight[i];  // `n.weight = weight` narrows the stored double (:1039)
        if (i > 0 && begin[i + 1] == begin[i])
        {
            SNK_REQUIRE(word_id[i] >= 0, "leaf %d has word id %d", i, word_id[i]);
            n_words = std::max(n_words, word_id[i] + 1);
        }
    }
    // a descent must end: every non-root chain leads to a leaf because parents precede... not guaranteed by the
    // format, so check that no node is its own ancestor (depth bounded by n)
    {
        std::vector<int> depth(n, -1);
        depth[0] = 0;
        for (int i = 1; i < n; ++i)
        {
            int steps = 0, j = i;
            while (depth[j] < 0 && steps <= n) j = parent[j], ++steps;
            SNK_REQUIRE(depth[j] >= 0, "parent links of node %d form a cycle", i);
            int d = depth[j] + steps;
            for (j = i; depth[j] < 0; j = parent[j]) depth[j] = d--;
        }
    }
    snk_voc* v = new snk_voc;
    v->device = device, v->k = k, v->L = L, v->n_nodes = n, v->n_words = n_words;
    bool ok = cudaMalloc(&v->d_child_begin, 4 * (size_t)(n + 1)) == cudaSuccess &&
              cudaMalloc(&v->d_child_id, 4 * (size_t)(n - 1)) == cudaSuccess &&
              cudaMalloc(&v->d_child_desc, 32 * (size_t)(n - 1)) == cudaSuccess &&
              cudaMalloc(&v->d_weight, 4 * (size_t)n) == cudaSuccess && cudaMalloc(&v->d_word_id, 4 * (size_t)n) == cudaSuccess;
    ok = ok && cudaMemcpy(v->d_child_begin, begin.data(), 4 * (size_t)(n + 1), cudaMemcpyHostToDevice) == cudaSuccess &&
         cudaMemcpy(v->d_child_id, child_id.data(), 4 * (size_t)(n - 1), cudaMemcpyHostToDevice) == cudaSuccess &&
         cudaMemcpy(v->d_child_desc, cdesc.data(), 32 * (size_t)(n - 1), cudaMemcpyHostToDevice) == cudaSuccess &&
         cudaMemcpy(v->d_weight, w.data(), 4 * (size_t)n, cudaMemcpyHostToDevice) == cudaSuccess &&
         cudaMemcpy(v->d_word_id, word_id, 4 * (size_t)n, cudaMemcpyHostToDevice) == cudaSuccess;
    if (!ok)
    {
        set_error("vocabulary upload failed: %s", cudaGetErrorString(cudaGetLastError()));
        snk_voc_destroy(v);
        return SNK_ERR_CUDA;
    }
    *out = v;
    return SNK_OK;
}
}  // namespace

int snk_voc_create(int device, int k, int L, int n_nodes, const int32_t* parent, const double* weight, const int32_t* word_id,
                   const uint64_t* desc, snk_voc** out)
{
    return voc_build(device, k, L, n_nodes, parent, weight, word_id, desc, out);
}

int snk_voc_load_raw(int device, const char* path, snk_voc** out)
{
    SNK_REQUIRE(path && out, "null argument");
    FILE* f = fopen(path, "rb");
    SNK_REQUIRE(f, "Could not load Voc file %s", path);  // loadRaw throws here (MiniBow2.h:1024-1027)
    int32_t hdr[4];
    uint64_t count = 0;
    bool ok = fread(hdr, 4, 4, f) == 4 && fread(&count, 8, 1, f) == 1 && count >= 2 && count < (1ull << 30);
    std::vector<int32_t> parent, word;
    std::vector<double> weight;
    std::vector<uint64_t> desc;
    if (ok)
    {
        parent.resize(count), word.resize(count), weight.resize(count), desc.resize(4 * count);
        std::vector<unsigned char> rec(52 * count);  // id i32, parent i32, weight f64, word_id i32, descriptor 32 B
        ok = fread(rec.data(), 52, count, f) == count;
        for (uint64_t i = 0; ok && i < count; ++i)
        {
            const unsigned char* r = &rec[52 * i];
            int32_t id;
            memcpy(&id, r, 4), memcpy(&parent[i], r + 4, 4), memcpy(&weight[i], r + 8, 8), memcpy(&word[i], r + 16, 4);
            memcpy(&desc[4 * i], r + 20, 32);
            ok = id == (int32_t)i;  // the reference indexes m_nodes by id (:1040)
        }
    }
    fclose(f);
    SNK_REQUIRE(ok, "%s is not a MiniBow2 raw vocabulary", path);
    return voc_build(device, hdr[0], hdr[1], (int)count, parent.data(), weight.data(), word.data(), desc.data(), out);
}

void snk_voc_destroy(snk_voc* v)
{
    if (!v) return;
    cudaSetDevice(v->device);
    if (v->d_child_begin) cudaFree(v->d_child_begin);
    if (v->d_child_id) cudaFree(v->d_child_id);
    if (v->d_child_desc) cudaFree(v->d_child_desc);
    if (v->d_weight) cudaFree(v->d_weight);
    if (v->d_word_id) cudaFree(v->d_word_id);
    delete v;
}

int snk_voc_info(const snk_voc* v, int* k, int* L, int* n_nodes, int* n_words)
{
    SNK_REQUIRE(v, "null vocabulary");
    if (k) *k = v->k;
    if (L) *L = v->L;
    if (n_nodes) *n_nodes = v->n_nodes;
    if (n_words) *n_words = v->n_words;
    return SNK_OK;
}

namespace
{
int bow_smem_attr(int device)
{
    static std::mutex attr_mutex;
    static bool attr_set[64] = {};
    std::lock_guard<std::mutex> lock(attr_mutex);
    if (device < 64 && !attr_set[device])
    {
        SNK_CUDA(cudaFuncSetAttribute(k_bow_assemble, cudaFuncAttributeMaxDynamicSharedMemorySize, BOW_MAX_FEATURES * 16));
        attr_set[device] = true;
    }
    return SNK_OK;
}
int pow2_at_least(int n)
{
    int p = 64;  // matches k_bow_assemble
    while (p < n) p <<= 1;
    return p;
}
}  // namespace

int snk_bow_transform(snk_ctx* c, const snk_voc* voc, const uint64_t* desc, int n, int levelsup, int32_t* bow_ids,
                      float* bow_vals, int* n_bow, int32_t* fv_node_ids, int32_t* fv_offsets, int32_t* fv_indices,
                      int* n_fv_nodes)
{
    SNK_REQUIRE(c && voc && n_bow && n_fv_nodes && fv_offsets, "null argument");
    SNK_REQUIRE(n >= 0 && n <= BOW_MAX_FEATURES, "n = %d outside [0, %d]", n, BOW_MAX_FEATURES);
    SNK_REQUIRE(voc->device == c->device, "vocabulary lives on device %d, context on %d", voc->device, c->device);
    *n_bow = *n_fv_nodes = 0;
    fv_offsets[0]        = 0;
    if (n == 0) return SNK_OK;
    SNK_REQUIRE(desc && bow_ids && bow_vals && fv_node_ids && fv_indices, "null output arrays");
    SNK_CUDA(cudaSetDevice(c->device));
    int rc = ctx_reserve(c, padded<uint64_t>(4 * (size_t)n) + 2 * padded<unsigned long long>(n) + 4 * padded<int>(n + 1) +
                                padded<float>(n) + padded<int>(4));
    if (rc) return rc;
    Carver cv{c->d_buf};
    auto* dd   = cv.take<uint64_t>(4 * (size_t)n);
    auto* dw   = cv.take<unsigned long long>(n);
    auto* df   = cv.take<unsigned long long>(n);
    int* dbi   = cv.take<int>(n + 1);
    float* dbv = cv.take<float>(n);
    int* dfn   = cv.take<int>(n + 1);
    int* dfo   = cv.take<int>(n + 1);
    int* dfi   = cv.take<int>(n + 1);
    int* dcnt  = cv.take<int>(4);
    cudaStream_t st = c->stream;
    if ((rc = bow_smem_attr(c->device))) return rc;
    SNK_CUDA(cudaMemcpyAsync(dd, desc, 32 * (size_t)n, cudaMemcpyHostToDevice, st));
    k_bow_descend<<<(n + 15) / 16, 256, 0, st>>>(voc->dev(), reinterpret_cast<const uint4*>(dd), nullptr, n, n, levelsup, dw, df);
    k_bow_assemble<<<1, BOW_NT, (size_t)pow2_at_least(n) * 16, st>>>(dw, df, nullptr, n, n, dbi, dbv, dfn, dfo, dfi, dcnt);
    SNK_CUDA(cudaGetLastError());
    SNK_CUDA(cudaMemcpyAsync(c->h_pin + 8, dcnt, 8, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(bow_ids, dbi, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(bow_vals, dbv, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(fv_node_ids, dfn, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(fv_offsets, dfo, 4 * (size_t)(n + 1), cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(fv_indices, dfi, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaStreamSynchronize(st));
    *n_bow      = c->h_pin[8];
    *n_fv_nodes = c->h_pin[9];
    return SNK_OK;
}

int snk_bow_transform_batch_device(snk_ctx* c, const snk_voc* voc, const uint64_t* desc_dev, const int32_t* counts_dev,
                                   int n_frames, int stride, int levelsup, int32_t* bow_ids_dev, float* bow_vals_dev,
                                   int32_t* fv_node_ids_dev, int32_t* fv_offsets_dev, int32_t* fv_indices_dev,
                                   int32_t* out_counts_dev)
{
    SNK_REQUIRE(c && voc && desc_dev && counts_dev && bow_ids_dev && bow_vals_dev && fv_node_ids_dev && fv_offsets_dev &&
                    fv_indices_dev && out_counts_dev,
                "null argument");
    SNK_REQUIRE(n_frames >= 0 && stride >= 1 && stride <= BOW_MAX_FEATURES, "stride = %d outside [1, %d]", stride, BOW_MAX_FEATURES);
    SNK_REQUIRE(voc->device == c->device, "vocabulary lives on device %d, context on %d", voc->device, c->device);
    if (n_frames == 0) return SNK_OK;
    SNK_CUDA(cudaSetDevice(c->device));
    int rc = ctx_reserve(c, 2 * padded<unsigned long long>((size_t)n_frames * stride));
    if (rc) return rc;
    if ((rc = bow_smem_attr(c->device))) return rc;
    Carver cv{c->d_buf};
    auto* dw = cv.take<unsigned long long>((size_t)n_frames * stride);
    auto* df = cv.take<unsigned long long>((size_t)n_frames * stride);
    cudaStream_t st = c->stream;
    k_bow_descend<<<dim3((stride + 15) / 16, n_frames), 256, 0, st>>>(voc->dev(), reinterpret_cast<const uint4*>(desc_dev),
                                                                      counts_dev, 0, stride, levelsup, dw, df);
    k_bow_assemble<<<n_frames, BOW_NT, (size_t)pow2_at_least(stride) * 16, st>>>(
        dw, df, counts_dev, 0, stride, bow_ids_dev, bow_vals_dev, fv_node_ids_dev, fv_offsets_dev, fv_indices_dev, out_counts_dev);
    SNK_CUDA(cudaGetLastError());
    return SNK_OK;
}

int snk_bow_score_batch(snk_ctx* c, const int32_t* q_ids, const float* q_vals, int nq, const int32_t* cand_ids,
                        const float* cand_vals, const int32_t* cand_offsets, int n_cand, float* scores)
{
    SNK_REQUIRE(c && nq >= 0 && n_cand >= 0, "bad argument");
    if (n_cand == 0) return SNK_OK;
    SNK_REQUIRE(cand_offsets && scores && (nq == 0 || (q_ids && q_vals)), "null arrays");
    const int total = cand_offsets[n_cand];
    SNK_REQUIRE(cand_offsets[0] == 0 && total >= 0 && (total == 0 || (cand_ids && cand_vals)), "bad candidate arrays");
    SNK_CUDA(cudaSetDevice(c->device));
    int rc = ctx_reserve(c, 2 * padded<int>(std::max(nq, 1)) + 2 * padded<int>(std::max(total, 1)) + padded<int>(n_cand + 1) +
                                padded<float>(n_cand));
    if (rc) return rc;
    Carver cv{c->d_buf};
    int* dqi   = cv.take<int>(std::max(nq, 1));
    float* dqv = cv.take<float>(std::max(nq, 1));
    int* dci   = cv.take<int>(std::max(total, 1));
    float* dcv = cv.take<float>(std::max(total, 1));
    int* dco   = cv.take<int>(n_cand + 1);
    float* ds  = cv.take<float>(n_cand);
    cudaStream_t st = c->stream;
    if (nq)
    {
        SNK_CUDA(cudaMemcpyAsync(dqi, q_ids, 4 * (size_t)nq, cudaMemcpyHostToDevice, st));
        SNK_CUDA(cudaMemcpyAsync(dqv, q_vals, 4 * (size_t)nq, cudaMemcpyHostToDevice, st));
    }
    if (total)
    {
        SNK_CUDA(cudaMemcpyAsync(dci, cand_ids, 4 * (size_t)total, cudaMemcpyHostToDevice, st));
        SNK_CUDA(cudaMemcpyAsync(dcv, cand_vals, 4 * (size_t)total, cudaMemcpyHostToDevice, st));
    }
    SNK_CUDA(cudaMemcpyAsync(dco, cand_offsets, 4 * (size_t)(n_cand + 1), cudaMemcpyHostToDevice, st));
    k_bow_score<<<(n_cand + 127) / 128, 128, 0, st>>>(dqi, dqv, nq, dci, dcv, dco, n_cand, ds);
    SNK_CUDA(cudaGetLastError());
    SNK_CUDA(cudaMemcpyAsync(scores, ds, 4 * (size_t)n_cand, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaStreamSynchronize(st));
    return SNK_OK;
}
