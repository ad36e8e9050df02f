// Host side of the matching context (snk_ctx) and the matcher entry points of include/snakeorb.h.
#include "match_kernels.cuh"
#include "project_kernels.cuh"
#include "image_kernels.cuh"
#include "bow_kernels.cuh"
#include "mapping_kernels.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>

using namespace snk;

struct snk_ctx
{
    int device              = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    // growable device scratch, carved sequentially per call
    unsigned char* d_buf = nullptr;
    size_t d_cap         = 0;
    int* d_err           = nullptr;
    int* h_pin           = nullptr;  // pinned, 16 ints
    unsigned char* h_stage = nullptr;  // pinned, growable: host side of batched uploads
    size_t h_stage_cap     = 0;
};

namespace
{
int ctx_reserve(snk_ctx* c, size_t bytes)
{
    if (bytes <= c->d_cap) return SNK_OK;
    SNK_CUDA(cudaStreamSynchronize(c->stream));
    if (c->d_buf) SNK_CUDA(cudaFree(c->d_buf));
    c->d_buf = nullptr;
    c->d_cap = 0;
    size_t cap = std::max(bytes + bytes / 4, (size_t)1 << 20);
    SNK_CUDA(cudaMalloc(&c->d_buf, cap));
    c->d_cap = cap;
    return SNK_OK;
}

struct Carver
{
    unsigned char* base;
    size_t off = 0;
    template <typename T>
    T* take(size_t n)
    {
        off     = align_up(off, (size_t)256);
        T* p    = reinterpret_cast<T*>(base + off);
        off += n * sizeof(T);
        return p;
    }
};
template <typename T>
size_t padded(size_t n)
{
    return align_up(n * sizeof(T), (size_t)256) + 256;
}

int check_err(snk_ctx* c, const char* what)
{
    SNK_CUDA(cudaMemcpyAsync(c->h_pin, c->d_err, 4, cudaMemcpyDeviceToHost, c->stream));
    SNK_CUDA(cudaStreamSynchronize(c->stream));
    if (c->h_pin[0])
    {
        set_error("%s: capacity error %d", what, c->h_pin[0]);
        cudaMemsetAsync(c->d_err, 0, 4, c->stream);
        return SNK_ERR_CAPACITY;
    }
    return SNK_OK;
}
}  // namespace

extern "C" {

int snk_ctx_create(int device, snk_ctx** out)
{
    SNK_REQUIRE(out, "null argument");
    int rc = select_device(device);
    if (rc) return rc;
    snk_ctx* c = new snk_ctx;
    c->device  = device;
    if (cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaMalloc(&c->d_err, 4) != cudaSuccess || cudaMemset(c->d_err, 0, 4) != cudaSuccess ||
        cudaHostAlloc(&c->h_pin, 64, cudaHostAllocDefault) != cudaSuccess)
    {
        set_error("context setup failed: %s", cudaGetErrorString(cudaGetLastError()));
        delete c;
        return SNK_ERR_CUDA;
    }
    c->stream = c->own_stream;
    *out      = c;
    return SNK_OK;
}

void snk_ctx_destroy(snk_ctx* c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    if (c->d_buf) cudaFree(c->d_buf);
    if (c->d_err) cudaFree(c->d_err);
    if (c->h_pin) cudaFreeHost(c->h_pin);
    if (c->h_stage) cudaFreeHost(c->h_stage);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
}

int snk_ctx_set_stream(snk_ctx* c, void* s)
{
    SNK_REQUIRE(c, "null context");
    c->stream = s ? (cudaStream_t)s : c->own_stream;
    return SNK_OK;
}

int snk_hamming_pairs(snk_ctx* c, const uint64_t* a, const uint64_t* b, int n, int32_t* dist)
{
    SNK_REQUIRE(c && a && b && dist && n >= 0, "bad argument");
    if (n == 0) return SNK_OK;
    SNK_CUDA(cudaSetDevice(c->device));
    int rc = ctx_reserve(c, 2 * padded<uint64_t>(4 * (size_t)n) + padded<int>(n));
    if (rc) return rc;
    Carver cv{c->d_buf};
    uint64_t* da = cv.take<uint64_t>(4 * (size_t)n);
    uint64_t* db = cv.take<uint64_t>(4 * (size_t)n);
    int* dd      = cv.take<int>(n);
    SNK_CUDA(cudaMemcpyAsync(da, a, 32 * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    SNK_CUDA(cudaMemcpyAsync(db, b, 32 * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    k_hamming_pairs<<<(n + 255) / 256, 256, 0, c->stream>>>(da, db, n, dd);
    SNK_CUDA(cudaMemcpyAsync(dist, dd, 4 * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
    SNK_CUDA(cudaStreamSynchronize(c->stream));
    return SNK_OK;
}

int snk_match_knn2_batch_device(snk_ctx* c, const uint64_t* d1, const int32_t* n_dev, int stride1, const uint64_t* d2,
                                const int32_t* m_dev, int stride2, int n_problems, snk_knn2* out)
{
    SNK_REQUIRE(c && d1 && d2 && out && n_dev && m_dev && n_problems >= 0, "bad argument");
    if (n_problems == 0 || stride1 == 0) return SNK_OK;
    SNK_CUDA(cudaSetDevice(c->device));
    k_knn2<<<dim3((stride1 + KNN_NT - 1) / KNN_NT, n_problems), KNN_NT, 0, c->stream>>>(d1, n_dev, 0, stride1, d2, m_dev,
                                                                                        0, stride2, out);
    SNK_CUDA(cudaGetLastError());
    return SNK_OK;
}

// How one n x m problem is spread over the machine: with at least two full waves of query blocks the thread-per-query
// kernel is used; below that the train set is split over warps (k_knn2_split + k_knn2_merge). qgroup = queries per block.
static bool knn2_split_plan(int device, int n, int m, int* qgroup, int* n_chunks)
{
    static int sms[64] = {};
    if (device < 64 && !sms[device]) cudaDeviceGetAttribute(&sms[device], cudaDevAttrMultiProcessorCount, device);
    const int nsm = (device < 64 && sms[device] > 0) ? sms[device] : 148;
    if (m <= 0 || m > KNS_MAX_M || (n + KNN_NT - 1) / KNN_NT >= 2 * nsm) return false;
    const int tblocks = (m + 32 * KNS_WARPS - 1) / (32 * KNS_WARPS);
    // about four blocks per SM, query groups in multiples of 32 (one coalesced partial store per 32 queries)
    int g = (int)(((long long)n * tblocks + 4LL * nsm - 1) / (4LL * nsm));
    g = std::min(256, std::max(32, (g + 31) / 32 * 32));
    *qgroup   = g;
    *n_chunks = (m + 31) / 32;
    return (size_t)*n_chunks * n * sizeof(uint2) <= ((size_t)64 << 20);
}

int snk_match_knn2(snk_ctx* c, const uint64_t* d1, int n, const uint64_t* d2, int m, snk_knn2* out)
{
    SNK_REQUIRE(c && out && n >= 0 && m >= 0 && (n == 0 || d1) && (m == 0 || d2), "bad argument");
    if (n == 0) return SNK_OK;
    SNK_CUDA(cudaSetDevice(c->device));
    int qgroup = 0, n_chunks = 0;
    const bool split = knn2_split_plan(c->device, n, m, &qgroup, &n_chunks);
    int rc = ctx_reserve(c, padded<uint64_t>(4 * (size_t)n) + padded<uint64_t>(4 * (size_t)std::max(m, 1)) + padded<snk_knn2>(n) +
                                (split ? padded<uint2>((size_t)n_chunks * n) : 0));
    if (rc) return rc;
    Carver cv{c->d_buf};
    uint64_t* a = cv.take<uint64_t>(4 * (size_t)n);
    uint64_t* b = cv.take<uint64_t>(4 * (size_t)std::max(m, 1));
    snk_knn2* o = cv.take<snk_knn2>(n);
    SNK_CUDA(cudaMemcpyAsync(a, d1, 32 * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    if (m) SNK_CUDA(cudaMemcpyAsync(b, d2, 32 * (size_t)m, cudaMemcpyHostToDevice, c->stream));
    if (split)
    {
        uint2* part = cv.take<uint2>((size_t)n_chunks * n);
        k_knn2_split<<<dim3((n_chunks + KNS_WARPS - 1) / KNS_WARPS, (n + qgroup - 1) / qgroup), KNS_WARPS * 32, (size_t)qgroup * 32,
                       c->stream>>>(a, n, b, m, qgroup, part);
        k_knn2_merge<<<(n + 127) / 128, 128, 0, c->stream>>>(part, n, n_chunks, o);
    }
    else
        k_knn2<<<dim3((n + KNN_NT - 1) / KNN_NT, 1), KNN_NT, 0, c->stream>>>(a, nullptr, n, n, b, nullptr, m, m, o);
    SNK_CUDA(cudaGetLastError());
    SNK_CUDA(cudaMemcpyAsync(out, o, sizeof(snk_knn2) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
    SNK_CUDA(cudaStreamSynchronize(c->stream));
    return SNK_OK;
}

// BruteForceMatcher::filterMatches (Features.h:394-414): a serial pass over n small records on the caller's thread —
// it is host bookkeeping in the reference too (the result is a std::vector built in index order).
int snk_filter_matches(const snk_knn2* k, int n, int threshold, float ratio, int32_t* ij, int* n_matches)
{
    SNK_REQUIRE(k && ij && n_matches && n >= 0, "bad argument");
    int cnt = 0;
    for (int i = 0; i < n; ++i)
    {
        if (k[i].dist0 > threshold) continue;
        if (float(k[i].dist0) > float(k[i].dist1) * ratio) continue;
        ij[2 * cnt]     = i;
        ij[2 * cnt + 1] = k[i].idx0;
        ++cnt;
    }
    *n_matches = cnt;
    return SNK_OK;
}

static int stereo_smem(int sort_cap, int nbins = ST_ROW_BINS) { return sort_cap * 4 + (nbins + 4) * 4; }
constexpr int ST_SMEM_CEILING = 200 * 1024;
// with the right keypoints' x (xt_bytes each) and octaves staged behind the counters; 0 if that does not fit
static int stereo_smem_staged(int sort_cap, int nbins, int xt_bytes)
{
    const size_t b = (((size_t)(sort_cap + nbins + 4) * 4 + 7) & ~(size_t)7) + (size_t)sort_cap * (xt_bytes + 4);
    return b <= (size_t)ST_SMEM_CEILING ? (int)b : 0;
}
// rows the counting sort of k_stereo has to cover for keypoints of an image with `rows` rows: a multiple of ST_NT
static int stereo_bins(int rows) { return std::min(ST_ROW_BINS, std::max(ST_NT, (rows + ST_NT) / ST_NT * ST_NT)); }

int snk_stereo_match(snk_ctx* c, const snk_stereo_params* p, const snk_keypoint_f64* left, const uint64_t* dl, int nl,
                     const snk_keypoint_f64* right, const uint64_t* dr, int nr, float* right_points, float* depth,
                     int* n_matches)
{
    SNK_REQUIRE(c && p && n_matches && nl >= 0 && nr >= 0, "bad argument");
    SNK_REQUIRE(nl == 0 || (left && dl && right_points && depth), "null left arrays");
    SNK_REQUIRE(nr == 0 || (right && dr), "null right arrays");
    SNK_REQUIRE(p->nlevels >= 1 && p->nlevels <= 16, "bad nlevels");
    *n_matches = 0;
    if (nl == 0) return SNK_OK;
    if (nr == 0)
    {
        for (int i = 0; i < nl; ++i) right_points[i] = depth[i] = -1000.f;
        return SNK_OK;
    }
    for (int i = 0; i < nl; ++i) SNK_REQUIRE(left[i].octave >= 0 && left[i].octave < p->nlevels, "octave out of range");
    SNK_REQUIRE(nr <= 16384, "more than 16384 right keypoints");
    SNK_CUDA(cudaSetDevice(c->device));
    const int stride = std::max(nl, nr);
    int rc = ctx_reserve(c, padded<snk_keypoint_f64>(2 * (size_t)stride) + padded<uint64_t>(8 * (size_t)stride) +
                                2 * padded<float>(stride) + padded<int>(4));
    if (rc) return rc;
    Carver cv{c->d_buf};
    snk_keypoint_f64* dk = cv.take<snk_keypoint_f64>(2 * (size_t)stride);
    uint64_t* dd         = cv.take<uint64_t>(8 * (size_t)stride);
    float* drp           = cv.take<float>(stride);
    float* ddp           = cv.take<float>(stride);
    int* dn              = cv.take<int>(4);
    cudaStream_t st      = c->stream;
    SNK_CUDA(cudaMemcpyAsync(dk, left, sizeof(snk_keypoint_f64) * (size_t)nl, cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaMemcpyAsync(dk + stride, right, sizeof(snk_keypoint_f64) * (size_t)nr, cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaMemcpyAsync(dd, dl, 32 * (size_t)nl, cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaMemcpyAsync(dd + 4 * (size_t)stride, dr, 32 * (size_t)nr, cudaMemcpyHostToDevice, st));
    int cap = 1;
    while (cap < nr) cap <<= 1;
    SNK_CUDA(cudaFuncSetAttribute(k_stereo<snk_keypoint_f64>, cudaFuncAttributeMaxDynamicSharedMemorySize, ST_SMEM_CEILING));
    SNK_CUDA(cudaMemsetAsync(dn, 0, 4, st));
    double ymax = 0;  // the right keypoints are host data here: the sort only has to cover the rows they use
    for (int i = 0; i < nr; ++i) ymax = (right[i].y > ymax) ? right[i].y : ymax;
    const int nbins = stereo_bins(ymax < 4096.0 ? (int)ymax + 1 : ST_ROW_BINS);
    const int staged = stereo_smem_staged(cap, nbins, 8);
    k_stereo<snk_keypoint_f64><<<dim3(1, std::min(8, (nl + ST_NT - 1) / ST_NT)), ST_NT, staged ? staged : stereo_smem(cap, nbins), st>>>(
        dk, dd, nullptr, stride, nl, nr, *p, cap, drp, ddp, dn, c->d_err, nbins, staged);
    SNK_CUDA(cudaMemcpyAsync(right_points, drp, 4 * (size_t)nl, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(depth, ddp, 4 * (size_t)nl, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(c->h_pin + 4, dn, 4, cudaMemcpyDeviceToHost, st));
    rc = check_err(c, "stereo_match");
    if (rc) return rc;
    *n_matches = c->h_pin[4];
    return SNK_OK;
}

int snk_stereo_matching(snk_ctx* c, const snk_rectify* rect_left, const snk_rectify* rect_right, const snk_stereo_params* p,
                        const snk_keypoint_f64* left, const uint64_t* dl, int nl, const snk_keypoint_f64* right,
                        const uint64_t* dr, int nr, float* right_points, float* depth, int* n_matches)
{
    SNK_REQUIRE(c && rect_left && rect_right && p && n_matches && nl >= 0 && nr >= 0, "bad argument");
    SNK_REQUIRE(nl == 0 || (left && dl && right_points && depth), "null left arrays");
    SNK_REQUIRE(nr == 0 || (right && dr), "null right arrays");
    SNK_REQUIRE(p->nlevels >= 1 && p->nlevels <= 16, "bad nlevels");
    *n_matches = 0;
    if (nl == 0) return SNK_OK;
    if (nr == 0)
    {
        for (int i = 0; i < nl; ++i) right_points[i] = depth[i] = -1000.f;
        return SNK_OK;
    }
    for (int i = 0; i < nl; ++i) SNK_REQUIRE(left[i].octave >= 0 && left[i].octave < p->nlevels, "octave out of range");
    SNK_REQUIRE(nr <= 16384, "more than 16384 right keypoints");
    SNK_CUDA(cudaSetDevice(c->device));
    const int stride = std::max(nl, nr);
    int rc = ctx_reserve(c, 2 * padded<snk_keypoint_f64>(2 * (size_t)stride) + padded<uint64_t>(8 * (size_t)stride) +
                                2 * padded<float>(stride) + padded<double>(2 * (size_t)stride) + padded<int>(4));
    if (rc) return rc;
    Carver cv{c->d_buf};
    snk_keypoint_f64* raw = cv.take<snk_keypoint_f64>(2 * (size_t)stride);
    snk_keypoint_f64* dk  = cv.take<snk_keypoint_f64>(2 * (size_t)stride);
    uint64_t* dd          = cv.take<uint64_t>(8 * (size_t)stride);
    float* drp            = cv.take<float>(stride);
    float* ddp            = cv.take<float>(stride);
    double* dnorm         = cv.take<double>(2 * (size_t)stride);  // normalized points: produced by the kernel, not needed here
    int* dn               = cv.take<int>(4);
    cudaStream_t st       = c->stream;
    SNK_CUDA(cudaMemcpyAsync(raw, left, sizeof(snk_keypoint_f64) * (size_t)nl, cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaMemcpyAsync(raw + stride, right, sizeof(snk_keypoint_f64) * (size_t)nr, cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaMemcpyAsync(dd, dl, 32 * (size_t)nl, cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaMemcpyAsync(dd + 4 * (size_t)stride, dr, 32 * (size_t)nr, cudaMemcpyHostToDevice, st));
    // the two Rectification::Forward loops (Preprocess.cpp:141-151): same per-point math as undistortKeypoints
    k_undistort<<<(nl + 127) / 128, 128, 0, st>>>(*rect_left, raw, nl, dk, dnorm);
    k_undistort<<<(nr + 127) / 128, 128, 0, st>>>(*rect_right, raw + stride, nr, dk + stride, dnorm);
    int cap = 1;
    while (cap < nr) cap <<= 1;
    SNK_CUDA(cudaFuncSetAttribute(k_stereo<snk_keypoint_f64>, cudaFuncAttributeMaxDynamicSharedMemorySize, ST_SMEM_CEILING));
    SNK_CUDA(cudaMemsetAsync(dn, 0, 4, st));
    const int staged = stereo_smem_staged(cap, ST_ROW_BINS, 8);
    k_stereo<snk_keypoint_f64><<<dim3(1, std::min(8, (nl + ST_NT - 1) / ST_NT)), ST_NT, staged ? staged : stereo_smem(cap), st>>>(
        dk, dd, nullptr, stride, nl, nr, *p, cap, drp, ddp, dn, c->d_err, ST_ROW_BINS, staged);
    SNK_CUDA(cudaMemcpyAsync(right_points, drp, 4 * (size_t)nl, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(depth, ddp, 4 * (size_t)nl, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(c->h_pin + 4, dn, 4, cudaMemcpyDeviceToHost, st));
    rc = check_err(c, "stereo_matching");
    if (rc) return rc;
    *n_matches = c->h_pin[4];
    return SNK_OK;
}

int snk_stereo_match_orb_device(snk_ctx* c, snk_orb* h, int n_pairs, const snk_stereo_params* p, float* right_points,
                                float* depth, int32_t* n_matches)
{
    SNK_REQUIRE(c && h && p && right_points && depth && n_matches && n_pairs >= 0, "bad argument");
    if (n_pairs == 0) return SNK_OK;
    const snk_keypoint_f32* kps;
    const uint64_t* desc;
    const int32_t* counts;
    int stride;
    int rc = snk_orb_device_results(h, &kps, &desc, &counts, &stride);
    if (rc) return rc;
    SNK_REQUIRE(2 * n_pairs <= snk_orb_last_batch(h), "n_pairs %d needs %d images, the last enqueue held %d", n_pairs,
                2 * n_pairs, snk_orb_last_batch(h));
    SNK_CUDA(cudaSetDevice(c->device));
    int cap = 1;
    while (cap < stride) cap <<= 1;
    SNK_REQUIRE(cap <= 16384, "more than 16384 keypoints per image");
    {
        // the attribute belongs to the device; set once per device, under a lock (handles may be used from several threads)
        static std::mutex mtx;
        static bool attr_set[64] = {};
        std::lock_guard<std::mutex> lock(mtx);
        if (c->device >= 64 || !attr_set[c->device])
        {
            SNK_CUDA(cudaFuncSetAttribute(k_stereo<snk_keypoint_f32>, cudaFuncAttributeMaxDynamicSharedMemorySize, ST_SMEM_CEILING));
            if (c->device < 64) attr_set[c->device] = true;
        }
    }
    // the kernel runs on the context's stream: it must start after the extraction that fills kps / desc / counts ...
    rc = snk_orb_order_stream_after(h, c->stream);
    if (rc) return rc;
    // enough blocks to fill the machine: several blocks per pair when the batch is small
    const int parts = std::max(1, std::min((stride + ST_NT - 1) / ST_NT, 1024 / std::max(n_pairs, 1)));
    SNK_CUDA(cudaMemsetAsync(n_matches, 0, 4 * (size_t)n_pairs, c->stream));
    int rows0 = ST_ROW_BINS;  // keypoint rows are level-0 rows: the image height bounds them
    if (snk_orb_get_level(h, 0, 0, 0, nullptr, &rows0, nullptr, nullptr) != SNK_OK) rows0 = ST_ROW_BINS;
    const int nbins = stereo_bins(rows0 + 1);
    const int staged = stereo_smem_staged(cap, nbins, 4);
    k_stereo<snk_keypoint_f32><<<dim3(n_pairs, parts), ST_NT, staged ? staged : stereo_smem(cap, nbins), c->stream>>>(
        kps, desc, counts, stride, 0, 0, *p, cap, right_points, depth, n_matches, c->d_err, nbins, staged);
    SNK_CUDA(cudaGetLastError());
    // ... and the handle's next enqueue must not overwrite them before this kernel has read them
    return snk_orb_order_after_stream(h, c->stream);
}

int snk_ctx_wait_orb(snk_ctx* c, snk_orb* h)
{
    SNK_REQUIRE(c && h, "null argument");
    return snk_orb_order_stream_after(h, c->stream);
}

int snk_orb_wait_ctx(snk_orb* h, snk_ctx* c)
{
    SNK_REQUIRE(c && h, "null argument");
    return snk_orb_order_after_stream(h, c->stream);
}

int snk_ctx_check(snk_ctx* c)
{
    SNK_REQUIRE(c, "null context");
    SNK_CUDA(cudaSetDevice(c->device));
    return check_err(c, "matcher kernels");
}

int snk_stereo_from_rgbd(snk_ctx* c, const snk_rgbd_params* p, const snk_keypoint_f64* kp_un, int n,
                         const float* depth_img, int rows, int cols, size_t pitch, float* right_points, float* depth,
                         int* n_matches)
{
    SNK_REQUIRE(c && p && n_matches && n >= 0, "bad argument");
    SNK_REQUIRE(n == 0 || (kp_un && right_points && depth), "null keypoint arrays");
    SNK_REQUIRE(depth_img && rows > 0 && cols > 0 && pitch >= (size_t)cols * 4 && pitch % 4 == 0, "bad depth image");
    *n_matches = 0;
    if (n == 0) return SNK_OK;
    SNK_CUDA(cudaSetDevice(c->device));
    const size_t img_floats = (size_t)rows * cols;
    int rc = ctx_reserve(c, padded<snk_keypoint_f64>(n) + padded<float>(img_floats) + 2 * padded<float>(n) + padded<int>(4));
    if (rc) return rc;
    Carver cv{c->d_buf};
    auto* dk  = cv.take<snk_keypoint_f64>(n);
    float* di = cv.take<float>(img_floats);
    float* dr = cv.take<float>(n);
    float* dd = cv.take<float>(n);
    int* dn   = cv.take<int>(4);
    cudaStream_t st = c->stream;
    SNK_CUDA(cudaMemcpyAsync(dk, kp_un, sizeof(snk_keypoint_f64) * (size_t)n, cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaMemcpy2DAsync(di, (size_t)cols * 4, depth_img, pitch, (size_t)cols * 4, rows, cudaMemcpyHostToDevice, st));
    SNK_CUDA(cudaMemsetAsync(dn, 0, 16, st));
    k_stereo_rgbd<<<(n + 127) / 128, 128, 0, st>>>(*p, dk, n, di, rows, cols, cols, dr, dd, dn);
    SNK_CUDA(cudaMemcpyAsync(right_points, dr, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(depth, dd, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(c->h_pin + 4, dn, 4, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaStreamSynchronize(st));
    *n_matches = c->h_pin[4];
    return SNK_OK;
}

int snk_undistort_keypoints(snk_ctx* c, const snk_rectify* r, const snk_keypoint_f64* kps, int n, snk_keypoint_f64* kp_un,
                            double* normalized_points)
{
    SNK_REQUIRE(c && r && n >= 0, "bad argument");
    SNK_REQUIRE(n == 0 || (kps && kp_un && normalized_points), "null keypoint arrays");
    SNK_REQUIRE(r->fx != 0 && r->fy != 0, "K_src with zero focal length");
    if (n == 0) return SNK_OK;
    SNK_CUDA(cudaSetDevice(c->device));
    int rc = ctx_reserve(c, 2 * padded<snk_keypoint_f64>(n) + padded<double>(2 * (size_t)n));
    if (rc) return rc;
    Carver cv{c->d_buf};
    auto* dk   = cv.take<snk_keypoint_f64>(n);
    auto* du   = cv.take<snk_keypoint_f64>(n);
    double* dn = cv.take<double>(2 * (size_t)n);
    cudaStream_t st = c->stream;
    SNK_CUDA(cudaMemcpyAsync(dk, kps, sizeof(snk_keypoint_f64) * (size_t)n, cudaMemcpyHostToDevice, st));
    k_undistort<<<(n + 127) / 128, 128, 0, st>>>(*r, dk, n, du, dn);
    SNK_CUDA(cudaGetLastError());
    SNK_CUDA(cudaMemcpyAsync(kp_un, du, sizeof(snk_keypoint_f64) * (size_t)n, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaMemcpyAsync(normalized_points, dn, 16 * (size_t)n, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaStreamSynchronize(st));
    return SNK_OK;
}

int snk_rgba_to_gray(snk_ctx* c, const uint8_t* rgba, int rows, int cols, size_t pitch, uint8_t* gray, size_t gray_pitch)
{
    SNK_REQUIRE(c && rows >= 0 && cols >= 0, "bad argument");
    if (rows == 0 || cols == 0) return SNK_OK;
    SNK_REQUIRE(rgba && gray && pitch >= (size_t)cols * 4 && gray_pitch >= (size_t)cols, "bad image arguments");
    SNK_CUDA(cudaSetDevice(c->device));
    const size_t sp = ((size_t)cols * 4 + 15) & ~(size_t)15, dp = ((size_t)cols + 3) & ~(size_t)3;
    int rc = ctx_reserve(c, padded<uint8_t>(sp * rows) + padded<uint8_t>(dp * rows));
    if (rc) return rc;
    Carver cv{c->d_buf};
    uint8_t* ds = cv.take<uint8_t>(sp * rows);
    uint8_t* dd = cv.take<uint8_t>(dp * rows);
    cudaStream_t st = c->stream;
    SNK_CUDA(cudaMemcpy2DAsync(ds, sp, rgba, pitch, (size_t)cols * 4, rows, cudaMemcpyHostToDevice, st));
    k_rgba_to_gray<<<dim3((cols + 511) / 512, rows, 1), 128, 0, st>>>(ds, 0, sp, dd, 0, (int)dp, cols, rows);
    SNK_CUDA(cudaGetLastError());
    SNK_CUDA(cudaMemcpy2DAsync(gray, gray_pitch, dd, dp, cols, rows, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaStreamSynchronize(st));
    return SNK_OK;
}

int snk_grid_bounds(int w, int h, const snk_camera* K, const double* dist, int resolution, snk_grid_view* out)
{
    SNK_REQUIRE(K && dist && out && w > 0 && h > 0 && resolution >= 1, "bad argument");
    SNK_REQUIRE(K->fx != 0 && K->fy != 0, "zero focal length");
    // host arithmetic in the reference's order; this translation unit's host code is built without FMA contraction
    // targets (plain x86-64), so it rounds like the reference's -O2 build
    double bminx = 0, bminy = 0, bmaxx = 0, bmaxy = 0;
    bool first = true;
    auto visit = [&](double px, double py) {
        const double nx0 = (px - K->cx) / K->fx, ny0 = (py - K->cy) / K->fy;  // unproject2
        double nx, ny;
        undistort_gn_hd(dist, nx0, ny0, nx, ny);
        const double ux = K->fx * nx + K->s * ny + K->cx, uy = K->fy * ny + K->cy;  // normalizedToImage
        if (first) bminx = bmaxx = ux, bminy = bmaxy = uy, first = false;
        bminx = std::min(ux, bminx), bminy = std::min(uy, bminy);
        bmaxx = std::max(ux, bmaxx), bmaxy = std::max(uy, bmaxy);
    };
    for (int i = 0; i < resolution + 1; ++i)
    {
        const float alpha = float(i) / resolution * w;  // float, as FeatureGrid2.h:70
        visit(alpha, 0);
        visit(alpha, h - 1);
        const float beta = float(i) / resolution * h;
        visit(0, beta);
        visit(w - 1, beta);
    }
    bminx -= 0.1, bminy -= 0.1, bmaxx += 0.1, bmaxy += 0.1;
    out->bmin_x = bminx, out->bmin_y = bminy, out->bmax_x = bmaxx, out->bmax_y = bmaxy;
    out->cell_size_inv_x = out->cell_size_inv_y = 1.0 / 20.0;
    out->rows  = (int)std::ceil((bmaxy - bminy) / 20.0);
    out->cols  = (int)std::ceil((bmaxx - bminx) / 20.0);
    out->cells = nullptr;
    return SNK_OK;
}

int snk_hamming_peak_selftest(snk_ctx* c, double* pair_ops_per_s)
{
    SNK_REQUIRE(c && pair_ops_per_s, "bad argument");
    SNK_CUDA(cudaSetDevice(c->device));
    int rc = ctx_reserve(c, 256);
    if (rc) return rc;
    cudaDeviceProp prop;
    SNK_CUDA(cudaGetDeviceProperties(&prop, c->device));
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 20000;
    cudaEvent_t e0, e1;
    SNK_CUDA(cudaEventCreate(&e0));
    SNK_CUDA(cudaEventCreate(&e1));
    double best = 0;
    for (int rep = 0; rep < 4; ++rep)
    {
        SNK_CUDA(cudaEventRecord(e0, c->stream));
        k_popc_peak<<<blocks, threads, 0, c->stream>>>(iters, 12345u + rep, reinterpret_cast<int*>(c->d_buf));
        SNK_CUDA(cudaEventRecord(e1, c->stream));
        SNK_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        SNK_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        if (rep) best = std::max(best, (double)blocks * threads * iters / (ms * 1e-3));
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *pair_ops_per_s = best;
    return SNK_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------------------------
// Feature cache files (host only)
// ------------------------------------------------------------------------------------------------------------------
extern "C" int snk_features_write(const char* path, const snk_keypoint_f64* kps, const uint64_t* desc, int n)
{
    SNK_REQUIRE(path && n >= 0 && (n == 0 || (kps && desc)), "bad argument");
    FILE* f = std::fopen(path, "wb");
    SNK_REQUIRE(f, "cannot open %s for writing", path);
    const uint64_t cnt = (uint64_t)n;  // BinaryFile::write(vector): size_t count, then the raw elements
    bool ok = std::fwrite(&cnt, 8, 1, f) == 1 && (n == 0 || std::fwrite(kps, sizeof(snk_keypoint_f64), n, f) == (size_t)n) &&
              std::fwrite(&cnt, 8, 1, f) == 1 && (n == 0 || std::fwrite(desc, 32, n, f) == (size_t)n);
    ok = (std::fclose(f) == 0) && ok;
    SNK_REQUIRE(ok, "short write to %s", path);
    return SNK_OK;
}

extern "C" int snk_features_read(const char* path, snk_keypoint_f64* kps, uint64_t* desc, int capacity, int* n_out)
{
    SNK_REQUIRE(path && n_out && capacity >= 0, "bad argument");
    FILE* f = std::fopen(path, "rb");
    SNK_REQUIRE(f, "cannot open %s", path);
    uint64_t nk = 0, nd = 0;
    int rc = SNK_OK;
    if (std::fread(&nk, 8, 1, f) != 1)
        rc = SNK_ERR_INVALID;
    else if (nk > (uint64_t)capacity)
        rc = SNK_ERR_CAPACITY;
    else if (nk && std::fread(kps, sizeof(snk_keypoint_f64), nk, f) != nk)
        rc = SNK_ERR_INVALID;
    else if (std::fread(&nd, 8, 1, f) != 1 || nd != nk)
        rc = SNK_ERR_INVALID;
    else if (nd && std::fread(desc, 32, nd, f) != nd)
        rc = SNK_ERR_INVALID;
    std::fclose(f);
    if (rc == SNK_ERR_CAPACITY)
    {
        set_error("%s holds %llu features, capacity is %d", path, (unsigned long long)nk, capacity);
        *n_out = (int)std::min<uint64_t>(nk, 0x7fffffffu);  // lets the caller size its arrays and read again
    }
    if (rc == SNK_ERR_INVALID) set_error("%s is not a feature cache file", path);
    if (rc) return rc;
    *n_out = (int)nk;
    return SNK_OK;
}

extern "C" int snk_keypoints_widen(const snk_keypoint_f32* in, int n, snk_keypoint_f64* out)
{
    SNK_REQUIRE(n >= 0 && (n == 0 || (in && out)), "bad argument");
    for (int i = 0; i < n; ++i)
    {
        out[i].x = in[i].x, out[i].y = in[i].y, out[i].size = in[i].size, out[i].angle = in[i].angle;
        out[i].response = in[i].response, out[i].octave = in[i].octave, out[i]._pad = 0;
    }
    return SNK_OK;
}

// ------------------------------------------------------------------------------------------------------------------
// kNN-2 matching for callers that must not allocate device memory
// ------------------------------------------------------------------------------------------------------------------
// The staging area that snk_match_knn2 obtains from cudaMalloc (device copies of both descriptor sets and of the
// result records) lives here in a fixed array that is part of the library image: every GPU that loads libsnakeorb.so
// carries these 320 KiB until the process exits. In host code the name `g_knn2_noalloc_scratch` is only a stand-in for
// the symbol, not a device address, so the address is looked up once per device with cudaGetSymbolAddress (below) and
// that pointer is what the kernel receives.
static __device__ __align__(256) unsigned char g_knn2_noalloc_scratch[SNK_KNN2_NOALLOC_BYTES];

namespace
{
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mutex;
unsigned char* g_noalloc_ptr[kMaxDevices] = {};  // device address of the scratch array, per device ordinal

// Makes `device` current; the first call for an ordinal also checks that it is an sm_100 part and looks up the
// array's address there (both answers are kept, so later calls cost one cudaSetDevice).
int noalloc_scratch(int device, unsigned char** out)
{
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (g_noalloc_ptr[device])
        SNK_CUDA(cudaSetDevice(device));
    else
    {
        int rc = select_device(device);
        if (rc) return rc;
        unsigned char* ptr = nullptr;
        cudaError_t e      = cudaGetSymbolAddress((void**)&ptr, g_knn2_noalloc_scratch);
        if (e != cudaSuccess || !ptr)
        {
            set_error("cudaGetSymbolAddress(g_knn2_noalloc_scratch) failed on device %d: %s", device, cudaGetErrorString(e));
            return SNK_ERR_CUDA;
        }
        g_noalloc_ptr[device] = ptr;
    }
    *out = g_noalloc_ptr[device];
    return SNK_OK;
}
}  // namespace

extern "C" int snk_match_knn2_noalloc(int device, void* cuda_stream, const uint64_t* d1, int n, const uint64_t* d2, int m,
                                      snk_knn2* out)
{
    SNK_REQUIRE(out && n >= 0 && m >= 0 && (n == 0 || d1) && (m == 0 || d2), "bad argument");
    SNK_REQUIRE(device >= 0 && device < kMaxDevices, "device ordinal %d out of range", device);
    SNK_REQUIRE(n <= SNK_KNN2_NOALLOC_MAX && m <= SNK_KNN2_NOALLOC_MAX,
                "snk_match_knn2_noalloc handles at most %d descriptors per side (got n=%d, m=%d); "
                "use snk_match_knn2 for larger sets",
                SNK_KNN2_NOALLOC_MAX, n, m);
    if (n == 0) return SNK_OK;
    unsigned char* base = nullptr;
    int rc              = noalloc_scratch(device, &base);
    if (rc) return rc;
    // fixed carve-up of the array: [d1 | d2 | results]
    uint64_t* a = reinterpret_cast<uint64_t*>(base);
    uint64_t* b = reinterpret_cast<uint64_t*>(base + (size_t)SNK_KNN2_NOALLOC_MAX * 32);
    snk_knn2* o = reinterpret_cast<snk_knn2*>(base + (size_t)SNK_KNN2_NOALLOC_MAX * 64);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    // Nothing in the array has to start at zero: the kernel reads only the first n / m descriptors, which the two
    // copies below have just overwritten, and it writes all n result records before they are copied back.
    SNK_CUDA(cudaMemcpyAsync(a, d1, 32 * (size_t)n, cudaMemcpyHostToDevice, st));
    if (m) SNK_CUDA(cudaMemcpyAsync(b, d2, 32 * (size_t)m, cudaMemcpyHostToDevice, st));
    k_knn2<<<dim3((n + KNN_NT - 1) / KNN_NT, 1), KNN_NT, 0, st>>>(a, nullptr, n, n, b, nullptr, m, m, o);
    SNK_CUDA(cudaGetLastError());
    SNK_CUDA(cudaMemcpyAsync(out, o, sizeof(snk_knn2) * (size_t)n, cudaMemcpyDeviceToHost, st));
    SNK_CUDA(cudaStreamSynchronize(st));
    return SNK_OK;
}

#include "project_host.inl"
#include "bow_host.inl"
#include "mapping_host.inl"
