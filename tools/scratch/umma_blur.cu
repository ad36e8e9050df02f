// Development harness for the tcgen05 blur experiment (verdict item 3b): one level, one kernel, checked against a
// plain CPU loop of the same fixed-point arithmetic.   nvcc -gencode arch=compute_100a,code=sm_100a -o umma_blur umma_blur.cu
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <vector>

#include "../../snake-slam_b200/csrc/blur_umma.cuh"

#define CK(x)                                                                                       \
    do                                                                                              \
    {                                                                                               \
        cudaError_t e_ = (x);                                                                       \
        if (e_ != cudaSuccess)                                                                      \
        {                                                                                           \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__);        \
            return 2;                                                                               \
        }                                                                                           \
    } while (0)

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int reflect(int p, int len)
{
    if (p < 0) p = -p;
    if (p >= len) p = 2 * (len - 1) - p;
    return p;
}

int main(int argc, char** argv)
{
    const int W = argc > 1 ? atoi(argv[1]) : 752, H = argc > 2 ? atoi(argv[2]) : 480, NIMG = argc > 3 ? atoi(argv[3]) : 4;
    const int reps = argc > 4 ? atoi(argv[4]) : 0;
    const int pitch = (W + 15) / 16 * 16;
    const size_t img_stride = (size_t)pitch * H;
    std::vector<uint8_t> src(img_stride * NIMG), want(img_stride * NIMG, 0), got(img_stride * NIMG, 0);
    srand(1);
    for (auto& b : src) b = (uint8_t)(rand() >> 7);
    static const int wt[4] = {56, 48, 34, 18};
    std::vector<int> row((size_t)W * H);
    for (int i = 0; i < NIMG; ++i)
    {
        const uint8_t* s = src.data() + i * img_stride;
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x)
            {
                int a = 0;
                for (int d = -3; d <= 3; ++d) a += wt[abs(d)] * s[(size_t)y * pitch + reflect(x + d, W)];
                row[(size_t)y * W + x] = a;
            }
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x)
            {
                int a = 0;
                for (int d = -3; d <= 3; ++d) a += wt[abs(d)] * row[(size_t)reflect(y + d, H) * W + x];
                want[i * img_stride + (size_t)y * pitch + x] = (uint8_t)((a + 32768) >> 16);
            }
    }
    uint8_t *d_src, *d_dst;
    int* d_err;
    CK(cudaMalloc(&d_src, src.size()));
    CK(cudaMalloc(&d_dst, src.size()));
    CK(cudaMalloc(&d_err, 4));
    CK(cudaMemset(d_err, 0, 4));
    CK(cudaMemset(d_dst, 0, src.size()));
    CK(cudaMemcpy(d_src, src.data(), src.size(), cudaMemcpyHostToDevice));
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    EncodeTiledFn fn = (EncodeTiledFn)p;
    snk::UmmaBlurMaps maps;
    const cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)NIMG};
    const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)img_stride};
    const cuuint32_t estr[3] = {1, 1, 1};
    const cuuint32_t box_in[3] = {snk::UB_K0, snk::UB_ROWS_IN, 1}, box_inx[3] = {snk::UB_K1, snk::UB_ROWS_IN, 1}, box_out[3] = {snk::UB_OUT_COLS, snk::UB_ROWS_OUT, 1};
    CUresult r = fn(&maps.src[0], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d_src, dims, strides, box_in, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return printf("encode src failed %d\n", (int)r), 2;
    r = fn(&maps.srcx[0], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d_src, dims, strides, box_inx, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
           CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return printf("encode srcx failed %d\n", (int)r), 2;
    r = fn(&maps.dst[0], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d_dst, dims, strides, box_out, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return printf("encode dst failed %d\n", (int)r), 2;
    // one task per column strip; the band table: the strips' column bands, then the row bands of every row tile
    std::vector<snk::UmmaBlurTask> tasks;
    std::vector<snk::UmmaBandJob> jobs;
    unsigned off = 0;
    const int n_row_tiles = (H + snk::UB_ROWS_OUT - 1) / snk::UB_ROWS_OUT;
    const unsigned wr_off = (unsigned)((W + snk::UB_OUT_COLS - 1) / snk::UB_OUT_COLS) * snk::UB_WC_BYTES;
    for (int x0 = 0; x0 < W; x0 += snk::UB_OUT_COLS)
    {
        tasks.push_back(snk::UmmaBlurTask{0, (short)x0, (short)W, (short)H, (short)n_row_tiles, (short)(x0 / snk::UB_OUT_COLS),
                                          (short)((W + snk::UB_OUT_COLS - 1) / snk::UB_OUT_COLS),
                                          (short)(x0 >= 3 && x0 + snk::UB_OUT_COLS + 3 <= W), off, wr_off});
        jobs.push_back(snk::UmmaBandJob{off, (short)x0, (short)W, 0, 0});
        off += snk::UB_WC_BYTES;
    }
    for (int t = 0; t < n_row_tiles; ++t)
    {
        jobs.push_back(snk::UmmaBandJob{off, (short)(t * snk::UB_ROWS_OUT), (short)H, 1, 0});
        off += snk::UB_WR_BYTES;
    }
    snk::UmmaBlurTask* d_tasks;
    snk::UmmaBandJob* d_jobs;
    unsigned char* d_bands;
    CK(cudaMalloc(&d_tasks, tasks.size() * sizeof(tasks[0])));
    CK(cudaMalloc(&d_jobs, jobs.size() * sizeof(jobs[0])));
    CK(cudaMalloc(&d_bands, off));
    CK(cudaMemcpy(d_tasks, tasks.data(), tasks.size() * sizeof(tasks[0]), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_jobs, jobs.data(), jobs.size() * sizeof(jobs[0]), cudaMemcpyHostToDevice));
    snk::k_umma_bands<<<(int)jobs.size(), 128>>>(d_jobs, d_bands);
    CK(cudaGetLastError());
    const int tiles_per_img = (int)tasks.size() * n_row_tiles;
    int n_sm = 0;
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, 0);
    const int grid = (int)std::min<long long>((long long)snk::UB_CTAS_PER_SM * n_sm, (long long)tiles_per_img * NIMG);
    CK(cudaFuncSetAttribute(snk::k_blur_umma, cudaFuncAttributeMaxDynamicSharedMemorySize, snk::UB_SMEM));
    snk::k_blur_umma<<<grid, snk::UB_NT, snk::UB_SMEM>>>(maps, d_tasks, (int)tasks.size(), NIMG, tiles_per_img, d_bands, d_err);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    int err = 0;
    CK(cudaMemcpy(&err, d_err, 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(got.data(), d_dst, got.size(), cudaMemcpyDeviceToHost));
    size_t bad = 0, first = (size_t)-1;
    for (int i = 0; i < NIMG; ++i)
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x)
            {
                const size_t o = i * img_stride + (size_t)y * pitch + x;
                if (got[o] != want[o])
                {
                    if (first == (size_t)-1) first = o;
                    ++bad;
                }
            }
    printf("W %d H %d images %d tasks %zu err %d mismatches %zu", W, H, NIMG, tasks.size(), err, bad);
    if (bad)
    {
        const size_t o = first % img_stride;
        printf(" first at img %zu y %zu x %zu: got %d want %d", first / img_stride, o / pitch, o % pitch, got[first], want[first]);
        int shown = 0;
        for (int y = 0; y < H && shown < 12; ++y)
            for (int x = 0; x < W && shown < 12; ++x)
                if (got[(size_t)y * pitch + x] != want[(size_t)y * pitch + x])
                    printf("\n  (%d,%d) got %d want %d", x, y, got[(size_t)y * pitch + x], want[(size_t)y * pitch + x]), ++shown;
    }
    printf("\n");
    if (reps > 0 && !bad && !err)
    {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0), cudaEventCreate(&e1);
        cudaEventRecord(e0);
        for (int k = 0; k < reps; ++k) snk::k_blur_umma<<<grid, snk::UB_NT, snk::UB_SMEM>>>(maps, d_tasks, (int)tasks.size(), NIMG, tiles_per_img, d_bands, d_err);
        cudaEventRecord(e1);
        CK(cudaDeviceSynchronize());
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        printf("%.4f ms per launch, %.1f Gpx/s\n", ms / reps, (double)W * H * NIMG / (ms / reps * 1e-3) * 1e-9);
    }
    return bad || err ? 1 : 0;
}
