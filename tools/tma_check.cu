// Stand-alone check of the TMA box load used by k_fast: 3-D u8 tensor (x, y, image), box 80 x H x 1, descriptor
// passed as a __grid_constant__ kernel parameter (single and inside a struct array). Finding recorded here: the
// innermost start coordinate must be 16-byte aligned (x0 = 339 raised "illegal instruction" on B200); boxes that
// reach outside the tensor are fine (zero filled).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 tools/tma_check.cu -o tools/tma_check
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>

struct Maps { CUtensorMap level[4]; };

template <bool NESTED>
__global__ void k(const __grid_constant__ CUtensorMap single, const __grid_constant__ Maps maps, int which, int x0, int y0, int img,
                  int box_h, uint8_t* out)
{
    __shared__ __align__(128) uint8_t tile[66][80];
    __shared__ __align__(8) unsigned long long bar;
    const unsigned bar_addr = (unsigned)__cvta_generic_to_shared(&bar), tile_addr = (unsigned)__cvta_generic_to_shared(&tile[0][0]);
    if (threadIdx.x == 0)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_addr));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0)
    {
        const CUtensorMap* m = NESTED ? &maps.level[which] : &single;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_addr), "r"(80 * box_h) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(tile_addr),
                     "l"(reinterpret_cast<unsigned long long>(m)), "r"(x0), "r"(y0), "r"(img), "r"(bar_addr)
                     : "memory");
    }
    unsigned done = 0;
    for (int spin = 0; !done && spin < (1 << 22); ++spin)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(bar_addr) : "memory");
    __syncthreads();
    for (int i = threadIdx.x; i < 80 * box_h; i += blockDim.x) out[i] = done ? tile[i / 80][i % 80] : 0xEE;
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e), __LINE__); return 2; } } while (0)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main()
{
    const int W = 752, H = 480, B = 3, pitch = 752, box_h = 38;
    std::vector<uint8_t> h((size_t)B * H * pitch);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (uint8_t)(i * 2654435761u >> 13);
    uint8_t *d, *o;
    CK(cudaMalloc(&d, h.size())); CK(cudaMalloc(&o, 80 * 66));
    CK(cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice));
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    if (!p) { printf("no encoder\n"); return 3; }
    CUtensorMap m; Maps ms{};
    cuuint64_t dims[3] = {W, H, B}, strides[2] = {(cuuint64_t)pitch, (cuuint64_t)pitch * H};
    cuuint32_t box[3] = {80, (cuuint32_t)box_h, 1}, es[3] = {1, 1, 1};
    CUresult r = ((EncodeTiledFn)p)(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode: %d\n", (int)r);
    if (r) return 4;
    ms.level[2] = m;
    int fails = 0;
    for (int nested = 0; nested < 2; ++nested)
        for (int t = 0; t < 3; ++t)
        {
            const int x0 = (t == 0) ? 32 : (t == 1 ? 704 : 336), y0 = (t == 0) ? 17 : (t == 1 ? 450 : 201), img = t;
            if (nested) k<true><<<1, 128>>>(m, ms, 2, x0, y0, img, box_h, o); else k<false><<<1, 128>>>(m, ms, 2, x0, y0, img, box_h, o);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("nested=%d t=%d: %s\n", nested, t, cudaGetErrorString(e)); return 5; }
            std::vector<uint8_t> got(80 * box_h);
            CK(cudaMemcpy(got.data(), o, got.size(), cudaMemcpyDeviceToHost));
            int bad = 0;
            for (int y = 0; y < box_h; ++y)
                for (int x = 0; x < 80; ++x)
                {
                    const int gx = x0 + x, gy = y0 + y;
                    const uint8_t want = (gx < W && gy < H) ? h[((size_t)img * H + gy) * pitch + gx] : 0;
                    bad += got[y * 80 + x] != want;
                }
            printf("nested=%d box at (%d,%d,img %d): %d mismatching bytes\n", nested, x0, y0, img, bad);
            fails += bad != 0;
        }
    printf(fails ? "TMA_CHECK FAILED\n" : "TMA_CHECK OK\n");
    return fails;
}
