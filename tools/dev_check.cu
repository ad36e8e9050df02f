// Device unit checks of the pure functions the kernels are built from, against host computations:
//   fast_arc_score  vs brute force          glibc ports (atan2f/sinf/cosf) vs the running libm
//   gnusort::sort   vs std::sort
// Build (here):  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -fmad=false -I snake-slam_b200/csrc tools/dev_check.cu -o tools/dev_check
// Run (GPU box): ./tools/dev_check
#include "orb_kernels.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>
namespace snk { void set_error(const char*, ...) {} int select_device(int) { return 0; } }
using namespace snk;

__global__ void k_arc(const uint8_t* patches, int n, int* out)
{
    __shared__ uint8_t tile[FAST_DIM][FAST_PITCH];
    for (int p = blockIdx.x; p < n; p += gridDim.x)
    {
        __syncthreads();
        if (threadIdx.x < 49) tile[threadIdx.x / 7][threadIdx.x % 7] = patches[p * 49 + threadIdx.x];
        __syncthreads();
        if (threadIdx.x == 0)
        {
            out[p] = fast_arc_score(tile, 3, 3);
        }
    }
}
__global__ void k_math(const float* a, const float* b, int n, float* o_atan2, float* o_sin, float* o_cos)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    o_atan2[i] = snkmath::atan2f_fdlibm(a[i], b[i]);
    float ang  = fabsf(a[i]) * 1e-5f;
    o_sin[i]   = snkmath::sinf_glibc(ang);
    o_cos[i]   = snkmath::cosf_glibc(ang);
}
__global__ void k_sort(uint32_t* keys, const int* offs, int nprob)
{
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < nprob) gnusort::sort(keys + offs[p], offs[p + 1] - offs[p]);
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 2; } } while (0)

int main()
{
    std::mt19937 rng(5);
    int fails = 0;
    {   // arc
        const int n = 200000;
        std::vector<uint8_t> h(n * 49);
        for (int p = 0; p < n; ++p)
            for (int k = 0; k < 49; ++k) h[p * 49 + k] = (p % 3 == 0) ? rng() % 256 : (p % 3 == 1 ? 100 + rng() % 30 : ((rng() % 4) ? 40 : 200));
        uint8_t* d; int* o;
        CK(cudaMalloc(&d, h.size())); CK(cudaMalloc(&o, n * 4));
        CK(cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice));
        k_arc<<<1024, 64>>>(d, n, o);
        std::vector<int> got(n);
        CK(cudaMemcpy(got.data(), o, n * 4, cudaMemcpyDeviceToHost));
        const int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1}, dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
        int bad = 0;
        for (int p = 0; p < n; ++p)
        {
            const uint8_t* t = &h[p * 49];
            int v = t[3 * 7 + 3], dd[16], best = -256;
            for (int k = 0; k < 16; ++k) dd[k] = v - t[(3 + dy[k]) * 7 + 3 + dx[k]];
            for (int s = 0; s < 16; ++s)
            {
                int mn = 999, mx = -999;
                for (int j = 0; j < 9; ++j) mn = std::min(mn, dd[(s + j) & 15]), mx = std::max(mx, dd[(s + j) & 15]);
                best = std::max(best, std::max(mn, -mx));
            }
            if (best != got[p] && bad++ < 5) printf("arc mismatch: want %d got %d\n", best, got[p]);
        }
        printf("arc: %d / %d mismatches\n", bad, n);
        fails += bad != 0;
    }
    {   // math
        const int n = 4000000;
        std::vector<float> a(n), b(n);
        for (int i = 0; i < n; ++i)
        {
            int m = (i & 1) ? 2048 : 1100000;
            a[i] = (float)((int)(rng() % (2 * m + 1)) - m);
            b[i] = (float)((int)(rng() % (2 * m + 1)) - m);
        }
        float *da, *db, *o1, *o2, *o3;
        CK(cudaMalloc(&da, n * 4)); CK(cudaMalloc(&db, n * 4)); CK(cudaMalloc(&o1, n * 4)); CK(cudaMalloc(&o2, n * 4)); CK(cudaMalloc(&o3, n * 4));
        CK(cudaMemcpy(da, a.data(), n * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(db, b.data(), n * 4, cudaMemcpyHostToDevice));
        k_math<<<(n + 255) / 256, 256>>>(da, db, n, o1, o2, o3);
        std::vector<float> r1(n), r2(n), r3(n);
        CK(cudaMemcpy(r1.data(), o1, n * 4, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(r2.data(), o2, n * 4, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(r3.data(), o3, n * 4, cudaMemcpyDeviceToHost));
        int b1 = 0, b2 = 0, b3 = 0;
        for (int i = 0; i < n; ++i)
        {
            float w1 = atan2f(a[i], b[i]), ang = fabsf(a[i]) * 1e-5f, w2 = sinf(ang), w3 = cosf(ang);
            if (memcmp(&w1, &r1[i], 4) && b1++ < 3) printf("atan2f(%g,%g): libm %a dev %a\n", a[i], b[i], w1, r1[i]);
            if (memcmp(&w2, &r2[i], 4) && b2++ < 3) printf("sinf(%a): libm %a dev %a\n", ang, w2, r2[i]);
            if (memcmp(&w3, &r3[i], 4) && b3++ < 3) printf("cosf(%a): libm %a dev %a\n", ang, w3, r3[i]);
        }
        printf("math: atan2f %d, sinf %d, cosf %d mismatches of %d\n", b1, b2, b3, n);
        fails += (b1 || b2 || b3);
    }
    {   // sort
        const int nprob = 4000;
        std::vector<int> offs(nprob + 1, 0);
        std::vector<uint32_t> keys;
        for (int p = 0; p < nprob; ++p)
        {
            int n = 1 + rng() % (p % 40 == 0 ? 3000 : 500), maxc = 2 + rng() % (p % 2 ? 5 : 300);
            for (int i = 0; i < n; ++i) keys.push_back(((uint32_t)(2 + rng() % maxc) << gnusort::SHIFT) | (uint32_t)i);
            offs[p + 1] = (int)keys.size();
        }
        std::vector<uint32_t> want = keys;
        for (int p = 0; p < nprob; ++p)
            std::sort(want.begin() + offs[p], want.begin() + offs[p + 1], [](uint32_t x, uint32_t y) { return (x >> gnusort::SHIFT) > (y >> gnusort::SHIFT); });
        uint32_t* dk; int* dof;
        CK(cudaMalloc(&dk, keys.size() * 4)); CK(cudaMalloc(&dof, offs.size() * 4));
        CK(cudaMemcpy(dk, keys.data(), keys.size() * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dof, offs.data(), offs.size() * 4, cudaMemcpyHostToDevice));
        k_sort<<<(nprob + 63) / 64, 64>>>(dk, dof, nprob);
        std::vector<uint32_t> got(keys.size());
        CK(cudaMemcpy(got.data(), dk, keys.size() * 4, cudaMemcpyDeviceToHost));
        int bad = 0;
        for (size_t i = 0; i < keys.size(); ++i) bad += got[i] != want[i];
        printf("sort: %d mismatching entries of %zu\n", bad, keys.size());
        fails += bad != 0;
    }
    printf(fails ? "DEV_CHECK FAILED\n" : "DEV_CHECK OK\n");
    return fails;
}
