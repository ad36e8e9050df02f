// How far CUDA's double log / pow are from the host's glibc on the inputs the projection searches can reach, and
// whether any DECISION taken from them differs. The reference computes (saiga/vision/util/ScalePyramid.h:44-57,78)
//     prediction = log(ref_depth * Scale(ref_level) / dist) / log(scale_factor)
//     PredictionConsistent(prediction, octave):  (prediction - octave)^2 < 1.2^2
//     r = float(double(r0) * pow(scale_factor, prediction))          (SnakeORBMatcher.cpp:449-452)
// with dist inside [min_distance, max_distance] of EstimateMinMaxDistance, i.e. the ratio inside
// [1 / scale_factor, Scale(L-1) * scale_factor]. The kernels (project_kernels.cuh:124-126, mapping_kernels.cuh:82) use
// CUDA's log / pow (documented <= 1 / 2 ulp off the correctly rounded value; glibc's are < 1 ulp off), so single values
// may differ in the last bit; this tool counts those and the decisions that flip because of them.
// Build (here):  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -fmad=false --extended-lambda --expt-relaxed-constexpr -Xcompiler -fopenmp tools/scale_check.cu -o tools/scale_check
// Run (GPU box): ./tools/scale_check [millions of samples, default 1024]
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

struct In
{
    double ratio;  // ref_depth * Scale(ref_level) / dist
    float r0;      // RadiusByViewingCos * th
};
struct Out
{
    double prediction;
    float r;
    uint32_t consistent;  // bit o: PredictionConsistent(prediction, o), o = 0..15
};

__host__ __device__ inline uint64_t mix(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
struct Bounds
{
    double v[16];  // scale_factor^(o +- 1.2), o = 0..7, computed once on the host
};
__host__ __device__ inline In make_input(uint64_t i, double sf, double smax, const Bounds& bnd)
{
    const uint64_t a = mix(i), b = mix(a);
    const double u = (double)(a >> 11) * (1.0 / 9007199254740992.0);
    const double lo = 1.0 / sf, hi = smax * sf;
    In in;
    // one sample in four is drawn next to an integer scale level +- 1.2 (where PredictionConsistent decides)
    if ((b & 3) == 0)
    {
        const double eps = ((double)((b >> 12) & 0xfffff) - 524288.0) * 2.220446049250313e-16;  // +- 2^19 ulp
        in.ratio = bnd.v[(b >> 2) % 16] * (1.0 + eps);
        if (!(in.ratio >= lo && in.ratio <= hi)) in.ratio = lo + u * (hi - lo);
    }
    else
        in.ratio = lo + u * (hi - lo);
    const float r0s[6] = {2.5f, 4.0f, 10.0f, 16.0f, 12.5f, 20.0f};
    in.r0 = r0s[(b >> 40) % 6];
    return in;
}
template <typename F1, typename F2>
__host__ __device__ inline Out evaluate(const In& in, double sf, double log_sf, F1 logf_, F2 powf_)
{
    Out o;
    o.prediction = logf_(in.ratio) / log_sf;
    o.r          = (float)((double)in.r0 * powf_(sf, o.prediction));
    o.consistent = 0;
    for (int k = 0; k < 16; ++k)
    {
        const double e = o.prediction - k;
        if (e * e < 1.2 * 1.2) o.consistent |= 1u << k;
    }
    return o;
}
__global__ void k_eval(uint64_t first, int n, double sf, double log_sf, double smax, Bounds bnd, Out* out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = evaluate(make_input(first + i, sf, smax, bnd), sf, log_sf, [](double x) { return log(x); },
                      [](double a, double b) { return pow(a, b); });
}

int main(int argc, char** argv)
{
    const long long total = (argc > 1 ? atoll(argv[1]) : 1024) * 1000000ll;
    const double sf = (double)1.2f, log_sf = std::log(sf);
    double smax = 1;
    for (int i = 1; i < 8; ++i) smax *= sf;
    Bounds bnd;
    for (int o = 0; o < 8; ++o) bnd.v[2 * o] = std::exp((o - 1.2) * log_sf), bnd.v[2 * o + 1] = std::exp((o + 1.2) * log_sf);
    const int chunk = 1 << 24;
    Out* d;
    cudaMalloc(&d, sizeof(Out) * chunk);
    std::vector<Out> h(chunk);
    long long pred_diff = 0, r_diff = 0, cons_diff = 0, max_ulp = 0, done = 0;
    for (uint64_t first = 0; (long long)first < total; first += chunk)
    {
        const int n = (int)std::min<long long>(chunk, total - first);
        k_eval<<<(n + 255) / 256, 256>>>(first, n, sf, log_sf, smax, bnd, d);
        if (cudaMemcpy(h.data(), d, sizeof(Out) * n, cudaMemcpyDeviceToHost) != cudaSuccess)
        {
            printf("CUDA error: %s\n", cudaGetErrorString(cudaGetLastError()));
            return 1;
        }
#pragma omp parallel for reduction(+ : pred_diff, r_diff, cons_diff) reduction(max : max_ulp)
        for (int i = 0; i < n; ++i)
        {
            const Out w = evaluate(make_input(first + i, sf, smax, bnd), sf, log_sf, [](double x) { return std::log(x); },
                                   [](double a, double b) { return std::pow(a, b); });
            int64_t a, b;
            memcpy(&a, &w.prediction, 8), memcpy(&b, &h[i].prediction, 8);
            if (a != b)
            {
                ++pred_diff;
                max_ulp = std::max<long long>(max_ulp, std::llabs(a - b));
            }
            if (memcmp(&w.r, &h[i].r, 4)) ++r_diff;
            if (w.consistent != h[i].consistent) ++cons_diff;
        }
        done += n;
    }
    printf("{\"samples\": %lld, \"prediction_last_bit_differences\": %lld, \"max_ulp\": %lld, "
           "\"radius_float_differences\": %lld, \"prediction_consistent_flips\": %lld}\n",
           done, pred_diff, max_ulp, r_diff, cons_diff);
    return 0;
}
