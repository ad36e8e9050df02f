// Validates snake-slam_b200/csrc/glibc_math.cuh (compiled for the host) against the running libm.
//   sinf/cosf : every float in [0, 6.5]  (angles reach the descriptor code as radians in [0, 2*pi])
//   atan2f    : all (m01, m10) on a dense integer lattice + >1e9 random integer moment pairs, then the
//               degrees()/wrap step of ORB::ComputeAngle
// Build: g++ -O2 -ffp-contract=off -mfma -fopenmp -I snake-slam_b200/csrc tools/check_glibc_math.cpp -o /tmp/chk -lm
#include "glibc_math.cuh"

#include <omp.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>

int main(int argc, char** argv)
{
    long long bad_sin = 0, bad_cos = 0, bad_atan = 0, n_atan = 0;
    const uint32_t hi = SNK_F2U(6.5f);
#pragma omp parallel for reduction(+ : bad_sin, bad_cos) schedule(static, 1 << 16)
    for (uint32_t u = 0; u <= hi; ++u)
    {
        float x = SNK_U2F(u);
        float a = sinf(x), b = snkmath::sinf_glibc(x);
        if (SNK_F2U(a) != SNK_F2U(b))
        {
            if (bad_sin < 5) printf("sin mismatch x=%a libm=%a mine=%a\n", x, a, b);
            ++bad_sin;
        }
        a = cosf(x), b = snkmath::cosf_glibc(x);
        if (SNK_F2U(a) != SNK_F2U(b))
        {
            if (bad_cos < 5) printf("cos mismatch x=%a libm=%a mine=%a\n", x, a, b);
            ++bad_cos;
        }
    }
    printf("sinf/cosf exhaustive over %u floats: %lld / %lld mismatches\n", hi + 1, bad_sin, bad_cos);

    long long reps = argc > 1 ? atoll(argv[1]) : 1200000000LL;
    const int M    = 1100000;  // |moment| bound: 255 * sum|u| over the disc is < 2^21
#pragma omp parallel reduction(+ : bad_atan, n_atan)
    {
        std::mt19937_64 rng(1234 + 77 * omp_get_thread_num());
#pragma omp for schedule(static)
        for (long long i = 0; i < reps; ++i)
        {
            uint64_t r = rng();
            int a = (int)(r % (2 * M + 1)) - M, b = (int)((r >> 32) % (2 * M + 1)) - M;
            if (i & 1)
            {  // small magnitudes exercise every reduction branch densely
                a = (int)(r % 4097) - 2048;
                b = (int)((r >> 32) % 4097) - 2048;
            }
            float y = (float)a, x = (float)b;
            float p = atan2f(y, x), q = snkmath::atan2f_fdlibm(y, x);
            ++n_atan;
            if (SNK_F2U(p) != SNK_F2U(q))
            {
                if (bad_atan < 5) printf("atan2 mismatch y=%d x=%d libm=%a mine=%a\n", a, b, p, q);
                ++bad_atan;
            }
        }
    }
    printf("atan2f over %lld integer pairs: %lld mismatches\n", n_atan, bad_atan);
    return (bad_sin || bad_cos || bad_atan) ? 1 : 0;
}
