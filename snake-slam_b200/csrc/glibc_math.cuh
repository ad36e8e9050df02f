// Bit-exact device ports of the three libm functions the reference's ORB code calls on the CPU
// (saiga/vision/features/OrbDescriptors.cpp:52 atan2f, :66-67 cosf / sinf; libm = glibc 2.39, x86-64).
// The rBRIEF sample positions are iRound() of expressions in these values, so descriptors are only
// bit-exact if the functions are. Algorithms restated from their published descriptions:
//   atan2f / atanf : Sun fdlibm single-precision (argument reduction to 4 breakpoints + odd/even degree-11 polynomial)
//   sinf / cosf    : ARM "optimized routines" double-precision polynomials with quadrant reduction by 2/pi * 2^24;
//                    glibc selects its FMA-compiled variant on every AVX2/FMA x86-64 host, i.e. the a*b+c steps
//                    below are single-rounded — SNK_FMA reproduces exactly those contractions.
// tools/check_glibc_math.cpp compiles this header for the host and compares against the running libm:
// exhaustively over every float in [0, 6.5] for sinf/cosf and over >1e9 integer moment pairs for atan2f.
//
// All arithmetic goes through the SNK_* wrappers so neither nvcc (-fmad) nor gcc (-ffp-contract) can re-associate
// or fuse anything that is not fused in the original.
#pragma once
#include <stdint.h>

#if defined(__CUDA_ARCH__)
#    define SNK_HD __device__ __forceinline__
#    define SNK_FMUL(a, b) __fmul_rn((a), (b))
#    define SNK_FADD(a, b) __fadd_rn((a), (b))
#    define SNK_FSUB(a, b) __fsub_rn((a), (b))
#    define SNK_FDIV(a, b) __fdiv_rn((a), (b))
#    define SNK_DMUL(a, b) __dmul_rn((a), (b))
#    define SNK_DADD(a, b) __dadd_rn((a), (b))
#    define SNK_DSUB(a, b) __dsub_rn((a), (b))
#    define SNK_FMA(a, b, c) __fma_rn((a), (b), (c))
#    define SNK_F2U(f) __float_as_uint(f)
#    define SNK_U2F(u) __uint_as_float(u)
#else
#    include <math.h>
#    include <string.h>
#    define SNK_HD static inline
#    define SNK_FMUL(a, b) ((float)(a) * (float)(b))
#    define SNK_FADD(a, b) ((float)(a) + (float)(b))
#    define SNK_FSUB(a, b) ((float)(a) - (float)(b))
#    define SNK_FDIV(a, b) ((float)(a) / (float)(b))
#    define SNK_DMUL(a, b) ((double)(a) * (double)(b))
#    define SNK_DADD(a, b) ((double)(a) + (double)(b))
#    define SNK_DSUB(a, b) ((double)(a) - (double)(b))
#    define SNK_FMA(a, b, c) fma((a), (b), (c))
static inline uint32_t SNK_F2U(float f)
{
    uint32_t u;
    memcpy(&u, &f, 4);
    return u;
}
static inline float SNK_U2F(uint32_t u)
{
    float f;
    memcpy(&f, &u, 4);
    return f;
}
#endif

namespace snkmath
{
// ------------------------------------------------------------------------------------------------
// atanf (fdlibm float)
// ------------------------------------------------------------------------------------------------
SNK_HD float atanf_fdlibm(float x)
{
    // decimal literals exactly as published (the hex values quoted in fdlibm's comments are not all the nearest floats)
    const float atanhi[4] = {(float)4.6364760399e-01, (float)7.8539812565e-01, (float)9.8279368877e-01,
                             (float)1.5707962513e+00};
    const float atanlo[4] = {(float)5.0121582440e-09, (float)3.7748947079e-08, (float)3.4473217170e-08,
                             (float)7.5497894159e-08};
    const float aT0 = (float)3.3333334327e-01, aT1 = (float)-2.0000000298e-01, aT2 = (float)1.4285714924e-01,
                aT3 = (float)-1.1111110449e-01, aT4 = (float)9.0908870101e-02, aT5 = (float)-7.6918758452e-02,
                aT6 = (float)6.6610731184e-02, aT7 = (float)-5.8335702866e-02, aT8 = (float)4.9768779427e-02,
                aT9 = (float)-3.6531571299e-02, aT10 = (float)1.6285819933e-02;
    const int32_t hx = (int32_t)SNK_F2U(x);
    const int32_t ix = hx & 0x7fffffff;
    int id;
    if (ix >= 0x4c000000)
    {  // |x| >= 2^25
        if (ix > 0x7f800000) return SNK_FADD(x, x);
        float r = SNK_FADD(atanhi[3], atanlo[3]);
        return hx > 0 ? r : -r;
    }
    if (ix < 0x3ee00000)
    {  // |x| < 0.4375
        if (ix < 0x31000000) return x;  // |x| < 2^-29
        id = -1;
    }
    else
    {
        x = SNK_U2F((uint32_t)ix);  // fabsf
        if (ix < 0x3f980000)
        {
            if (ix < 0x3f300000)
            {
                id = 0;
                x  = SNK_FDIV(SNK_FSUB(SNK_FMUL(2.0f, x), 1.0f), SNK_FADD(2.0f, x));
            }
            else
            {
                id = 1;
                x  = SNK_FDIV(SNK_FSUB(x, 1.0f), SNK_FADD(x, 1.0f));
            }
        }
        else
        {
            if (ix < 0x401c0000)
            {
                id = 2;
                x  = SNK_FDIV(SNK_FSUB(x, 1.5f), SNK_FADD(1.0f, SNK_FMUL(1.5f, x)));
            }
            else
            {
                id = 3;
                x  = SNK_FDIV(-1.0f, x);
            }
        }
    }
    const float z = SNK_FMUL(x, x);
    const float w = SNK_FMUL(z, z);
    float s1 = SNK_FMUL(w, aT10);
    s1       = SNK_FMUL(w, SNK_FADD(aT8, s1));
    s1       = SNK_FMUL(w, SNK_FADD(aT6, s1));
    s1       = SNK_FMUL(w, SNK_FADD(aT4, s1));
    s1       = SNK_FMUL(w, SNK_FADD(aT2, s1));
    s1       = SNK_FMUL(z, SNK_FADD(aT0, s1));
    float s2 = SNK_FMUL(w, aT9);
    s2       = SNK_FMUL(w, SNK_FADD(aT7, s2));
    s2       = SNK_FMUL(w, SNK_FADD(aT5, s2));
    s2       = SNK_FMUL(w, SNK_FADD(aT3, s2));
    s2       = SNK_FMUL(w, SNK_FADD(aT1, s2));
    const float xs = SNK_FMUL(x, SNK_FADD(s1, s2));
    if (id < 0) return SNK_FSUB(x, xs);
    const float zz = SNK_FSUB(atanhi[id], SNK_FSUB(SNK_FSUB(xs, atanlo[id]), x));
    return (hx < 0) ? -zz : zz;
}

// ------------------------------------------------------------------------------------------------
// atan2f (fdlibm float) — finite inputs only matter here (integer moments), specials kept for completeness
// ------------------------------------------------------------------------------------------------
SNK_HD float atan2f_fdlibm(float y, float x)
{
    const float pi_o_4 = (float)7.8539818525e-01, pi_o_2 = (float)1.5707963705e+00, pi = (float)3.1415927410e+00,
                pi_lo = (float)-8.7422776573e-08;
    const int32_t hx = (int32_t)SNK_F2U(x), hy = (int32_t)SNK_F2U(y);
    const int32_t ix = hx & 0x7fffffff, iy = hy & 0x7fffffff;
    if (ix > 0x7f800000 || iy > 0x7f800000) return SNK_FADD(x, y);
    if (hx == 0x3f800000) return atanf_fdlibm(y);
    const int m = ((hy >> 31) & 1) | ((hx >> 30) & 2);
    if (iy == 0)
    {
        switch (m)
        {
            case 0:
            case 1:
                return y;
            case 2:
                return pi;  // pi + tiny
            default:
                return -pi;
        }
    }
    if (ix == 0) return (hy < 0) ? -pi_o_2 : pi_o_2;
    if (ix == 0x7f800000)
    {
        if (iy == 0x7f800000)
        {
            switch (m)
            {
                case 0:
                    return pi_o_4;
                case 1:
                    return -pi_o_4;
                case 2:
                    return SNK_FMUL(3.0f, pi_o_4);
                default:
                    return SNK_FMUL(-3.0f, pi_o_4);
            }
        }
        else
        {
            switch (m)
            {
                case 0:
                    return 0.0f;
                case 1:
                    return -0.0f;
                case 2:
                    return pi;
                default:
                    return -pi;
            }
        }
    }
    if (iy == 0x7f800000) return (hy < 0) ? -pi_o_2 : pi_o_2;
    const int32_t k = (iy - ix) >> 23;
    float z;
    if (k > 60)
        z = SNK_FADD(pi_o_2, SNK_FMUL(0.5f, pi_lo));
    else if (hx < 0 && k < -60)
        z = 0.0f;
    else
        z = atanf_fdlibm(SNK_U2F(SNK_F2U(SNK_FDIV(y, x)) & 0x7fffffffu));
    switch (m)
    {
        case 0:
            return z;
        case 1:
            return SNK_U2F(SNK_F2U(z) ^ 0x80000000u);
        case 2:
            return SNK_FSUB(pi, SNK_FSUB(z, pi_lo));
        default:
            return SNK_FSUB(SNK_FSUB(z, pi_lo), pi);
    }
}

// ------------------------------------------------------------------------------------------------
// sinf / cosf  (double polynomials; |x| < 120 is all the path can produce: angles are in [0, 2*pi])
// ------------------------------------------------------------------------------------------------
struct SinCosCoef
{
    double c0, c1, c2, c3, c4, s1, s2, s3;
};

SNK_HD float sincos_poly(double x, double x2, bool negate_cos_table, int n)
{
    // quadrant-0/1 table; the second table of the original is the same with c0..c4 negated
    const double C0 = 1.0, C1 = -0x1.ffffffd0c621cp-2, C2 = 0x1.55553e1068f19p-5, C3 = -0x1.6c087e89a359dp-10,
                 C4 = 0x1.99343027bf8c3p-16;
    const double S1 = -0x1.555545995a603p-3, S2 = 0x1.1107605230bc4p-7, S3 = -0x1.994eb3774cf24p-13;
    if ((n & 1) == 0)
    {
        const double x3 = SNK_DMUL(x, x2);
        const double s1 = SNK_FMA(x2, S3, S2);
        const double x7 = SNK_DMUL(x3, x2);
        const double s  = SNK_FMA(x3, S1, x);
        return (float)SNK_FMA(x7, s1, s);
    }
    const double sg = negate_cos_table ? -1.0 : 1.0;
    const double x4 = SNK_DMUL(x2, x2);
    const double c2 = SNK_FMA(x2, sg * C4, sg * C3);
    const double c1 = SNK_FMA(x2, sg * C2, sg * C1);
    const double x6 = SNK_DMUL(x4, x2);
    const double c  = SNK_FMA(x2, c1, sg * C0);
    return (float)SNK_FMA(x6, c2, c);
}

SNK_HD uint32_t abstop12(float x) { return (SNK_F2U(x) >> 20) & 0x7ff; }

// returns sin (want_cos = 0) or cos (want_cos = 1) of y, |y| < 120
SNK_HD float sincosf_glibc(float y, int want_cos)
{
    const double HPI_INV = 0x1.45F306DC9C883p+23;  // 2/pi * 2^24
    const double HPI     = 0x1.921FB54442D18p0;
    double x             = (double)y;
    if (abstop12(y) < 0x3f4u)  // |y| < pi/4 (abstop12(0x1.921FB6p-1f))
    {
        const double x2 = SNK_DMUL(x, x);
        if (abstop12(y) < 0x398u)  // |y| < 2^-12
            return want_cos ? 1.0f : y;
        return sincos_poly(x, x2, false, want_cos);
    }
    // reduce_fast
    const double r = SNK_DMUL(x, HPI_INV);
    const int n    = ((int32_t)r + 0x800000) >> 24;
    x              = SNK_FMA(-(double)n, HPI, x);
    const double s = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;  // sign[] = {1,-1,-1,1}
    return sincos_poly(SNK_DMUL(x, s), SNK_DMUL(x, x), (n & 2) != 0, n ^ want_cos);
}

SNK_HD float sinf_glibc(float y) { return sincosf_glibc(y, 0); }
SNK_HD float cosf_glibc(float y) { return sincosf_glibc(y, 1); }
}  // namespace snkmath
