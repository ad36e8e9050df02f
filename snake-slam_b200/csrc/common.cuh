// Shared host-side plumbing of libsnakeorb: error text, CUDA call checking, small device helpers.
#pragma once
#include "../../include/snakeorb.h"

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstring>

namespace snk
{
void set_error(const char* fmt, ...);

// returns SNK_ERR_CUDA from the enclosing function when a runtime call fails
#define SNK_CUDA(call)                                                                                        \
    do                                                                                                        \
    {                                                                                                         \
        cudaError_t e__ = (call);                                                                             \
        if (e__ != cudaSuccess)                                                                               \
        {                                                                                                     \
            snk::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__);      \
            return SNK_ERR_CUDA;                                                                              \
        }                                                                                                     \
    } while (0)

#define SNK_REQUIRE(cond, ...)          \
    do                                  \
    {                                   \
        if (!(cond))                    \
        {                               \
            snk::set_error(__VA_ARGS__); \
            return SNK_ERR_INVALID;     \
        }                               \
    } while (0)

// Select `device`, verify it is an sm_100 part (the library ships sm_100a code only; there is no fallback).
int select_device(int device);

template <typename T>
static inline T align_up(T v, T a)
{
    return (v + a - 1) / a * a;
}
}  // namespace snk

// ---- device helpers ------------------------------------------------------------------------------------------
__device__ __forceinline__ int warp_sum(int v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_incl_scan(int v, int lane)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
        int t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    return v;
}
