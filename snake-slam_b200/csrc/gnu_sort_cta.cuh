// std::sort as libstdc++ executes it (see gnu_sort.cuh), evaluated by a whole CTA with the identical permutation.
//
//   * introsort's partition phase only depends on comparison outcomes, and the sub-ranges of one recursion level are
//     disjoint: they are partitioned concurrently, one warp per range, level by level (the depth budget that triggers
//     the heapsort fallback travels with each range).
//   * one unguarded Hoare partition is data parallel once written on the ORIGINAL range: with l_1 < l_2 < ... the
//     positions whose value is not before the pivot and r_1 > r_2 > ... those the pivot is not before, the serial
//     loop swaps (l_t, r_t) for t = 1..T, T = #{t : l_t < r_t}, and returns min(l_{T+1}, r_T) (l_1 if T = 0).
//   * the ranges of at most 16 elements the recursion leaves behind are mutually ordered (everything left of a cut is
//     not after the pivot, everything right of it not before), so the closing insertion pass never moves an element
//     across a cut: it is an independent (stable) insertion sort per leaf range, one thread each.
//
// Used by k_distribute (node order, FeatureDistribution.cpp:124) and k_bow_assemble (MiniBow2.h:104).
#pragma once
#include "gnu_sort.cuh"

namespace gnusort
{
// v[first, last) with last - first > 16; returns the cut of std::__unguarded_partition_pivot. Lp / Rq: scratch with
// one u16 per element of the range (indexed like v). All 32 lanes of the warp call.
template <class K, class C>
__device__ int warp_partition_pivot(K* v, int first, int last, uint16_t* Lp, uint16_t* Rq, int lane)
{
    if (lane == 0) median_to_first<K, C>(v, first, first + 1, first + (last - first) / 2, last - 1);
    __syncwarp();
    const K P   = v[first];
    const int s = first + 1, e = last;
    const unsigned lt = (1u << lane) - 1u;
    int nL = 0, nR = 0;
    for (int base = s; base < e; base += 32)
    {
        const int i      = base + lane;
        const bool valid = i < e;
        const K x        = valid ? v[i] : K(0);
        const bool isL = valid && !C::before(x, P), isR = valid && !C::before(P, x);
        const unsigned mL = __ballot_sync(0xffffffffu, isL), mR = __ballot_sync(0xffffffffu, isR);
        if (isL) Lp[s + nL + __popc(mL & lt)] = (uint16_t)i;
        if (isR) Rq[s + nR + __popc(mR & lt)] = (uint16_t)i;
        nL += __popc(mL);
        nR += __popc(mR);
    }
    __syncwarp();
    // T = number of leading t with l_t < r_t (monotone: l increases, r decreases)
    const int m = min(nL, nR);
    int T = 0;
    for (int base = 0; base < m; base += 32)
    {
        const int t   = base + lane;
        const bool ok = t < m && Lp[s + t] < Rq[s + nR - 1 - t];
        const unsigned bm = __ballot_sync(0xffffffffu, ok);
        if (bm == 0xffffffffu)
            T += 32;
        else
        {
            T += __ffs(~bm) - 1;
            break;
        }
    }
    for (int t = lane; t < T; t += 32)
    {
        const int a = Lp[s + t], b = Rq[s + nR - 1 - t];
        const K x = v[a];
        v[a]      = v[b];
        v[b]      = x;
    }
    int cut;
    if (T == 0)
        cut = Lp[s];
    else
    {
        const int rT = Rq[s + nR - T];
        cut          = (T < nL && Lp[s + T] < rT) ? (int)Lp[s + T] : rT;
    }
    __syncwarp();
    return cut;
}

// scratch bytes cta_sort needs for n elements (n < 65536)
__host__ __device__ constexpr int cta_sort_scratch_bytes(int n) { return 16 + 16 * (n / 17 + 2) + 4 * (n / 2 + 2) + 4 * n + 8; }

// std::sort(v, v + n) by a CTA of NT threads; every thread calls, the array is sorted when the call returns
// (ends with a barrier). scratch: shared memory, 8-byte aligned, cta_sort_scratch_bytes(n) bytes.
template <class K, class C, int NT>
__device__ void cta_sort(K* v, int n, unsigned char* scratch, int tid)
{
    if (n <= 1) return;  // uniform
    const int lane = tid & 31, wid = tid >> 5;
    const int S    = n / 17 + 2;
    int* ctrl      = reinterpret_cast<int*>(scratch);  // [0], [1]: ranges of the two level lists, [2]: leaf ranges
    uint2* segs    = reinterpret_cast<uint2*>(scratch + 16);
    uint32_t* leaf = reinterpret_cast<uint32_t*>(segs + 2 * S);
    uint16_t* Lp   = reinterpret_cast<uint16_t*>(leaf + (n / 2 + 2));
    uint16_t* Rq   = Lp + n;
    if (tid == 0)
    {
        int lg = 0;
        while ((n >> (lg + 1)) != 0) ++lg;
        ctrl[0] = ctrl[1] = ctrl[2] = 0;
        if (n > 16)
        {
            segs[0] = make_uint2((unsigned)n << 16, (unsigned)(2 * lg));  // first | last << 16, depth budget
            ctrl[0] = 1;
        }
        else
        {
            leaf[0] = (unsigned)n << 16;
            ctrl[2] = 1;
        }
    }
    __syncthreads();
    for (int cur = 0; ctrl[cur] > 0; cur ^= 1)
    {
        const int ns = ctrl[cur];
        uint2* mine  = segs + cur * S;
        uint2* nxt   = segs + (cur ^ 1) * S;
        for (int sidx = wid; sidx < ns; sidx += NT / 32)
        {
            const uint2 sg  = mine[sidx];
            const int first = (int)(sg.x & 0xffffu), last = (int)(sg.x >> 16);
            int depth       = (int)sg.y;
            if (depth == 0)
            {
                if (lane == 0) heap_sort<K, C>(v, first, last);  // std::__partial_sort(first, last, last)
                __syncwarp();
                continue;
            }
            --depth;
            const int cut = warp_partition_pivot<K, C>(v, first, last, Lp, Rq, lane);
            if (lane == 0)
            {
                if (last - cut > 16)
                    nxt[atomicAdd(&ctrl[cur ^ 1], 1)] = make_uint2((unsigned)cut | ((unsigned)last << 16), depth);
                else if (last - cut > 1)
                    leaf[atomicAdd(&ctrl[2], 1)] = (unsigned)cut | ((unsigned)last << 16);
                if (cut - first > 16)
                    nxt[atomicAdd(&ctrl[cur ^ 1], 1)] = make_uint2((unsigned)first | ((unsigned)cut << 16), depth);
                else if (cut - first > 1)
                    leaf[atomicAdd(&ctrl[2], 1)] = (unsigned)first | ((unsigned)cut << 16);
            }
        }
        __syncthreads();
        if (tid == 0) ctrl[cur] = 0;
        __syncthreads();
    }
    const int nl = ctrl[2];
    for (int i = tid; i < nl; i += NT)
    {
        const unsigned r = leaf[i];
        insertion<K, C>(v, (int)(r & 0xffffu), (int)(r >> 16));
    }
    __syncthreads();
}
}  // namespace gnusort
