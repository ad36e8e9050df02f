// std::sort as libstdc++ (GCC 13) executes it, restated for device code.
//
// QuadtreeFeatureDistributor::Distribute (saiga/vision/features/FeatureDistribution.cpp:124) orders its inner nodes
// with std::sort on the keypoint count alone. std::sort is not stable, the counts are small integers with many ties,
// and the resulting order decides which nodes are split before the quota is reached and in which order keypoints are
// emitted. The permutation libstdc++ produces is therefore part of the reference's result and has to be reproduced:
// introsort (median-of-3 to front, unguarded Hoare partition, depth limit 2*floor(log2 n), heapsort fallback, threshold
// 16) followed by the guarded/unguarded insertion pass. The permutation depends only on the comparison outcomes, so
// sorting 32-bit (count << SHIFT | slot) keys with a comparator that looks at the count only moves the same elements to
// the same places as sorting the node structs. Keys are (count << SHIFT | slot), SHIFT = 12: up to 4096 nodes per
// sort (the node lists never exceed quota + 8 entries) and counts below 2^20.
//
// The functions are templates over the key type K and an ordering C (C::before(a, b) == "a sorts before b") so that
// other unstable std::sort calls of the reference can reuse them: MiniBow2's FeatureVector::setFeatures sorts
// (node id, feature index) pairs by node id alone (saiga/vision/slam/MiniBow2.h:104), see bow_kernels.cuh.
//
// tools/check_gnu_sort.cpp runs this code on the host against std::sort on millions of tie-heavy arrays.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#    define SNK_SORT_HD __host__ __device__ __forceinline__
#    define SNK_SORT_MEMBER static __host__ __device__ __forceinline__
#else
#    define SNK_SORT_HD static inline
#    define SNK_SORT_MEMBER static inline
#endif

namespace gnusort
{
constexpr int SHIFT = 12;
// the distributor's order: descending count, count in the bits above SHIFT
struct DescCount
{
    SNK_SORT_MEMBER bool before(uint32_t a, uint32_t b) { return (a >> SHIFT) > (b >> SHIFT); }
};
// ascending signed high word of a 64-bit key (payload in the low word)
struct AscHigh32
{
    SNK_SORT_MEMBER bool before(uint64_t a, uint64_t b) { return (int32_t)(a >> 32) < (int32_t)(b >> 32); }
};

template <class K, class C>
SNK_SORT_HD void swp(K* v, int i, int j)
{
    K t = v[i];
    v[i]       = v[j];
    v[j]       = t;
}

template <class K, class C>
SNK_SORT_HD void linear_insert_unguarded(K* v, int last)
{
    K val = v[last];
    int next     = last - 1;
    while (C::before(val, v[next]))
    {
        v[last] = v[next];
        last    = next;
        --next;
    }
    v[last] = val;
}

template <class K, class C>
SNK_SORT_HD void insertion(K* v, int first, int last)
{
    if (first == last) return;
    for (int i = first + 1; i != last; ++i)
    {
        if (C::before(v[i], v[first]))
        {
            K val = v[i];
            for (int k = i; k > first; --k) v[k] = v[k - 1];
            v[first] = val;
        }
        else
            linear_insert_unguarded<K, C>(v, i);
    }
}

template <class K, class C>
SNK_SORT_HD void push_heap(K* v, int first, int hole, int top, K value)
{
    int parent = (hole - 1) / 2;
    while (hole > top && C::before(v[first + parent], value))
    {
        v[first + hole] = v[first + parent];
        hole            = parent;
        parent          = (hole - 1) / 2;
    }
    v[first + hole] = value;
}

template <class K, class C>
SNK_SORT_HD void adjust_heap(K* v, int first, int hole, int len, K value)
{
    const int top = hole;
    int second    = hole;
    while (second < (len - 1) / 2)
    {
        second = 2 * (second + 1);
        if (C::before(v[first + second], v[first + second - 1])) --second;
        v[first + hole] = v[first + second];
        hole            = second;
    }
    if ((len & 1) == 0 && second == (len - 2) / 2)
    {
        second          = 2 * (second + 1);
        v[first + hole] = v[first + second - 1];
        hole            = second - 1;
    }
    push_heap<K, C>(v, first, hole, top, value);
}

template <class K, class C>
SNK_SORT_HD void heap_sort(K* v, int first, int last)
{
    const int len = last - first;
    if (len >= 2)
    {
        int parent = (len - 2) / 2;
        while (true)
        {
            K value = v[first + parent];
            adjust_heap<K, C>(v, first, parent, len, value);
            if (parent == 0) break;
            --parent;
        }
    }
    while (last - first > 1)
    {
        --last;
        K value = v[last];
        v[last]        = v[first];
        adjust_heap<K, C>(v, first, 0, last - first, value);
    }
}

template <class K, class C>
SNK_SORT_HD void median_to_first(K* v, int result, int a, int b, int c)
{
    if (C::before(v[a], v[b]))
    {
        if (C::before(v[b], v[c]))
            swp<K, C>(v, result, b);
        else if (C::before(v[a], v[c]))
            swp<K, C>(v, result, c);
        else
            swp<K, C>(v, result, a);
    }
    else if (C::before(v[a], v[c]))
        swp<K, C>(v, result, a);
    else if (C::before(v[b], v[c]))
        swp<K, C>(v, result, c);
    else
        swp<K, C>(v, result, b);
}

template <class K, class C>
SNK_SORT_HD int partition_unguarded(K* v, int first, int last, int pivot)
{
    // the pivot slot lies outside [first, last) and is never written here: read it once (one dependent shared-memory
    // load less per step of the single-threaded device sort)
    const K pv = v[pivot];
    while (true)
    {
        K a = v[first];
        while (C::before(a, pv)) a = v[++first];
        --last;
        K b = v[last];
        while (C::before(pv, b)) b = v[--last];
        if (!(first < last)) return first;
        v[first] = b;
        v[last]  = a;
        ++first;
    }
}

// std::sort(v, v + n). The recursion of the introsort loop (right part recursive, left part iterative) is unrolled
// onto an explicit stack; the order in which sub-ranges are processed does not influence the result because they are
// disjoint, but the depth budget each sub-range inherits does, so it is carried along.
template <class K, class C>
SNK_SORT_HD void sort_as(K* v, int n)
{
    if (n <= 1) return;
    int lg = 0;
    while ((n >> (lg + 1)) != 0) ++lg;  // floor(log2 n)
    int stack_first[64], stack_last[64], stack_depth[64];
    int sp          = 0;
    stack_first[sp] = 0, stack_last[sp] = n, stack_depth[sp] = 2 * lg, ++sp;
    while (sp)
    {
        --sp;
        int first = stack_first[sp], last = stack_last[sp], depth = stack_depth[sp];
        while (last - first > 16)
        {
            if (depth == 0)
            {
                heap_sort<K, C>(v, first, last);
                break;
            }
            --depth;
            const int mid = first + (last - first) / 2;
            median_to_first<K, C>(v, first, first + 1, mid, last - 1);
            const int cut = partition_unguarded<K, C>(v, first + 1, last, first);
            stack_first[sp] = cut, stack_last[sp] = last, stack_depth[sp] = depth, ++sp;
            last = cut;
        }
    }
    if (n > 16)
    {
        insertion<K, C>(v, 0, 16);
        for (int i = 16; i != n; ++i) linear_insert_unguarded<K, C>(v, i);
    }
    else
        insertion<K, C>(v, 0, n);
}
// the quadtree distributor's call (keys = count << SHIFT | slot)
SNK_SORT_HD void sort(uint32_t* v, int n) { sort_as<uint32_t, DescCount>(v, n); }
}  // namespace gnusort
