// MiniBow2::TemplatedVocabulary::transform on the device (saiga/vision/slam/MiniBow2.h:883-983).
//
//   k_bow_descend   greedy descent of the vocabulary tree, 16 lanes per descriptor, children of a node contiguous
//   k_bow_assemble  one CTA per frame: FeatureVector::setFeatures and BowVector::set/normalize
//
// The reference sorts the (node id, feature index) pairs with std::sort on the node id alone; std::sort is unstable
// and the node ids repeat heavily, so the order of the features inside a node — which later decides the candidate
// order of the BoW matchers — is whatever libstdc++'s introsort leaves behind. That permutation is reproduced
// exactly and in parallel by gnusort::cta_sort (gnu_sort_cta.cuh).
// tests/test_gpu_bow.py compares against std::sort through the oracle on tie-heavy inputs of every size class.
#pragma once
#include "common.cuh"
#include "gnu_sort_cta.cuh"

namespace snk
{
struct VocDev
{
    int k, L, n_nodes, n_words;
    const int* child_begin;    // [n_nodes + 1] CSR over the children, in loadRaw's order
    const int* child_id;       // [n_nodes - 1] node id per CSR slot
    const uint4* child_desc;   // [2 * (n_nodes - 1)] descriptor per CSR slot (children of a node are contiguous)
    const float* weight;       // [n_nodes]
    const int* word_id;        // [n_nodes]
};

constexpr int BOW_MAX_FEATURES = 8192;

// ---- descent: transform(feature, levelsup), MiniBow2.h:937-983 --------------------------------------------------
// words[i] = (word id or -1) << 32 | weight bits;  feats[i] = (node id or -1) << 32 | i
// Batched: blockIdx.y = frame; frame f has counts[f] (or n_fixed when counts is null) descriptors at desc + 2*f*stride,
// its outputs at words/feats + f*stride.
__global__ void __launch_bounds__(256) k_bow_descend(VocDev V, const uint4* __restrict__ desc, const int* __restrict__ counts,
                                                     int n_fixed, int stride, int levelsup,
                                                     unsigned long long* __restrict__ words,
                                                     unsigned long long* __restrict__ feats)
{
    const int sub = threadIdx.x & 15;
    const int i   = (blockIdx.x * 256 + threadIdx.x) >> 4;
    const unsigned group = 0xffffu << (threadIdx.x & 16);
    const int n = counts ? min(counts[blockIdx.y], stride) : n_fixed;
    if (i >= n) return;  // whole 16-lane groups leave together
    desc += 2 * (size_t)blockIdx.y * stride;
    words += (size_t)blockIdx.y * stride;
    feats += (size_t)blockIdx.y * stride;
    const uint4 a = desc[2 * i], b = desc[2 * i + 1];
    const int nid_level = V.L - levelsup;
    int node = 0, nid = 0, level = 0;
    int begin = V.child_begin[0], cnt = V.child_begin[1] - begin;
    do
    {
        ++level;
        unsigned best = 0xffffffffu;  // distance << 20 | child slot: the first minimum wins (strict < at :961)
        for (int c0 = 0; c0 < cnt; c0 += 16)
        {
            const int c = c0 + sub;
            if (c < cnt)
            {
                const uint4 p = V.child_desc[2 * (begin + c)], q = V.child_desc[2 * (begin + c) + 1];
                const int d = __popc(a.x ^ p.x) + __popc(a.y ^ p.y) + __popc(a.z ^ p.z) + __popc(a.w ^ p.w) +
                              __popc(b.x ^ q.x) + __popc(b.y ^ q.y) + __popc(b.z ^ q.z) + __popc(b.w ^ q.w);
                best = min(best, ((unsigned)d << 20) | (unsigned)c);
            }
        }
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(group, best, o));
        node  = V.child_id[begin + (best & 0xfffffu)];
        begin = V.child_begin[node];
        cnt   = V.child_begin[node + 1] - begin;
        if (level == nid_level) nid = node;
    } while (cnt > 0);
    if (sub == 0)
    {
        const float w = V.weight[node];
        const int wid = (w > 0.f) ? V.word_id[node] : -1;
        const int fn  = (w > 0.f) ? nid : -1;
        words[i] = ((unsigned long long)(unsigned)wid << 32) | __float_as_uint(w);
        feats[i] = ((unsigned long long)(unsigned)fn << 32) | (unsigned)i;
    }
}

// bitonic sort of np2 (power of two) 64-bit keys in shared memory by the whole CTA
__device__ __forceinline__ void cta_bitonic_u64(unsigned long long* key, int np2, int tid, int nt)
{
    for (int k = 2; k <= np2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1)
        {
            for (int i = tid; i < np2; i += nt)
            {
                const int ixj = i ^ j;
                if (ixj > i)
                {
                    const unsigned long long a = key[i], b = key[ixj];
                    const bool up = (i & k) == 0;
                    if ((a > b) == up)
                    {
                        key[i]   = b;
                        key[ixj] = a;
                    }
                }
            }
            __syncthreads();
        }
}

constexpr int BOW_NT = 256;

// exclusive rank of `flag` among the threads of the CTA for one 256-wide chunk; total returned in *total
__device__ __forceinline__ int cta_rank(bool flag, int tid, int* s_warp, int* total)
{
    const unsigned m = __ballot_sync(0xffffffffu, flag);
    const int lane = tid & 31, wid = tid >> 5;
    if (lane == 0) s_warp[wid] = __popc(m);
    __syncthreads();
    int before = 0, all = 0;
#pragma unroll
    for (int w = 0; w < BOW_NT / 32; ++w)
    {
        const int c = s_warp[w];
        if (w < wid) before += c;
        all += c;
    }
    __syncthreads();
    *total = all;
    return before + __popc(m & ((1u << lane) - 1u));
}

// dynamic shared memory: key[np2] u64 | 8 * np2 bytes of sort scratch (reused as np2 floats)
// Batched like k_bow_descend: blockIdx.x = frame, outputs of frame f at f*stride (fv_offsets: f*(stride+1), counts: 2f).
__global__ void __launch_bounds__(BOW_NT) k_bow_assemble(const unsigned long long* __restrict__ words,
                                                          const unsigned long long* __restrict__ feats,
                                                          const int* __restrict__ in_counts, int n_fixed, int stride,
                                                          int* __restrict__ bow_ids, float* __restrict__ bow_vals,
                                                          int* __restrict__ fv_nodes, int* __restrict__ fv_offsets,
                                                          int* __restrict__ fv_indices, int* __restrict__ counts)
{
    const int n = in_counts ? min(in_counts[blockIdx.x], stride) : n_fixed;
    int np2     = 64;  // >= 64 keeps 8 * np2 >= gnusort::cta_sort_scratch_bytes(n)
    while (np2 < n) np2 <<= 1;
    {
        const size_t f = blockIdx.x;
        words += f * stride, feats += f * stride, bow_ids += f * stride, bow_vals += f * stride, fv_nodes += f * stride;
        fv_offsets += f * (stride + 1), fv_indices += f * stride, counts += 2 * f;
    }
    extern __shared__ __align__(16) unsigned char bow_smem[];
    unsigned long long* key = reinterpret_cast<unsigned long long*>(bow_smem);
    unsigned char* scratch  = reinterpret_cast<unsigned char*>(key + np2);  // 8 * np2 bytes
    float* vals             = reinterpret_cast<float*>(scratch);
    __shared__ int s_warp[BOW_NT / 32];
    __shared__ int s_n;
    __shared__ float s_norm;
    const int tid = threadIdx.x, lane = tid & 31;

    // ================= FeatureVector::setFeatures (MiniBow2.h:101-126) =================
    for (int i = tid; i < n; i += BOW_NT) key[i] = feats[i];
    __syncthreads();
    gnusort::cta_sort<unsigned long long, gnusort::AscHigh32, BOW_NT>(key, n, scratch, tid);
    {
        // entries of zero-weight words (node id -1) sort first and are skipped (:111)
        int skip_local = 0;
        for (int i = tid; i < n; i += BOW_NT) skip_local += ((int)(key[i] >> 32) == -1);
        skip_local = warp_sum(skip_local);
        if (tid == 0) s_n = 0;
        __syncthreads();
        if (lane == 0 && skip_local) atomicAdd(&s_n, skip_local);
        __syncthreads();
        const int skip = s_n;
        __syncthreads();
        int nodes_before = 0;
        for (int base = 0; base < n; base += BOW_NT)
        {
            const int i = base + tid;
            bool head = false;
            int id = 0;
            if (i < n && i >= skip)
            {
                id   = (int)(key[i] >> 32);
                head = (i == skip) || ((int)(key[i - 1] >> 32) != id);
                fv_indices[i - skip] = (int)(unsigned)key[i];
            }
            int total;
            const int r = cta_rank(head, tid, s_warp, &total);
            if (head)
            {
                fv_nodes[nodes_before + r]   = id;
                fv_offsets[nodes_before + r] = i - skip;
            }
            nodes_before += total;
        }
        if (tid == 0)
        {
            fv_offsets[nodes_before] = n - skip;
            counts[1]                = nodes_before;
        }
    }
    __syncthreads();

    // ================= BowVector::set + normalize (MiniBow2.h:46-93) =================
    // equal word ids carry equal weights, so the order std::sort leaves inside a word does not influence the sums
    for (int i = tid; i < np2; i += BOW_NT)
    {
        unsigned long long k2 = ~0ull;
        if (i < n)
        {
            const unsigned long long w = words[i];
            k2 = ((unsigned long long)((long long)(int)(w >> 32) + 1) << 32) | (w & 0xffffffffull);
        }
        key[i] = k2;
    }
    __syncthreads();
    cta_bitonic_u64(key, np2, tid, BOW_NT);
    int words_before = 0;
    for (int base = 0; base < n; base += BOW_NT)
    {
        const int i = base + tid;
        bool head = false;
        unsigned id1 = 0;
        if (i < n)
        {
            id1  = (unsigned)(key[i] >> 32);
            head = id1 != 0u && (i == 0 || (unsigned)(key[i - 1] >> 32) != id1);
        }
        int total;
        const int r = cta_rank(head, tid, s_warp, &total);
        if (head)
        {
            // back().second += f.second once per further occurrence, in float (:66)
            const float w = __uint_as_float((unsigned)key[i]);
            float acc     = w;
            for (int j = i + 1; j < n && (unsigned)(key[j] >> 32) == id1; ++j) acc += w;
            bow_ids[words_before + r] = (int)id1 - 1;
            vals[words_before + r]    = acc;
        }
        words_before += total;
    }
    __syncthreads();
    if (tid == 0)
    {
        float norm = 0.f;
        for (int i = 0; i < words_before; ++i) norm += fabsf(vals[i]);  // serial, in ascending word order (:86)
        s_norm    = norm;
        counts[0] = words_before;
    }
    __syncthreads();
    const float norm = s_norm;
    for (int i = tid; i < words_before; i += BOW_NT) bow_vals[i] = (norm > 0.f) ? vals[i] / norm : vals[i];
}

// FeatureVector::score (MiniBow2.h:134-170): L1 score of one query BowVector against many, one thread per candidate.
// The reference adds the terms of the common words in ascending word order in float: the merge stays serial per pair
// (its lower_bound jumps only skip non-common words, a plain two-pointer walk visits the same matches in the same order).
__global__ void k_bow_score(const int* __restrict__ q_ids, const float* __restrict__ q_vals, int nq,
                            const int* __restrict__ c_ids, const float* __restrict__ c_vals,
                            const int* __restrict__ c_offsets, int n_cand, float* __restrict__ scores)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_cand) return;
    int i = 0, j = c_offsets[c];
    const int je = c_offsets[c + 1];
    float score = 0.f;
    while (i < nq && j < je)
    {
        const int a = q_ids[i], b = c_ids[j];
        if (a == b)
        {
            const float vi = q_vals[i], wi = c_vals[j];
            score = __fadd_rn(score, __fsub_rn(__fsub_rn(fabsf(__fsub_rn(vi, wi)), fabsf(vi)), fabsf(wi)));
            ++i, ++j;
        }
        else if (a < b)
            ++i;
        else
            ++j;
    }
    scores[c] = __fmul_rn(score, -0.5f);
}
}  // namespace snk
