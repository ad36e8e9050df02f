// TEST INFRASTRUCTURE ONLY — see bow_ref.h. Restates MiniBow2.h; the std::sort calls are kept as std::sort because
// the permutation libstdc++ produces for equal keys is part of the reference's result (feature order inside a
// FeatureVector node decides the candidate order of the BoW matchers).
#include "bow_ref.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>

namespace bowref
{
namespace
{
inline int hamming256(const uint64_t* a, const uint64_t* b)
{
    return __builtin_popcountll(a[0] ^ b[0]) + __builtin_popcountll(a[1] ^ b[1]) + __builtin_popcountll(a[2] ^ b[2]) +
           __builtin_popcountll(a[3] ^ b[3]);
}
}  // namespace

bool Vocabulary::from_arrays(int k_, int L_, int n, const int32_t* par, const double* w, const int32_t* wid,
                             const uint64_t* d)
{
    k = k_, L = L_;
    parent.assign(par, par + n);
    word_id.assign(wid, wid + n);
    weight.resize(n);
    for (int i = 0; i < n; ++i) weight[i] = (float)w[i];
    desc.assign(d, d + 4 * (size_t)n);
    children.assign(n, {});
    // loadRaw: every node except the root is appended to its parent's child list in file order (MiniBow2.h:1040)
    for (int i = 1; i < n; ++i)
    {
        if (par[i] < 0 || par[i] >= n) return false;
        children[par[i]].push_back(i);
    }
    n_words = 0;
    for (int i = 1; i < n; ++i)
        if (children[i].empty()) n_words = std::max(n_words, wid[i] + 1);
    return n > 1 && !children[0].empty();
}

bool Vocabulary::load_raw(const std::string& path)
{
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return false;
    int32_t hdr[4];
    uint64_t count = 0;
    bool ok = fread(hdr, 4, 4, f) == 4 && fread(&count, 8, 1, f) == 1;
    std::vector<int32_t> par, wid;
    std::vector<double> w;
    std::vector<uint64_t> d;
    for (uint64_t i = 0; ok && i < count; ++i)
    {
        int32_t id, p, word;
        double weight_;
        uint64_t dd[4];
        ok = fread(&id, 4, 1, f) == 1 && fread(&p, 4, 1, f) == 1 && fread(&weight_, 8, 1, f) == 1 &&
             fread(&word, 4, 1, f) == 1 && fread(dd, 8, 4, f) == 4 && id == (int32_t)i;
        par.push_back(p), wid.push_back(word), w.push_back(weight_);
        d.insert(d.end(), dd, dd + 4);
    }
    fclose(f);
    return ok && from_arrays(hdr[0], hdr[1], (int)count, par.data(), w.data(), wid.data(), d.data());
}

bool Vocabulary::save_raw(const std::string& path) const
{
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) return false;
    const int32_t hdr[4] = {k, L, 0, 0};
    const uint64_t count = parent.size();
    fwrite(hdr, 4, 4, f);
    fwrite(&count, 8, 1, f);
    std::vector<std::pair<int32_t, int32_t>> words(n_words);
    for (size_t i = 0; i < count; ++i)
    {
        const int32_t id = (int32_t)i;
        const double w   = weight[i];
        fwrite(&id, 4, 1, f), fwrite(&parent[i], 4, 1, f), fwrite(&w, 8, 1, f), fwrite(&word_id[i], 4, 1, f);
        fwrite(&desc[4 * i], 8, 4, f);
        if (i > 0 && children[i].empty()) words[word_id[i]] = {word_id[i], id};
    }
    const uint64_t nw = words.size();
    fwrite(&nw, 8, 1, f);
    fwrite(words.data(), 8, words.size(), f);
    fclose(f);
    return true;
}

void transform(const Vocabulary& v, const uint64_t* desc, int n, int levelsup, std::vector<int32_t>& bow_ids,
               std::vector<float>& bow_vals, std::vector<int32_t>& fv_nodes, std::vector<int32_t>& fv_offsets,
               std::vector<int32_t>& fv_indices)
{
    std::vector<std::pair<int32_t, float>> words(n);
    std::vector<std::pair<int32_t, int32_t>> feats(n);
    const int nid_level = v.L - levelsup;
    for (int i = 0; i < n; ++i)
    {
        // transform(feature, levelsup), MiniBow2.h:937-983: greedy descent, first minimum wins
        int nid = 0, final_id = 0, level = 0;
        do
        {
            ++level;
            const auto& ch = v.children[final_id];
            final_id       = ch[0];
            int best       = hamming256(desc + 4 * (size_t)i, &v.desc[4 * (size_t)final_id]);
            for (size_t c = 1; c < ch.size(); ++c)
            {
                const int d = hamming256(desc + 4 * (size_t)i, &v.desc[4 * (size_t)ch[c]]);
                if (d < best) best = d, final_id = ch[c];
            }
            if (level == nid_level) nid = final_id;
        } while (!v.children[final_id].empty());
        const float w = v.weight[final_id];
        if (w > 0)
            words[i] = {v.word_id[final_id], w}, feats[i] = {nid, i};
        else
            words[i] = {-1, w}, feats[i] = {-1, i};
    }
    // BowVector::set (MiniBow2.h:46-76) + normalize (:82-93)
    bow_ids.clear(), bow_vals.clear();
    std::sort(words.begin(), words.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
    int current = -1;
    for (auto& f : words)
    {
        if (f.first == -1) continue;
        if (f.first != current)
        {
            bow_ids.push_back(f.first), bow_vals.push_back(f.second);
            current = f.first;
        }
        else
            bow_vals.back() += f.second;
    }
    float norm = 0.0f;
    for (float x : bow_vals) norm += std::abs(x);
    if (norm > 0.0)
        for (float& x : bow_vals) x /= norm;
    // FeatureVector::setFeatures (MiniBow2.h:101-126)
    fv_nodes.clear(), fv_offsets.clear(), fv_indices.clear();
    std::sort(feats.begin(), feats.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
    current = -1;
    for (auto& f : feats)
    {
        if (f.first == -1) continue;
        if (f.first != current)
        {
            fv_nodes.push_back(f.first);
            fv_offsets.push_back((int32_t)fv_indices.size());
            current = f.first;
        }
        fv_indices.push_back(f.second);
    }
    fv_offsets.push_back((int32_t)fv_indices.size());
}

float score(const int32_t* ids1, const float* v1, int n1, const int32_t* ids2, const float* v2, int n2)
{
    int i = 0, j = 0;
    float s = 0;
    while (i < n1 && j < n2)
    {
        if (ids1[i] == ids2[j])
        {
            s += std::abs(v1[i] - v2[j]) - std::abs(v1[i]) - std::abs(v2[j]);
            ++i, ++j;
        }
        else if (ids1[i] < ids2[j])
            i = (int)(std::lower_bound(ids1 + i, ids1 + n1, ids2[j]) - ids1);
        else
            j = (int)(std::lower_bound(ids2 + j, ids2 + n2, ids1[i]) - ids2);
    }
    return s * -0.5f;
}
}  // namespace bowref
