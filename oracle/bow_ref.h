// TEST INFRASTRUCTURE ONLY — CPU restatement of MiniBow2's vocabulary file format and transform
// (External/saiga/src/saiga/vision/slam/MiniBow2.h). Never linked into the product.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace bowref
{
struct Vocabulary
{
    int k = 0, L = 0;
    std::vector<int32_t> parent, word_id;
    std::vector<float> weight;  // loadRaw narrows the stored double (MiniBow2.h:1037-1039)
    std::vector<uint64_t> desc; // 4 per node
    std::vector<std::vector<int32_t>> children;
    int n_words = 0;
    bool load_raw(const std::string& path);  // MiniBow2.h:1020-1052
    bool from_arrays(int k, int L, int n, const int32_t* parent, const double* weight, const int32_t* word_id,
                     const uint64_t* desc);
    bool save_raw(const std::string& path) const;  // MiniBow2.h:1056-1075
};

// TemplatedVocabulary::transform (MiniBow2.h:883-983) -> BowVector (ids, vals) and FeatureVector as CSR
void transform(const Vocabulary& v, const uint64_t* desc, int n, int levelsup, std::vector<int32_t>& bow_ids,
               std::vector<float>& bow_vals, std::vector<int32_t>& fv_nodes, std::vector<int32_t>& fv_offsets,
               std::vector<int32_t>& fv_indices);
// FeatureVector::score (MiniBow2.h:134-170)
float score(const int32_t* ids1, const float* v1, int n1, const int32_t* ids2, const float* v2, int n2);
}  // namespace bowref
