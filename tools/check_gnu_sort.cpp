// Validates snake-slam_b200/csrc/gnu_sort.cuh against the running libstdc++ std::sort on tie-heavy inputs.
// Build: g++ -O2 -I snake-slam_b200/csrc tools/check_gnu_sort.cpp -o /tmp/chk_sort
#include "gnu_sort.cuh"

#include <algorithm>
#include <cstdio>
#include <random>
#include <vector>

struct Node
{
    float a, b, c, d;
    int from, to;
    int id;
    int count() const { return to - from; }
};

int main()
{
    std::mt19937 rng(7);
    long long cases = 0, bad = 0;
    for (int rep = 0; rep < 300000; ++rep)
    {
        int n       = 1 + rng() % (rep % 50 == 0 ? 4000 : 600);
        int maxc    = 2 + rng() % (rep % 3 == 0 ? 4 : (rep % 3 == 1 ? 40 : 3000));
        int pattern = rng() % 4;
        std::vector<Node> nodes(n);
        std::vector<uint32_t> keys(n);
        for (int i = 0; i < n; ++i)
        {
            int c = 2 + rng() % maxc;
            if (pattern == 1) c = 2 + (i * maxc) / n;            // ascending
            if (pattern == 2) c = 2 + ((n - i) * maxc) / n;      // descending
            if (pattern == 3 && (rng() % 8)) c = 2;              // mostly equal
            nodes[i] = {0, 0, 0, 0, 0, c, i};
            keys[i]  = ((uint32_t)c << gnusort::SHIFT) | (uint32_t)i;
        }
        std::sort(nodes.begin(), nodes.end(), [](const Node& x, const Node& y) { return x.count() > y.count(); });
        gnusort::sort(keys.data(), n);
        ++cases;
        for (int i = 0; i < n; ++i)
            if ((int)(keys[i] & 0xfff) != nodes[i].id)
            {
                ++bad;
                if (bad < 5) printf("mismatch n=%d maxc=%d pattern=%d at %d\n", n, maxc, pattern, i);
                break;
            }
    }
    // adversarial: median-of-3 killer-like sequences to hit the heapsort fallback
    for (int n = 17; n < 3000; n += 7)
    {
        std::vector<Node> nodes(n);
        std::vector<uint32_t> keys(n);
        for (int i = 0; i < n; ++i)
        {
            int c    = (i % 2 == 0) ? (i / 2 + 2) : (n - i / 2 + 2);
            nodes[i] = {0, 0, 0, 0, 0, c, i};
            keys[i]  = ((uint32_t)c << gnusort::SHIFT) | (uint32_t)i;
        }
        std::sort(nodes.begin(), nodes.end(), [](const Node& x, const Node& y) { return x.count() > y.count(); });
        gnusort::sort(keys.data(), n);
        ++cases;
        for (int i = 0; i < n; ++i)
            if ((int)(keys[i] & 0xfff) != nodes[i].id)
            {
                ++bad;
                break;
            }
    }
    // MiniBow2 FeatureVector::setFeatures: std::sort of (node id, feature index) pairs on the node id alone, ids
    // from a small set (k^(L-levelsup) nodes) plus -1 entries — 64-bit keys, ascending signed high word
    for (int rep = 0; rep < 60000; ++rep)
    {
        int n     = 1 + rng() % (rep % 20 == 0 ? 8000 : 2500);
        int nodes = 1 + rng() % (rep % 2 ? 100 : 1000);
        std::vector<std::pair<int, int>> pairs(n);
        std::vector<uint64_t> keys(n);
        for (int i = 0; i < n; ++i)
        {
            int id   = (rng() % 16 == 0) ? -1 : (int)(rng() % nodes) * 7 + 3;
            pairs[i] = {id, i};
            keys[i]  = ((uint64_t)(uint32_t)id << 32) | (uint32_t)i;
        }
        std::sort(pairs.begin(), pairs.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
        gnusort::sort_as<uint64_t, gnusort::AscHigh32>(keys.data(), n);
        ++cases;
        for (int i = 0; i < n; ++i)
            if ((int)(uint32_t)keys[i] != pairs[i].second || (int32_t)(keys[i] >> 32) != pairs[i].first)
            {
                ++bad;
                if (bad < 5) printf("pair mismatch n=%d at %d\n", n, i);
                break;
            }
    }
    printf("%lld cases, %lld mismatches\n", cases, bad);
    return bad != 0;
}
