// TEST INFRASTRUCTURE — CPU restatement of the vocabulary-tree descent of
// DBoW2::TemplatedVocabulary<FORB>::transform(feature, word_id, weight, nid, levelsup)
// (reference Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1239-1283) with the node numbering of
// loadFromTextFile (:1362-1450).  Pinned against the reference's own DBoW2 compiled verbatim
// (oracle/_ref/ref_match_driver bowtransform, tests/test_oracle_refmatch.py).
#include <cstdint>
#include <vector>
#include "orb_oracle.h"

extern "C" void orc_bow_transform(int levels, int n_nodes, const int32_t* parent, const uint8_t* is_leaf, const uint8_t* node_desc,
                                  const uint8_t* desc, int n, int levelsup, int32_t* word_id, int32_t* node_id) {
    std::vector<std::vector<int>> children(n_nodes);
    std::vector<int> word(n_nodes, -1);
    int n_words = 0;
    for (int i = 1; i < n_nodes; ++i) { children[parent[i]].push_back(i); if (is_leaf[i]) word[i] = n_words++; }
    const int nid_level = levels - levelsup;
    for (int f = 0; f < n; ++f) {
        const uint8_t* q = desc + (size_t)f * 32;
        int final_id = 0, current_level = 0, nid = -1;
        if (nid_level <= 0) nid = 0;                                   // :1251
        do {
            ++current_level;
            const std::vector<int>& nodes = children[final_id];
            final_id = nodes[0];
            int best_d = orc_descriptor_distance(q, node_desc + (size_t)final_id * 32);
            for (size_t k = 1; k < nodes.size(); ++k) {
                const int d = orc_descriptor_distance(q, node_desc + (size_t)nodes[k] * 32);
                if (d < best_d) { best_d = d; final_id = nodes[k]; }    // :1270-1274
            }
            if (current_level == nid_level) nid = final_id;             // :1277-1278
        } while (!children[final_id].empty());                          // isLeaf(): no children
        word_id[f] = word[final_id];
        node_id[f] = nid < 0 ? final_id : nid;   // the reference leaves *nid unset when a leaf sits above that level; convention: the leaf
    }
}
