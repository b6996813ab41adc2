/* TEST STAND-IN for the reference's Thirdparty/DBoW2/DBoW2/BowVector.h (same include path, same guard): lets tests/cpp compile
 * include/ORBVocabulary.h outside the reference tree.  Inside the reference tree the real DBoW2 header is found instead. */
#include <cmath>
#include <map>
#include <vector>
#ifndef __D_T_BOW_VECTOR__
#define __D_T_BOW_VECTOR__
namespace DBoW2 {
typedef unsigned int WordId;
typedef double WordValue;
typedef unsigned int NodeId;
enum LNorm { L1, L2 };
enum WeightingType { TF_IDF, TF, IDF, BINARY };
enum ScoringType { L1_NORM, L2_NORM, CHI_SQUARE, KL, BHATTACHARYYA, DOT_PRODUCT };
// Thirdparty/DBoW2/DBoW2/BowVector.h:56-104, BowVector.cpp:26-84
class BowVector : public std::map<WordId, WordValue> {
public:
    void addWeight(WordId id, WordValue v) {
        iterator vit = this->lower_bound(id);
        if (vit != this->end() && !(this->key_comp()(id, vit->first))) vit->second += v;
        else this->insert(vit, value_type(id, v));
    }
    void addIfNotExist(WordId id, WordValue v) {
        iterator vit = this->lower_bound(id);
        if (vit == this->end() || (this->key_comp()(id, vit->first))) this->insert(vit, value_type(id, v));
    }
    void normalize(LNorm norm_type) {
        double norm = 0.0;
        if (norm_type == L1) { for (iterator it = begin(); it != end(); ++it) norm += std::fabs(it->second); }
        else { for (iterator it = begin(); it != end(); ++it) norm += it->second * it->second; norm = std::sqrt(norm); }
        if (norm > 0.0) for (iterator it = begin(); it != end(); ++it) it->second /= norm;
    }
};
}  // namespace DBoW2
#endif
