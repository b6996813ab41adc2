/* TEST STAND-IN for the reference's Thirdparty/DBoW2/DBoW2/FeatureVector.h (same include path, same guard): lets tests/cpp compile
 * include/ORBVocabulary.h outside the reference tree.  Inside the reference tree the real DBoW2 header is found instead. */
#include <cmath>
#include <map>
#include <vector>
#ifndef __D_T_FEATURE_VECTOR__
#define __D_T_FEATURE_VECTOR__
namespace DBoW2 {
// Thirdparty/DBoW2/DBoW2/FeatureVector.h:22-49, FeatureVector.cpp:31-47
class FeatureVector : public std::map<NodeId, std::vector<unsigned int> > {
public:
    void addFeature(NodeId id, unsigned int i_feature) {
        iterator vit = this->lower_bound(id);
        if (vit != this->end() && vit->first == id) vit->second.push_back(i_feature);
        else { vit = this->insert(vit, value_type(id, std::vector<unsigned int>())); vit->second.push_back(i_feature); }
    }
};
}  // namespace DBoW2
#endif
