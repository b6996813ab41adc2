/**
 * B200 drop-in for the reference's include/ORBVocabulary.h:
 *   typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> ORBVocabulary;
 * as far as the front-end uses it — loadFromTextFile (src/System.cc:67-68,
 * Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1362-1450) and
 * transform(features, BowVector&, FeatureVector&, levelsup) (src/Frame.cc:439-449, src/KeyFrame.cc:60-71,
 * TemplatedVocabulary.h:1128-1207).  The tree descent of every descriptor runs on the GPU through the
 * C ABI (orb_bow_transform, include/orb_b200.h); BowVector / FeatureVector are assembled here in the
 * reference's order of double operations, so the weights are bit-equal.  No CPU fallback: without a
 * CUDA device loadFromTextFile throws.
 *
 * When the reference's DBoW2 headers (BowVector.h / FeatureVector.h) are included first, their types
 * are used; otherwise the minimal equivalents below stand in.
 */
#ifndef ORB_B200_ORBVOCABULARY_H
#define ORB_B200_ORBVOCABULARY_H

#include <cmath>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>
#include "orb_b200.h"
#include "orb_b200_cv.h"

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

namespace ORB_SLAM2 {

class ORBVocabulary {
public:
    explicit ORBVocabulary(int max_features = 1 << 16, int device = 0)
        : m_k(0), m_L(0), m_scoring(DBoW2::L1_NORM), m_weighting(DBoW2::TF_IDF), h_(NULL), max_features_(max_features), device_(device) {}
    ~ORBVocabulary() { if (h_) orb_vocabulary_destroy(h_); }

    // TemplatedVocabulary.h:1362-1450.  Returns false on a malformed header like the reference; empty lines
    // are skipped (the reference's eof() loop turns a trailing newline into a node with unspecified fields).
    bool loadFromTextFile(const std::string& filename) {
        std::ifstream f(filename.c_str());
        if (!f.is_open()) return false;
        std::string s;
        std::getline(f, s);
        std::stringstream ss(s);
        int n1 = -1, n2 = -1;
        m_k = m_L = -1;
        ss >> m_k >> m_L >> n1 >> n2;
        if (m_k < 0 || m_k > 20 || m_L < 1 || m_L > 10 || n1 < 0 || n1 > 5 || n2 < 0 || n2 > 3) return false;
        m_scoring = (DBoW2::ScoringType)n1; m_weighting = (DBoW2::WeightingType)n2;
        std::vector<int32_t> parent(1, 0); std::vector<uint8_t> leaf(1, 0), desc(32, 0);
        word_weight_.clear();
        while (std::getline(f, s)) {
            std::stringstream sn(s);
            int pid, isLeaf;
            if (!(sn >> pid >> isLeaf)) continue;
            parent.push_back(pid); leaf.push_back(isLeaf > 0 ? 1 : 0);
            for (int i = 0; i < 32; ++i) { int b = 0; sn >> b; desc.push_back((uint8_t)b); }
            double w = 0; sn >> w;
            if (isLeaf > 0) word_weight_.push_back(w);
        }
        if (h_) { orb_vocabulary_destroy(h_); h_ = NULL; }
        const int st = orb_vocabulary_create(m_L, (int)parent.size(), parent.data(), leaf.data(), desc.data(), max_features_, device_, &h_);
        if (st != ORB_OK) throw std::runtime_error(std::string("ORBVocabulary (B200): ") + orb_last_error());
        return true;
    }

    bool empty() const { return word_weight_.empty(); }
    unsigned int size() const { return (unsigned int)word_weight_.size(); }
    int getBranchingFactor() const { return m_k; }
    int getDepthLevels() const { return m_L; }
    DBoW2::WordValue getWordWeight(DBoW2::WordId id) const { return word_weight_[id]; }
    DBoW2::ScoringType getScoringType() const { return m_scoring; }
    DBoW2::WeightingType getWeightingType() const { return m_weighting; }

    // TemplatedVocabulary.h:1128-1207
    void transform(const std::vector<cv::Mat>& features, DBoW2::BowVector& v, DBoW2::FeatureVector& fv, int levelsup) const {
        v.clear(); fv.clear();
        if (empty()) return;
        const int n = (int)features.size();
        std::vector<uint8_t> rows((size_t)n * 32);
        for (int i = 0; i < n; ++i) std::memcpy(&rows[(size_t)i * 32], features[i].data, 32);
        std::vector<int32_t> word(n > 0 ? n : 1), node(n > 0 ? n : 1);
        const int st = orb_bow_transform(h_, rows.data(), n, levelsup, word.data(), node.data());
        if (st != ORB_OK) throw std::runtime_error(std::string("ORBVocabulary::transform (B200): ") + orb_last_error());
        // GeneralScoring::mustNormalize (ScoringObject.h:42-78): every scoring but the dot product normalises, L2 only for L2_NORM
        const bool must = m_scoring != DBoW2::DOT_PRODUCT;
        const DBoW2::LNorm norm = m_scoring == DBoW2::L2_NORM ? DBoW2::L2 : DBoW2::L1;
        const bool tf = m_weighting == DBoW2::TF || m_weighting == DBoW2::TF_IDF;
        for (int i = 0; i < n; ++i) {
            const DBoW2::WordValue w = word_weight_[word[i]];
            if (w > 0) {   // not stopped
                if (tf) v.addWeight((DBoW2::WordId)word[i], w); else v.addIfNotExist((DBoW2::WordId)word[i], w);
                fv.addFeature((DBoW2::NodeId)node[i], (unsigned int)i);
            }
        }
        if (tf && !v.empty() && !must) {
            const double nd = (double)v.size();
            for (DBoW2::BowVector::iterator vit = v.begin(); vit != v.end(); ++vit) vit->second /= nd;
        }
        if (must) v.normalize(norm);
    }

    orb_vocabulary* handle() const { return h_; }

private:
    ORBVocabulary(const ORBVocabulary&);
    ORBVocabulary& operator=(const ORBVocabulary&);
    int m_k, m_L;
    DBoW2::ScoringType m_scoring;
    DBoW2::WeightingType m_weighting;
    std::vector<double> word_weight_;   // by word id
    orb_vocabulary* h_;
    int max_features_, device_;
};

}  // namespace ORB_SLAM2

#endif
