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
 * score(a, b) (src/KeyFrameDatabase.cc:137,267, src/LoopClosing.cc:145) is the host-side merge of ScoringObject.cpp, chosen by
 * the file's scoring type.  BowVector / FeatureVector are the reference's own DBoW2 containers, included by the path the
 * reference uses (tests/cpp carries stand-ins under the same relative path for builds outside the reference tree).
 * One object may be used from several threads (the library serialises the staged transform per handle).
 */
#ifndef ORBVOCABULARY_H   /* the guard of the file this one replaces */
#define ORBVOCABULARY_H

#include <cfloat>
#include <cmath>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>
#include "orb_b200.h"
#include "orb_b200_cv.h"

#include "Thirdparty/DBoW2/DBoW2/BowVector.h"
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"

#ifdef ORB_B200_WITH_OPENCV
// Inside the reference tree: the header this file replaces pulled in DBoW2's TemplatedVocabulary.h, whose file-scope
// `using namespace std;` (TemplatedVocabulary.h:36) the reference's own headers rely on (Frame.h, Map.h, KeyFrameDatabase.h name
// vector / list / set unqualified).  Kept for source compatibility there, and only there.
using namespace std;
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

    // One feature: word id, weight and the node passed at level L - levelsup (TemplatedVocabulary.h:1239-1283)
    void transform(const cv::Mat& feature, DBoW2::WordId& id, DBoW2::WordValue& weight, DBoW2::NodeId* nid = NULL, int levelsup = 0) const {
        int32_t word = 0, node = 0;
        const int st = orb_bow_transform(h_, feature.data, 1, levelsup, &word, &node);
        if (st != ORB_OK) throw std::runtime_error(std::string("ORBVocabulary::transform (B200): ") + orb_last_error());
        id = (DBoW2::WordId)word; weight = word_weight_[word];
        if (nid) *nid = (DBoW2::NodeId)node;
    }
    DBoW2::WordId transform(const cv::Mat& feature) const {
        if (empty()) return 0;
        DBoW2::WordId id; DBoW2::WordValue w;
        transform(feature, id, w);
        return id;
    }

    // Similarity of two bag-of-words vectors under the vocabulary's scoring type (TemplatedVocabulary.h:1213-1218 ->
    // ScoringObject.cpp:21-313).  Terms are accumulated over the ids both vectors hold, in ascending id order, like the reference's
    // merge (which skips ahead with lower_bound: the visited pairs and their order are the same as a plain two-pointer walk).
    double score(const DBoW2::BowVector& a, const DBoW2::BowVector& b) const {
        DBoW2::BowVector::const_iterator ia = a.begin(), ib = b.begin();
        double acc = 0;
        const bool kl = m_scoring == DBoW2::KL;
        static const double kLogEps = std::log(DBL_EPSILON);   // GeneralScoring::LOG_EPS = log(DBL_EPSILON)
        while (ia != a.end() && ib != b.end()) {
            if (ia->first == ib->first) {
                const double vi = ia->second, wi = ib->second;
                switch (m_scoring) {
                    case DBoW2::L1_NORM: acc += std::fabs(vi - wi) - std::fabs(vi) - std::fabs(wi); break;
                    case DBoW2::L2_NORM: case DBoW2::DOT_PRODUCT: acc += vi * wi; break;
                    case DBoW2::CHI_SQUARE: if (vi + wi != 0.0) acc += vi * wi / (vi + wi); break;
                    case DBoW2::KL: if (vi != 0 && wi != 0) acc += vi * std::log(vi / wi); break;
                    case DBoW2::BHATTACHARYYA: acc += std::sqrt(vi * wi); break;
                }
                ++ia; ++ib;
            } else if (ia->first < ib->first) {
                if (kl) acc += ia->second * (std::log(ia->second) - kLogEps);   // a word only the first vector holds (:170-174)
                ++ia;
            } else {
                ++ib;
            }
        }
        switch (m_scoring) {
            case DBoW2::L1_NORM: return -acc / 2.0;
            case DBoW2::L2_NORM: return acc >= 1 ? 1.0 : 1.0 - std::sqrt(1.0 - acc);
            case DBoW2::CHI_SQUARE: return 2. * acc;
            case DBoW2::KL:
                for (; ia != a.end(); ++ia) if (ia->second != 0) acc += ia->second * (std::log(ia->second) - kLogEps);
                return acc;
            default: return acc;
        }
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
