/*
 * ORBVocabulary.h -- host-side mirror of ORB_SLAM2::ORBVocabulary (include/ORBVocabulary.h:31-32 =
 * DBoW2::TemplatedVocabulary<cv::Mat, FORB>) for the calls the tracking / mapping / loop-closing threads make:
 *
 *   loadFromTextFile(file)                    src/System.cc:70                 TemplatedVocabulary.h:1358-1443
 *   transform(features, BowVec, FeatVec, 4)   src/Frame.cc:459-466, src/KeyFrame.cc:59-69   TemplatedVocabulary.h:1127-1190
 *   score(BowVec, BowVec)                     src/LoopClosing.cc:157, src/KeyFrameDatabase.cc:159,288   ScoringObject.cpp
 *   size() / empty() / getBranchingFactor() / getDepthLevels()
 *
 * The tree lives in HBM behind liborbb200.so (include/orbb200.h, "DBoW2 vocabulary transform"); transform() is one
 * device call for all features of the frame, transformBatch() one for many frames.  score() walks two host std::maps
 * and stays on the host, term by term as ScoringObject.cpp writes it.  No CPU path for the transform: a failing device
 * call throws std::runtime_error with orbb200_last_error().
 *
 * Drop-in: `typedef orbb200::ORBVocabulary ORBVocabulary;` inside namespace ORB_SLAM2.  With the
 * real DBoW2 headers on the include path define ORBB200_USE_DBOW2 so that BowVector / FeatureVector are DBoW2's own
 * types; otherwise the two std::map subclasses below stand in for them (same key / value types, BowVector.h:56-57,
 * FeatureVector.h:23-24).
 */
#ifndef ORBB200_ORBVOCABULARY_H
#define ORBB200_ORBVOCABULARY_H

#include <cfloat>
#include <cmath>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "../orbb200.h"
#ifdef ORBB200_USE_OPENCV
#include <opencv2/core/core.hpp>
#else
#include "cv_shim.h"
#endif
#ifdef ORBB200_USE_DBOW2
#include "Thirdparty/DBoW2/DBoW2/BowVector.h"
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"
#else
namespace DBoW2 {
typedef unsigned int WordId;
typedef double WordValue;
typedef unsigned int NodeId;
enum WeightingType { TF_IDF, TF, IDF, BINARY };
enum ScoringType { L1_NORM, L2_NORM, CHI_SQUARE, KL, BHATTACHARYYA, DOT_PRODUCT };
class BowVector : public std::map<WordId, WordValue> {};
class FeatureVector : public std::map<NodeId, std::vector<unsigned int> > {};
} // namespace DBoW2
#endif

namespace orbb200 {

class ORBVocabulary {
public:
    explicit ORBVocabulary(int device = -1) : v_(0), device_(device), k_(0), L_(0), scoring_(0), weighting_(0), nodes_(0), words_(0) {}
    ~ORBVocabulary() { orbb200_vocabulary_destroy(v_); }

    /* TemplatedVocabulary::loadFromTextFile: false on an unreadable / malformed file (System.cc:71-76 then exits) */
    bool loadFromTextFile(const std::string &filename)
    {
        orbb200_vocabulary_destroy(v_);
        v_ = 0;
        words_ = nodes_ = 0;
        const int rc = orbb200_vocabulary_load_text(&v_, filename.c_str(), device_);
        if (rc == ORBB200_ERR_CUDA) fail("loadFromTextFile");
        if (rc != ORBB200_OK) return false;
        orbb200_vocabulary_info(v_, &k_, &L_, &scoring_, &weighting_, &nodes_, &words_);
        return true;
    }
    /* the same tree from the loader's arrays (orbb200_vocabulary_create) */
    void create(int k, int L, int scoring, int weighting, int nnodes, const int *parent, const unsigned char *is_leaf,
                const unsigned char *desc, const double *weight)
    {
        orbb200_vocabulary_destroy(v_);
        v_ = 0;
        if (orbb200_vocabulary_create(&v_, k, L, scoring, weighting, nnodes, parent, is_leaf, desc, weight, device_) != ORBB200_OK)
            fail("create");
        orbb200_vocabulary_info(v_, &k_, &L_, &scoring_, &weighting_, &nodes_, &words_);
    }

    unsigned int size() const { return (unsigned int)words_; }  /* m_words.size() (:83) */
    bool empty() const { return words_ == 0; }                  /* :89 */
    int getBranchingFactor() const { return k_; }               /* :146 */
    int getDepthLevels() const { return L_; }                   /* :152 */
    DBoW2::WeightingType getWeightingType() const { return (DBoW2::WeightingType)weighting_; }
    DBoW2::ScoringType getScoringType() const { return (DBoW2::ScoringType)scoring_; }
    orbb200_vocabulary *handle() const { return v_; }

    /* transform(features, v, fv, levelsup) with the reference's argument: one 1x32 row header per feature
     * (Converter::toDescriptorVector, src/Converter.cc:27-35) */
    void transform(const std::vector<cv::Mat> &features, DBoW2::BowVector &v, DBoW2::FeatureVector &fv, int levelsup) const
    {
        std::vector<unsigned char> rows(features.size() * 32);
        for (size_t i = 0; i < features.size(); i++) std::memcpy(&rows[i * 32], features[i].ptr<unsigned char>(), 32);
        std::vector<DBoW2::BowVector> bv(1);
        std::vector<DBoW2::FeatureVector> fvs(1);
        const int n = (int)features.size();
        run(rows.empty() ? 0 : &rows[0], &n, 1, levelsup, bv, fvs);
        v.swap(bv[0]);
        fv.swap(fvs[0]);
    }
    /* the same on the N x 32 descriptor matrix itself (Frame::mDescriptors), skipping the per-row headers */
    void transform(const cv::Mat &descriptors, DBoW2::BowVector &v, DBoW2::FeatureVector &fv, int levelsup) const
    {
        std::vector<cv::Mat> one(1, descriptors);
        std::vector<DBoW2::BowVector> bv;
        std::vector<DBoW2::FeatureVector> fvs;
        transformBatch(one, bv, fvs, levelsup);
        v.swap(bv[0]);
        fv.swap(fvs[0]);
    }
    /* Frame::ComputeBoW for many frames in one device call: descriptors[i] is frame i's N_i x 32 matrix */
    void transformBatch(const std::vector<cv::Mat> &descriptors, std::vector<DBoW2::BowVector> &v,
                        std::vector<DBoW2::FeatureVector> &fv, int levelsup) const
    {
        std::vector<int> counts(descriptors.size());
        size_t total = 0;
        for (size_t i = 0; i < descriptors.size(); i++) total += (size_t)(counts[i] = descriptors[i].rows);
        std::vector<unsigned char> rows(total * 32);
        size_t o = 0;
        for (size_t i = 0; i < descriptors.size(); i++)
            for (int r = 0; r < descriptors[i].rows; r++, o += 32) std::memcpy(&rows[o], descriptors[i].ptr<unsigned char>(r), 32);
        v.assign(descriptors.size(), DBoW2::BowVector());
        fv.assign(descriptors.size(), DBoW2::FeatureVector());
        run(rows.empty() ? 0 : &rows[0], counts.empty() ? 0 : &counts[0], (int)counts.size(), levelsup, v, fv);
    }

    /* TemplatedVocabulary::score (:1194-1198) -> the scoring object's score(), ScoringObject.cpp:23-315 */
    double score(const DBoW2::BowVector &v1, const DBoW2::BowVector &v2) const { return score(scoring_, v1, v2); }
    /* the six scoring classes by their ScoringType value (host arithmetic only, usable without a device) */
    static double score(int scoring_, const DBoW2::BowVector &v1, const DBoW2::BowVector &v2)
    {
        typedef DBoW2::BowVector::const_iterator It;
        It a = v1.begin(), b = v2.begin();
        const It ae = v1.end(), be = v2.end();
        static const double LOG_EPS = std::log(DBL_EPSILON); /* ScoringObject.cpp:18 */
        double score = 0;
        while (a != ae && b != be) {
            const double vi = a->second, wi = b->second;
            if (a->first == b->first) {
                switch (scoring_) {
                case 0: score += std::fabs(vi - wi) - std::fabs(vi) - std::fabs(wi); break; /* L1 :40 */
                case 1: score += vi * wi; break;                                            /* L2 :91 */
                case 2: if (vi + wi != 0.0) score += vi * wi / (vi + wi); break;            /* chi square :150 */
                case 3: if (vi != 0 && wi != 0) score += vi * std::log(vi / wi); break;     /* KL :195 */
                case 4: score += std::sqrt(vi * wi); break;                                 /* Bhattacharyya :246 */
                default: score += vi * wi; break;                                           /* dot product :291 */
                }
                ++a;
                ++b;
            } else if (a->first < b->first) {
                if (scoring_ == 3) { /* KL counts the words only v1 has (:202-206) */
                    score += vi * (std::log(vi) - LOG_EPS);
                    ++a;
                } else {
                    a = v1.lower_bound(b->first);
                }
            } else {
                b = v2.lower_bound(a->first);
            }
        }
        switch (scoring_) {
        case 0: return -score / 2.0;                                      /* :63 */
        case 1: return score >= 1 ? 1.0 : 1.0 - std::sqrt(1.0 - score);   /* :117-120 */
        case 2: return 2. * score;                                        /* :172 */
        case 3:
            for (; a != ae; ++a)
                if (a->second != 0) score += a->second * (std::log(a->second) - LOG_EPS); /* :216-218 */
            return score;
        default: return score;
        }
    }

private:
    ORBVocabulary(const ORBVocabulary &);
    ORBVocabulary &operator=(const ORBVocabulary &);

    static void fail(const char *what) { throw std::runtime_error(std::string("orbb200::ORBVocabulary::") + what + ": " + orbb200_last_error()); }

    void run(const unsigned char *rows, const int *counts, int nimg, int levelsup, std::vector<DBoW2::BowVector> &v,
             std::vector<DBoW2::FeatureVector> &fv) const
    {
        if (!v_ || nimg == 0) return; /* if(empty()) return; leaves both maps cleared */
        size_t total = 0;
        for (int i = 0; i < nimg; i++) total += (size_t)counts[i];
        std::vector<int> bowOff(nimg + 1), fvOff(nimg + 1), bowId(total + 1), fvNode(total + 1), fvFeatOff(total + 2), fvFeat(total + 1);
        std::vector<double> bowVal(total + 1);
        if (orbb200_compute_bow_batch(v_, rows, counts, nimg, levelsup, &bowOff[0], &bowId[0], &bowVal[0], &fvOff[0], &fvNode[0],
                                      &fvFeatOff[0], &fvFeat[0]) != ORBB200_OK)
            fail("transform");
        for (int i = 0; i < nimg; i++) { /* already in key order: hinted inserts are O(1) each */
            for (int j = bowOff[i]; j < bowOff[i + 1]; j++)
                v[i].insert(v[i].end(), DBoW2::BowVector::value_type((DBoW2::WordId)bowId[j], bowVal[j]));
            for (int j = fvOff[i]; j < fvOff[i + 1]; j++) {
                DBoW2::FeatureVector::iterator it =
                    fv[i].insert(fv[i].end(), DBoW2::FeatureVector::value_type((DBoW2::NodeId)fvNode[j], std::vector<unsigned int>()));
                it->second.assign(fvFeat.begin() + fvFeatOff[j], fvFeat.begin() + fvFeatOff[j + 1]);
            }
        }
    }

    orbb200_vocabulary *v_;
    int device_, k_, L_, scoring_, weighting_, nodes_, words_;
};

} // namespace orbb200

#endif /* ORBB200_ORBVOCABULARY_H */
