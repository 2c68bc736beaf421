// dbow2_ref.cpp -- TEST INFRASTRUCTURE.  C entry points around the reference's OWN DBoW2 vocabulary code (compiled from
// /root/reference/Thirdparty/DBoW2 where it lies, see oracle/Makefile) so that tests can pin oracle/vocab_oracle.py and
// the CUDA transform against it, and bench.py --impl reference can time it.  Nothing here restates an algorithm: every
// call lands in TemplatedVocabulary<cv::Mat, FORB> (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h).
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "DBoW2/FORB.h"
#include "DBoW2/TemplatedVocabulary.h"

// yml persistence is outside the path: cv::FileStorage is an aborting stub in oracle/cv_shim_impl.cpp

namespace {
typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> Voc; // include/ORBVocabulary.h:31-32
struct OpenVoc : public Voc { // the per-feature transform is protected (TemplatedVocabulary.h:297-298)
    OpenVoc(int k, int L, DBoW2::WeightingType w, DBoW2::ScoringType s) : Voc(k, L, w, s) {}
    void one(const cv::Mat &f, DBoW2::WordId &id, DBoW2::WordValue &w, DBoW2::NodeId *nid, int levelsup) const
    {
        transform(f, id, w, nid, levelsup);
    }
    int nodes() const { return (int)m_nodes.size(); }
    int scoring() const { return (int)m_scoring; }
    int weighting() const { return (int)m_weighting; }
};
std::vector<cv::Mat> rows_of(const uint8_t *desc, int n) // Converter::toDescriptorVector (src/Converter.cc:27-35)
{
    std::vector<cv::Mat> v;
    v.reserve(n);
    for (int i = 0; i < n; i++) v.push_back(cv::Mat(1, 32, CV_8U, (void *)(desc + (size_t)i * 32)));
    return v;
}
} // namespace

extern "C" {
void *dbow2_ref_new(int k, int L, int weighting, int scoring)
{
    return new OpenVoc(k, L, (DBoW2::WeightingType)weighting, (DBoW2::ScoringType)scoring);
}
void dbow2_ref_free(void *v) { delete (OpenVoc *)v; }
int dbow2_ref_load_text(void *v, const char *path) { return ((OpenVoc *)v)->loadFromTextFile(path) ? 0 : -1; }
void dbow2_ref_save_text(void *v, const char *path) { ((OpenVoc *)v)->saveToTextFile(path); }
// HKmeans training with the reference's create() on `nimg` images of counts[i] descriptors each
void dbow2_ref_create(void *v, const uint8_t *desc, const int *counts, int nimg, int seed)
{
    DUtils::Random::SeedRandOnce(seed);
    std::vector<std::vector<cv::Mat> > tf(nimg);
    size_t o = 0;
    for (int i = 0; i < nimg; i++) {
        for (int j = 0; j < counts[i]; j++, o++) tf[i].push_back(cv::Mat(1, 32, CV_8U, (void *)(desc + o * 32)).clone());
    }
    ((OpenVoc *)v)->create(tf);
}
void dbow2_ref_info(void *v, int *k, int *L, int *nodes, int *words, int *scoring, int *weighting)
{
    OpenVoc *o = (OpenVoc *)v;
    *k = o->getBranchingFactor(); *L = o->getDepthLevels(); *nodes = o->nodes(); *words = (int)o->size();
    *scoring = o->scoring(); *weighting = o->weighting();
}
// per-feature descent (TemplatedVocabulary.h:1218-1259); nid is preset to -1 so that "never written" is visible
void dbow2_ref_transform_features(void *v, const uint8_t *desc, int n, int levelsup, int *word, double *weight, int *nid)
{
    OpenVoc *o = (OpenVoc *)v;
    for (int i = 0; i < n; i++) {
        cv::Mat f(1, 32, CV_8U, (void *)(desc + (size_t)i * 32));
        DBoW2::WordId id = 0; DBoW2::WordValue w = 0; DBoW2::NodeId node = (DBoW2::NodeId)-1;
        o->one(f, id, w, &node, levelsup);
        word[i] = (int)id; weight[i] = w; nid[i] = (int)node;
    }
}
// Frame::ComputeBoW (src/Frame.cc:459-466): transform(features, BowVector, FeatureVector, levelsup).  Flattened outputs:
// bow_id/bow_val[<= n] in map order; fv_node[<= n] in map order with fv_off[nfv+1] into fv_feat[<= n].
void dbow2_ref_transform(void *v, const uint8_t *desc, int n, int levelsup, int *nbow, int *bow_id, double *bow_val, int *nfv,
                         int *fv_node, int *fv_off, int *fv_feat)
{
    DBoW2::BowVector bv; DBoW2::FeatureVector fv;
    std::vector<cv::Mat> d = rows_of(desc, n);
    ((OpenVoc *)v)->transform(d, bv, fv, levelsup);
    int i = 0;
    for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it, ++i) { bow_id[i] = (int)it->first; bow_val[i] = it->second; }
    *nbow = i;
    int j = 0, o = 0;
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++j) {
        fv_node[j] = (int)it->first; fv_off[j] = o;
        for (size_t q = 0; q < it->second.size(); q++) fv_feat[o++] = (int)it->second[q];
    }
    fv_off[j] = o; *nfv = j;
}
double dbow2_ref_score(void *v, int na, const int *ida, const double *va, int nb, const int *idb, const double *vb)
{
    DBoW2::BowVector a, b;
    for (int i = 0; i < na; i++) a.addWeight(ida[i], va[i]);
    for (int i = 0; i < nb; i++) b.addWeight(idb[i], vb[i]);
    return ((OpenVoc *)v)->score(a, b);
}
}
