/* TEST INFRASTRUCTURE — C wrapper around the reference's OWN vocabulary transform: the vendored DBoW2
 * (cslam/thirdparty/DBoW2/DBoW2/{TemplatedVocabulary.h, FORB.cpp, BowVector.cpp, FeatureVector.cpp, ScoringObject.cpp})
 * compiled UNMODIFIED over the mini cv::Mat of oracle/matchshim, instantiated exactly as cslam/include/cslam/ORBVocabulary.h
 * does. Frame::ComputeBoW (cslam/src/Frame.cpp:268-275) is one call: transform(descriptors, mBowVec, mFeatVec, 4).
 * The real ORBvoc.txt is a missing blob; tests load a small seeded vocabulary written in the same text format
 * (TemplatedVocabulary.h:1338-1420). Nothing here is shipped or called by the product path. */
#include <cslam/ORBVocabulary.h>

#include <cstring>

namespace {
/* the single-feature transform (TemplatedVocabulary.h:1218) is protected; no members are added */
struct VocabAccess : cslam::ORBVocabulary { using cslam::ORBVocabulary::transform; };
}

extern "C" {

void* rv_load(const char* path) {
    cslam::ORBVocabulary* v = new cslam::ORBVocabulary();
    if (!v->loadFromTextFile(path) || v->empty()) { delete v; return nullptr; }
    return v;
}
void rv_free(void* h) { delete (cslam::ORBVocabulary*)h; }
int rv_words(const void* h) { return (int)((const cslam::ORBVocabulary*)h)->size(); }

/* transform(features, BowVector&, FeatureVector&, levelsup), TemplatedVocabulary.h:1127-1197.
 * bow_word/bow_value [n]: the BowVector's entries in map order, *n_bow of them; fv_node [n]: node id of feature i in the
 * FeatureVector, -1 if the feature is in no node (stopped word); fv_order [n]: the features in FeatureVector iteration
 * order (node ascending, insertion order inside a node), *n_fv of them. */
void rv_transform(const void* h, int n, const uint8_t* desc, int levelsup, int* n_bow, unsigned* bow_word, double* bow_value,
                  int* fv_node, int* n_fv, int* fv_order) {
    const cslam::ORBVocabulary& V = *(const cslam::ORBVocabulary*)h;
    std::vector<cv::Mat> feats(n);
    for (int i = 0; i < n; i++) {                    /* Converter::toDescriptorVector: one 1x32 CV_8U row per feature */
        feats[i].create(1, 32, CV_8U);
        std::memcpy(feats[i].data, desc + (size_t)32 * i, 32);
    }
    DBoW2::BowVector bv;
    DBoW2::FeatureVector fv;
    V.transform(feats, bv, fv, levelsup);
    int k = 0;
    for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it, ++k) { bow_word[k] = it->first; bow_value[k] = it->second; }
    *n_bow = k;
    for (int i = 0; i < n; i++) fv_node[i] = -1;
    k = 0;
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it)
        for (size_t j = 0; j < it->second.size(); j++) { fv_node[it->second[j]] = (int)it->first; fv_order[k++] = (int)it->second[j]; }
    *n_fv = k;
}

/* word id, weight and the node `levelsup` levels above the leaf for single features, TemplatedVocabulary.h:1218-1261 */
void rv_transform_features(const void* h, int n, const uint8_t* desc, int levelsup, unsigned* word, double* weight, unsigned* node) {
    const VocabAccess& V = *static_cast<const VocabAccess*>((const cslam::ORBVocabulary*)h);
    cv::Mat f(1, 32, CV_8U);
    for (int i = 0; i < n; i++) {
        std::memcpy(f.data, desc + (size_t)32 * i, 32);
        DBoW2::WordId w; DBoW2::WordValue wt; DBoW2::NodeId nid = 0;
        V.transform(f, w, wt, &nid, levelsup);
        word[i] = w; weight[i] = wt; node[i] = nid;
    }
}

}  // extern "C"
