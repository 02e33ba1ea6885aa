/* ref_dbow_shim.cpp — C entry points around the reference's vendored DBoW2, compiled where it lies.  TEST INFRASTRUCTURE ONLY.
 *
 * oracle/Makefile (target _ref/libdbow_ref.so) compiles /root/reference/Thirdparty/DBoW2/DBoW2/{BowVector,FeatureVector,
 * ScoringObject,FORB}.cpp unmodified and instantiates TemplatedVocabulary<FORB::TDescriptor, FORB> (TemplatedVocabulary.h) — the
 * ORBVocabulary type of /root/reference/include/ORBVocabulary.h — against oracle/refstub/ (cv::Mat of cvstub.hpp, an empty
 * boost::serialization).  The vocabulary is read with the reference's own loadFromTextFile (TemplatedVocabulary.h:1338-1426) and
 * queried with transform(features, BowVector, FeatureVector, levelsup) (:1127-1193), the call of Frame::ComputeBoW
 * (src/Frame.cc:724-731).  tests/test_ref_pinned.py compares oracle/bow_oracle.cpp with it bit for bit (doubles included). */
#include <cstring>
#include <vector>

#include "refstub/cvstub.hpp"
#include "FORB.h"
#include "TemplatedVocabulary.h"

typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> ORBVocabulary;   /* include/ORBVocabulary.h:29-30 */

extern "C" {

void* ref_voc_load_text(const char* path) {
    ORBVocabulary* v = new ORBVocabulary();
    if (!v->loadFromTextFile(path)) { delete v; return nullptr; }
    return v;
}

void ref_voc_destroy(void* h) { delete (ORBVocabulary*)h; }

int ref_voc_words(void* h) { return (int)((ORBVocabulary*)h)->size(); }

/* transform(vCurrentDesc, mBowVec, mFeatVec, levelsup).  Outputs in map (key) order: bow_words / bow_values [*nbow];
 * fv_nodes [*nfv], fv_starts [*nfv + 1], fv_feats (concatenated feature indices).  Capacities: n entries each (+1 for starts). */
void ref_voc_transform(void* h, const uint8_t* desc, int n, int levelsup, int32_t* bow_words, double* bow_values, int32_t* nbow, int32_t* fv_nodes,
                       int32_t* fv_starts, int32_t* fv_feats, int32_t* nfv) {
    std::vector<cv::Mat> feats;
    feats.reserve(n);
    for (int i = 0; i < n; i++) {
        cv::Mat m(1, 32, CV_8U);
        std::memcpy(m.data, desc + (size_t)i * 32, 32);
        feats.push_back(m);
    }
    DBoW2::BowVector bv;
    DBoW2::FeatureVector fv;
    ((ORBVocabulary*)h)->transform(feats, bv, fv, levelsup);
    int k = 0;
    for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it, ++k) { bow_words[k] = (int32_t)it->first; bow_values[k] = it->second; }
    *nbow = k;
    int e = 0, pos = 0;
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++e) {
        fv_nodes[e] = (int32_t)it->first;
        fv_starts[e] = pos;
        for (unsigned f : it->second) fv_feats[pos++] = (int32_t)f;
    }
    fv_starts[e] = pos;
    *nfv = e;
}

/* ORBVocabulary::score(v1, v2) on two sparse vectors given in key order (KeyFrameDatabase / LoopClosing use it on mBowVec). */
double ref_voc_score(void* h, const int32_t* w1, const double* v1, int n1, const int32_t* w2, const double* v2, int n2) {
    DBoW2::BowVector a, b;
    for (int i = 0; i < n1; i++) a.insert(a.end(), std::make_pair((DBoW2::WordId)w1[i], v1[i]));
    for (int i = 0; i < n2; i++) b.insert(b.end(), std::make_pair((DBoW2::WordId)w2[i], v2[i]));
    return ((ORBVocabulary*)h)->score(a, b);
}

}  // extern "C"
