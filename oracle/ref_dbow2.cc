// ORACLE / test infrastructure: C wrapper around the UNMODIFIED reference DBoW2 (compiled from the sources where they
// lie under /root/reference/Thirdparty/DBoW2, see oracle/Makefile target `ref`, output oracle/_ref/libref_dbow2.so).
// It exposes exactly what ORB_SLAM uses on the BoW path:
//   ORBVocabulary (include/ORBVocabulary.h:31-32) = TemplatedVocabulary<FORB::TDescriptor, FORB>
//   loadFromTextFile / saveToTextFile / create / transform(features, BowVector, FeatureVector, levelsup)
//   (Frame::ComputeBoW src/Frame.cc:279-286) and FORB::distance (== ORBmatcher::DescriptorDistance).
// OpenCV is replaced by oracle/ref_shim/opencv/cv.h (types only).  Used to pin oracle/bow_oracle.cc and to generate
// tests/golden/bow_*; never part of the product.
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "DBoW2/FORB.h"
#include "DBoW2/TemplatedVocabulary.h"

typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> ORBVocabulary;

static std::vector<cv::Mat> to_mats(const uint8_t* desc, int n) {
  std::vector<cv::Mat> v(n);
  for (int i = 0; i < n; i++) {
    v[i].create(1, 32, CV_8U);
    std::memcpy(v[i].ptr<unsigned char>(), desc + (size_t)i * 32, 32);
  }
  return v;
}

extern "C" {

void* ref_voc_load_text(const char* path) {
  ORBVocabulary* v = new ORBVocabulary();
  if (!v->loadFromTextFile(path)) { delete v; return nullptr; }
  return v;
}

// TemplatedVocabulary::create on `nimg` training images of `per` descriptors each, then saveToTextFile.
void* ref_voc_create(const uint8_t* desc, int nimg, int per, int k, int L, int weighting, int scoring) {
  std::vector<std::vector<cv::Mat>> feats(nimg);
  for (int i = 0; i < nimg; i++) feats[i] = to_mats(desc + (size_t)i * per * 32, per);
  ORBVocabulary* v = new ORBVocabulary(k, L, (DBoW2::WeightingType)weighting, (DBoW2::ScoringType)scoring);
  v->create(feats);
  return v;
}

void ref_voc_save_text(void* voc, const char* path) { static_cast<ORBVocabulary*>(voc)->saveToTextFile(path); }
void ref_voc_free(void* voc) { delete static_cast<ORBVocabulary*>(voc); }
int ref_voc_stop_words(void* voc, double min_weight) { return static_cast<ORBVocabulary*>(voc)->stopWords(min_weight); }

void ref_voc_info(void* voc, int32_t out[5]) {
  ORBVocabulary* v = static_cast<ORBVocabulary*>(voc);
  out[0] = v->getBranchingFactor(); out[1] = v->getDepthLevels(); out[2] = (int)v->getScoringType();
  out[3] = (int)v->getWeightingType(); out[4] = (int)v->size();
}

// transform(features, BowVector, FeatureVector, levelsup).  Outputs flattened in map order:
//   bow_word / bow_weight [*n_bow];  fv_node [*n_nodes], fv_start [*n_nodes + 1], fv_items [<= n].
void ref_voc_transform(void* voc, const uint8_t* desc, int n, int levelsup, int32_t* bow_word, double* bow_weight,
                       int32_t* n_bow, int32_t* fv_node, int32_t* fv_start, int32_t* fv_items, int32_t* n_nodes) {
  ORBVocabulary* v = static_cast<ORBVocabulary*>(voc);
  std::vector<cv::Mat> feats = to_mats(desc, n);
  DBoW2::BowVector bv;
  DBoW2::FeatureVector fv;
  v->transform(feats, bv, fv, levelsup);
  int i = 0;
  for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it, ++i) {
    bow_word[i] = (int32_t)it->first;
    bow_weight[i] = it->second;
  }
  *n_bow = i;
  int k = 0, pos = 0;
  for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++k) {
    fv_node[k] = (int32_t)it->first;
    fv_start[k] = pos;
    for (size_t j = 0; j < it->second.size(); j++) fv_items[pos++] = (int32_t)it->second[j];
  }
  fv_start[k] = pos;
  *n_nodes = k;
}

double ref_voc_score(void* voc, const int32_t* w1, const double* v1, int n1, const int32_t* w2, const double* v2,
                     int n2) {
  DBoW2::BowVector a, b;
  for (int i = 0; i < n1; i++) a.insert(a.end(), DBoW2::BowVector::value_type(w1[i], v1[i]));
  for (int i = 0; i < n2; i++) b.insert(b.end(), DBoW2::BowVector::value_type(w2[i], v2[i]));
  return static_cast<ORBVocabulary*>(voc)->score(a, b);
}

// FORB::distance on n independent pairs (the code ORBmatcher::DescriptorDistance repeats, src/ORBmatcher.cc:1794-1810)
void ref_forb_distance(const uint8_t* a, const uint8_t* b, int n, int32_t* out) {
  cv::Mat ma(1, 32, CV_8U), mb(1, 32, CV_8U);
  for (int i = 0; i < n; i++) {
    std::memcpy(ma.ptr<unsigned char>(), a + (size_t)i * 32, 32);
    std::memcpy(mb.ptr<unsigned char>(), b + (size_t)i * 32, 32);
    out[i] = DBoW2::FORB::distance(ma, mb);
  }
}

}  // extern "C"
