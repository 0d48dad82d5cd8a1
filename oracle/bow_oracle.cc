// ORACLE / test infrastructure (never linked into the product).  CPU restatement of the BoW transform the reference
// runs per frame (N2):
//   Frame::ComputeBoW                          src/Frame.cc:279-286   (transform(descriptors, BowVec, FeatVec, 4))
//   TemplatedVocabulary::transform (features)  Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1126-1194
//   TemplatedVocabulary::transform (1 feature) TemplatedVocabulary.h:1217-1259   (tree descent)
//   TemplatedVocabulary::loadFromTextFile      TemplatedVocabulary.h:1337-1424   (ORBvoc.txt format)
//   BowVector::addWeight/addIfNotExist/normalize  BowVector.cpp:34-85
//   FeatureVector::addFeature                  FeatureVector.cpp:31-47
//   FORB::distance                             FORB.cpp:81-101
//   score (L1)                                 ScoringObject.cpp L1Scoring::score
// PINNED against the reference itself: oracle/_ref/libref_dbow2.so is the unmodified DBoW2 compiled from
// /root/reference (oracle/Makefile target `ref`); tests/golden/bow_* hold a vocabulary the reference created and
// saved and the transforms it produced; tests/test_bow_oracle.py compares.
//
// Two places where the reference has undefined behaviour and this restatement picks a defined one:
//  * loadFromTextFile's `while(!f.eof())` turns the file's trailing newline into one more node whose parent id is an
//    uninitialised int.  Here (and in the product) a blank line ends the file.
//  * transform leaves *nid unset when the descent hits a leaf above level L - levelsup (only possible in a ragged
//    tree).  Here (and in the product) the leaf itself is reported.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

namespace {

struct Voc {
  int k = 0, L = 0, scoring = 0, weighting = 0, nwords = 0;
  std::vector<int32_t> parent;       // [nnodes], parent[0] = -1
  std::vector<uint8_t> desc;         // [nnodes][32] (root: zeros, never compared)
  std::vector<double> weight;        // [nnodes]
  std::vector<int32_t> word;         // [nnodes], -1 for inner nodes
  std::vector<std::vector<int32_t>> children;
};

inline int hamming(const uint8_t* a, const uint8_t* b) {
  int d = 0;
  for (int i = 0; i < 32; i++) d += __builtin_popcount((unsigned)(a[i] ^ b[i]));
  return d;
}

bool finish(Voc* v, const std::vector<uint8_t>& is_leaf) {
  const int n = (int)v->parent.size();
  v->children.assign(n, {});
  v->word.assign(n, -1);
  v->nwords = 0;
  for (int i = 1; i < n; i++) {
    const int p = v->parent[i];
    if (p < 0 || p >= i) return false;  // a parent always precedes its children in the file
    v->children[p].push_back(i);
    if (is_leaf[i]) v->word[i] = v->nwords++;
  }
  for (int i = 1; i < n; i++)
    if ((is_leaf[i] != 0) != v->children[i].empty()) return false;  // isLeaf() is "no children"
  return n > 1;
}

}  // namespace

extern "C" {

void* bo_voc_load_text(const char* path) {
  std::ifstream f(path);
  if (!f) return nullptr;
  std::string line;
  if (!std::getline(f, line)) return nullptr;
  Voc* v = new Voc();
  if (sscanf(line.c_str(), "%d %d %d %d", &v->k, &v->L, &v->scoring, &v->weighting) != 4 || v->k < 0 || v->k > 20 ||
      v->L < 1 || v->L > 10 || v->scoring < 0 || v->scoring > 5 || v->weighting < 0 || v->weighting > 3) {
    delete v;
    return nullptr;
  }
  v->parent.push_back(-1);
  v->desc.assign(32, 0);
  v->weight.push_back(0.0);
  std::vector<uint8_t> leaf(1, 0);
  while (std::getline(f, line)) {
    const char* p = line.c_str();
    char* e;
    long pid = strtol(p, &e, 10);
    if (e == p) break;  // blank line: end of file
    p = e;
    long lf = strtol(p, &e, 10);
    p = e;
    v->parent.push_back((int32_t)pid);
    leaf.push_back(lf > 0);
    for (int i = 0; i < 32; i++) {
      long b = strtol(p, &e, 10);
      p = e;
      v->desc.push_back((uint8_t)b);
    }
    v->weight.push_back(strtod(p, &e));
  }
  if (!finish(v, leaf)) { delete v; return nullptr; }
  return v;
}

void* bo_voc_from_arrays(int k, int L, int scoring, int weighting, int nnodes, const int32_t* parent,
                         const uint8_t* desc, const double* weight, const uint8_t* is_leaf) {
  Voc* v = new Voc();
  v->k = k; v->L = L; v->scoring = scoring; v->weighting = weighting;
  v->parent.assign(parent, parent + nnodes);
  v->desc.assign(desc, desc + (size_t)nnodes * 32);
  v->weight.assign(weight, weight + nnodes);
  std::vector<uint8_t> leaf(is_leaf, is_leaf + nnodes);
  if (!finish(v, leaf)) { delete v; return nullptr; }
  return v;
}

void bo_voc_free(void* voc) { delete static_cast<Voc*>(voc); }

// info = {k, L, scoring, weighting, nwords, nnodes}
void bo_voc_info(void* voc, int32_t info[6]) {
  Voc* v = static_cast<Voc*>(voc);
  info[0] = v->k; info[1] = v->L; info[2] = v->scoring; info[3] = v->weighting; info[4] = v->nwords;
  info[5] = (int32_t)v->parent.size();
}

void bo_voc_arrays(void* voc, int32_t* parent, uint8_t* desc, double* weight, uint8_t* is_leaf) {
  Voc* v = static_cast<Voc*>(voc);
  const size_t n = v->parent.size();
  memcpy(parent, v->parent.data(), n * 4);
  memcpy(desc, v->desc.data(), n * 32);
  memcpy(weight, v->weight.data(), n * 8);
  for (size_t i = 0; i < n; i++) is_leaf[i] = v->word[i] >= 0;
}

// one feature down the tree (TemplatedVocabulary.h:1217-1259): first strict minimum among the children, in child order
static void descend(const Voc* v, const uint8_t* d, int levelsup, int32_t* word, double* w, int32_t* nid) {
  const int nid_level = v->L - levelsup;
  int32_t node = 0, at = -1;
  if (nid_level <= 0) at = 0;
  int level = 0;
  do {
    ++level;
    const std::vector<int32_t>& ch = v->children[node];
    node = ch[0];
    int best = hamming(d, &v->desc[(size_t)node * 32]);
    for (size_t c = 1; c < ch.size(); c++) {
      const int dist = hamming(d, &v->desc[(size_t)ch[c] * 32]);
      if (dist < best) { best = dist; node = ch[c]; }
    }
    if (level == nid_level) at = node;
  } while (!v->children[node].empty());
  if (at < 0) at = node;  // ragged tree: leaf above nid_level (the reference leaves *nid unset)
  *word = v->word[node];
  *w = v->weight[node];
  *nid = at;
}

// per-feature results of the descent (word, weight, node at level L - levelsup)
void bo_descend(void* voc, const uint8_t* desc, int n, int levelsup, int32_t* word, double* weight, int32_t* node) {
  const Voc* v = static_cast<Voc*>(voc);
  for (int i = 0; i < n; i++) descend(v, desc + (size_t)i * 32, levelsup, &word[i], &weight[i], &node[i]);
}

// transform(features, BowVector, FeatureVector, levelsup), outputs flattened in std::map (ascending id) order:
//   bow_word / bow_weight [*n_bow];  fv_node [*n_nodes], fv_start [*n_nodes + 1], fv_items.
void bo_transform(void* voc, const uint8_t* desc, int n, int levelsup, int32_t* bow_word, double* bow_weight,
                  int32_t* n_bow, int32_t* fv_node, int32_t* fv_start, int32_t* fv_items, int32_t* n_nodes) {
  const Voc* v = static_cast<Voc*>(voc);
  const bool must = v->scoring != 5;           // DotProductScoring does not normalise (ScoringObject.h:72-87)
  const bool l2 = v->scoring == 1;
  const bool tf = v->weighting == 0 || v->weighting == 1;  // TF_IDF, TF accumulate; IDF, BINARY insert once
  // dense accumulators stand in for the two std::maps; entries are emitted in ascending id order
  std::vector<double> acc(v->nwords, 0.0);
  std::vector<uint8_t> seen(v->nwords, 0);
  std::vector<std::vector<int32_t>> per_node(v->parent.size());
  for (int i = 0; i < n; i++) {
    int32_t word, node;
    double w;
    descend(v, desc + (size_t)i * 32, levelsup, &word, &w, &node);
    if (w > 0) {  // not a stopped word
      if (!seen[word]) { seen[word] = 1; acc[word] = w; }
      else if (tf) acc[word] += w;
      per_node[node].push_back(i);
    }
  }
  int nb = 0;
  for (int wd = 0; wd < v->nwords; wd++)
    if (seen[wd]) { bow_word[nb] = wd; bow_weight[nb] = acc[wd]; nb++; }
  if (tf && nb && !must) {
    const double nd = (double)nb;
    for (int i = 0; i < nb; i++) bow_weight[i] /= nd;
  }
  if (must) {
    double norm = 0.0;
    if (!l2) { for (int i = 0; i < nb; i++) norm += std::fabs(bow_weight[i]); }
    else { for (int i = 0; i < nb; i++) norm += bow_weight[i] * bow_weight[i]; norm = std::sqrt(norm); }
    if (norm > 0.0) for (int i = 0; i < nb; i++) bow_weight[i] /= norm;
  }
  *n_bow = nb;
  int k = 0, pos = 0;
  for (size_t nd = 0; nd < per_node.size(); nd++) {
    if (per_node[nd].empty()) continue;
    fv_node[k] = (int32_t)nd;
    fv_start[k] = pos;
    for (int32_t it : per_node[nd]) fv_items[pos++] = it;
    k++;
  }
  fv_start[k] = pos;
  *n_nodes = k;
}

// L1Scoring::score (ScoringObject.cpp): merge of two ascending word lists;
//   score = -sum(|a - b| - |a| - |b|) / 2 over common words
double bo_score_l1(const int32_t* w1, const double* v1, int n1, const int32_t* w2, const double* v2, int n2) {
  double s = 0.0;
  int i = 0, j = 0;
  while (i < n1 && j < n2) {
    if (w1[i] == w2[j]) { s += std::fabs(v1[i] - v2[j]) - std::fabs(v1[i]) - std::fabs(v2[j]); i++; j++; }
    else if (w1[i] < w2[j]) i++;
    else j++;
  }
  return -s / 2.0;
}

}  // extern "C"
