"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes front-end of oracle/bow_oracle.cc (the BoW transform, N2: Frame::ComputeBoW src/Frame.cc:279-286 ->
TemplatedVocabulary::transform, Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1126-1259) and, when it was built in the
authoring container, of oracle/_ref/libref_dbow2.so = the unmodified reference DBoW2 (`RefVocabulary`).
Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import oracle as _O

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(_HERE, "_ref", "libref_dbow2.so")

SCORING = ("L1_NORM", "L2_NORM", "CHI_SQUARE", "KL", "BHATTACHARYYA", "DOT_PRODUCT")   # BowVector.h:45-53
WEIGHTING = ("TF_IDF", "TF", "IDF", "BINARY")                                          # BowVector.h:36-42

_P = C.c_void_p
_I = C.c_int


def _lib(fast=False):
    L = _O.lib(fast)
    if not getattr(L, "_bow_ready", False):
        L.bo_voc_load_text.restype = _P
        L.bo_voc_load_text.argtypes = [C.c_char_p]
        L.bo_voc_from_arrays.restype = _P
        L.bo_voc_from_arrays.argtypes = [_I, _I, _I, _I, _I, _P, _P, _P, _P]
        L.bo_voc_free.argtypes = [_P]
        L.bo_voc_info.argtypes = [_P, _P]
        L.bo_voc_arrays.argtypes = [_P, _P, _P, _P, _P]
        L.bo_descend.argtypes = [_P, _P, _I, _I, _P, _P, _P]
        L.bo_transform.argtypes = [_P, _P, _I, _I, _P, _P, _P, _P, _P, _P, _P]
        L.bo_score_l1.restype = C.c_double
        L.bo_score_l1.argtypes = [_P, _P, _I, _P, _P, _I]
        L._bow_ready = True
    return L


def _transform(fn, handle, desc, levelsup):
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    n = len(desc)
    bw = np.zeros(max(n, 1), np.int32); bv = np.zeros(max(n, 1), np.float64); nb = np.zeros(1, np.int32)
    fn_ = np.zeros(max(n, 1), np.int32); fs = np.zeros(n + 1, np.int32); fi = np.zeros(max(n, 1), np.int32)
    nn = np.zeros(1, np.int32)
    fn(handle, desc.ctypes.data, n, levelsup, bw.ctypes.data, bv.ctypes.data, nb.ctypes.data, fn_.ctypes.data,
       fs.ctypes.data, fi.ctypes.data, nn.ctypes.data)
    nb, nn = int(nb[0]), int(nn[0])
    return (bw[:nb].copy(), bv[:nb].copy()), (fn_[:nn].copy(), fs[:nn + 1].copy(), fi[:fs[nn]].copy())


class Vocabulary:
    """The restated vocabulary tree.  Node ids and word ids are numbered as loadFromTextFile numbers them."""

    def __init__(self, handle, fast=False):
        if not handle:
            raise ValueError("vocabulary could not be built (bad file / inconsistent tree)")
        self._L = _lib(fast)
        self._h = handle
        info = np.zeros(6, np.int32)
        self._L.bo_voc_info(handle, info.ctypes.data)
        self.k, self.L, self.scoring, self.weighting, self.nwords, self.nnodes = (int(v) for v in info)

    @classmethod
    def load_text(cls, path, fast=False):
        return cls(_lib(fast).bo_voc_load_text(os.fsencode(path)), fast)

    @classmethod
    def from_arrays(cls, k, L, scoring, weighting, parent, desc, weight, is_leaf, fast=False):
        parent = np.ascontiguousarray(parent, np.int32); desc = np.ascontiguousarray(desc, np.uint8)
        weight = np.ascontiguousarray(weight, np.float64); is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        return cls(_lib(fast).bo_voc_from_arrays(k, L, scoring, weighting, len(parent), parent.ctypes.data,
                                                 desc.ctypes.data, weight.ctypes.data, is_leaf.ctypes.data), fast)

    def arrays(self):
        """(parent[nnodes] i32, desc[nnodes,32] u8, weight[nnodes] f64, is_leaf[nnodes] u8)"""
        n = self.nnodes
        parent = np.zeros(n, np.int32); desc = np.zeros((n, 32), np.uint8); weight = np.zeros(n, np.float64)
        leaf = np.zeros(n, np.uint8)
        self._L.bo_voc_arrays(self._h, parent.ctypes.data, desc.ctypes.data, weight.ctypes.data, leaf.ctypes.data)
        return parent, desc, weight, leaf

    def descend(self, desc, levelsup=4):
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(desc)
        word = np.zeros(n, np.int32); w = np.zeros(n, np.float64); node = np.zeros(n, np.int32)
        self._L.bo_descend(self._h, desc.ctypes.data, n, levelsup, word.ctypes.data, w.ctypes.data, node.ctypes.data)
        return word, w, node

    def transform(self, desc, levelsup=4):
        """-> ((word_id, weight), (node_id, start, items)) in ascending id order"""
        return _transform(self._L.bo_transform, self._h, desc, levelsup)

    def __del__(self):
        try:
            self._L.bo_voc_free(self._h)
        except Exception:
            pass


def score_l1(b1, b2) -> float:
    w1 = np.ascontiguousarray(b1[0], np.int32); v1 = np.ascontiguousarray(b1[1], np.float64)
    w2 = np.ascontiguousarray(b2[0], np.int32); v2 = np.ascontiguousarray(b2[1], np.float64)
    return float(_lib().bo_score_l1(w1.ctypes.data, v1.ctypes.data, len(w1), w2.ctypes.data, v2.ctypes.data, len(w2)))


def kfdb_candidates(query_bow, db_bows, exclude=None):
    """The candidate scoring of KeyFrameDatabase::DetectRelocalisationCandidates / DetectLoopCandidates (reference
    src/KeyFrameDatabase.cc:198-255, :75-139) on flat BowVectors (word ids ascending, weights): shared-word count per
    keyframe (what walking the inverted file accumulates in mnRelocWords / mnLoopWords), maxCommonWords over the keyframes
    that are not excluded, minCommonWords = int(float(max) * 0.8f), and float(L1Scoring::score) for keyframes with
    more than minCommonWords shared words.  Returns (common, max_common, min_common, score[-1 where not scored])."""
    n = len(db_bows)
    qw = np.asarray(query_bow[0], np.int32)
    common = np.array([len(np.intersect1d(qw, np.asarray(b[0], np.int32), assume_unique=True)) for b in db_bows], np.int32)
    ex = np.zeros(n, bool) if exclude is None else np.asarray(exclude).astype(bool)
    mx = int(common[~ex].max()) if (~ex).any() else 0
    mn = int(np.float32(mx) * np.float32(0.8))
    score = np.full(n, -1, np.float32)
    for k in range(n):
        if not ex[k] and common[k] > mn:
            score[k] = np.float32(score_l1(query_bow, db_bows[k]))
    return common, mx, mn, score


def save_text(path, k, L, scoring, weighting, parent, desc, weight, is_leaf, trailing_newline=False):
    """Writes the ORBvoc.txt layout saveToTextFile produces (TemplatedVocabulary.h:1429-1449); weights with 17
    significant digits so a round trip is exact."""
    lines = [f"{k} {L}  {scoring} {weighting}"]
    for i in range(1, len(parent)):
        lines.append(f"{int(parent[i])} {int(is_leaf[i])} " + " ".join(str(int(b)) for b in desc[i]) +
                     f"  {float(weight[i]):.17g}")
    with open(path, "w") as f:
        f.write("\n".join(lines) + ("\n" if trailing_newline else ""))


def synth_vocabulary(k=10, L=3, seed=0, scoring=0, weighting=0, stop_frac=0.02, flip0=96, ragged=False):
    """A synthetic tree in loadFromTextFile numbering order is not required by the format (only parent-before-child);
    this one is laid out the way HKmeansStep numbers nodes: the k children of a node get consecutive ids, depth first
    (TemplatedVocabulary.h:640-800).  Child descriptors are the parent's with a shrinking number of random bit flips,
    so descents are not ties all the way down.  Leaf weights look like idf values; a fraction is 0 (stopped words).
    ragged=True prunes some subtrees so that leaves also appear above level L."""
    rng = np.random.default_rng(seed)
    parent, desc, leaf, level = [-1], [np.zeros(32, np.uint8)], [0], [0]

    def grow(node, lvl, bits):
        ids = []
        for _ in range(k):
            b = bits.copy()
            nflip = max(2, flip0 >> (lvl * 1))
            b[rng.choice(256, nflip, replace=False)] ^= 1
            parent.append(node); desc.append(np.packbits(b)); level.append(lvl + 1)
            is_leaf = lvl + 1 == L or (ragged and lvl + 1 >= 1 and rng.random() < 0.15)
            leaf.append(1 if is_leaf else 0)
            ids.append((len(parent) - 1, b, is_leaf))
            if not is_leaf:
                grow(len(parent) - 1, lvl + 1, b)
        return ids

    grow(0, 0, rng.integers(0, 2, 256).astype(np.uint8))
    parent = np.asarray(parent, np.int32); desc = np.stack(desc); leaf = np.asarray(leaf, np.uint8)
    weight = np.zeros(len(parent), np.float64)
    nl = int(leaf.sum())
    w = np.log(rng.uniform(1.5, 400.0, nl))
    w[rng.random(nl) < stop_frac] = 0.0
    weight[leaf == 1] = w
    return dict(k=k, L=L, scoring=scoring, weighting=weighting, parent=parent, desc=desc, weight=weight, is_leaf=leaf)


def synth_vocabulary_fast(k=10, L=6, seed=0, scoring=0, weighting=0):
    """Full k-ary tree of depth L built level by level with numpy (the bench's ORBvoc-sized tree: k=10, L=6 ->
    1,111,111 nodes), breadth-first numbering (parents precede children; children of a node are consecutive)."""
    rng = np.random.default_rng(seed)
    bits = rng.integers(0, 2, (1, 256)).astype(np.uint8)
    parents = [np.array([-1], np.int32)]; descs = [np.zeros((1, 32), np.uint8)]
    first = 0; count = 1
    for lvl in range(L):
        nch = count * k
        par = np.repeat(np.arange(first, first + count, dtype=np.int32), k)
        cb = np.repeat(bits, k, axis=0)
        nflip = max(2, 96 >> lvl)
        cols = rng.integers(0, 256, (nch, nflip))
        rows = np.repeat(np.arange(nch), nflip)
        np.bitwise_xor.at(cb, (rows, cols.ravel()), 1)
        parents.append(par); descs.append(np.packbits(cb, axis=1))
        first += count; count = nch; bits = cb
    parent = np.concatenate(parents); desc = np.concatenate(descs)
    leaf = np.zeros(len(parent), np.uint8); leaf[first:] = 1
    weight = np.zeros(len(parent), np.float64)
    weight[first:] = np.log(rng.uniform(1.5, 400.0, count))
    return dict(k=k, L=L, scoring=scoring, weighting=weighting, parent=parent, desc=desc, weight=weight, is_leaf=leaf)


# ---- the compiled reference (authoring container only; tests that need it skip when the .so is absent) -------------
def have_ref() -> bool:
    return os.path.exists(REF_LIB)


_ref = None


def _ref_lib():
    global _ref
    if _ref is None:
        L = C.CDLL(REF_LIB)
        L.ref_voc_load_text.restype = _P
        L.ref_voc_load_text.argtypes = [C.c_char_p]
        L.ref_voc_create.restype = _P
        L.ref_voc_create.argtypes = [_P, _I, _I, _I, _I, _I, _I]
        L.ref_voc_save_text.argtypes = [_P, C.c_char_p]
        L.ref_voc_free.argtypes = [_P]
        L.ref_voc_stop_words.restype = _I
        L.ref_voc_stop_words.argtypes = [_P, C.c_double]
        L.ref_voc_info.argtypes = [_P, _P]
        L.ref_voc_transform.argtypes = [_P, _P, _I, _I, _P, _P, _P, _P, _P, _P, _P]
        L.ref_voc_score.restype = C.c_double
        L.ref_voc_score.argtypes = [_P, _P, _P, _I, _P, _P, _I]
        L.ref_forb_distance.argtypes = [_P, _P, _I, _P]
        _ref = L
    return _ref


class RefVocabulary:
    """ORBVocabulary of the reference itself (include/ORBVocabulary.h:31-32)."""

    def __init__(self, handle):
        if not handle:
            raise ValueError("reference vocabulary load/create failed")
        self._L = _ref_lib()
        self._h = handle
        info = np.zeros(5, np.int32)
        self._L.ref_voc_info(handle, info.ctypes.data)
        self.k, self.L, self.scoring, self.weighting, self.nwords = (int(v) for v in info)

    @classmethod
    def load_text(cls, path):
        return cls(_ref_lib().ref_voc_load_text(os.fsencode(path)))

    @classmethod
    def create(cls, training, k, L, weighting=0, scoring=0):
        """training: [nimg, per, 32] uint8 -> TemplatedVocabulary::create (k-means++ seeds come from rand())."""
        t = np.ascontiguousarray(training, np.uint8)
        return cls(_ref_lib().ref_voc_create(t.ctypes.data, t.shape[0], t.shape[1], k, L, weighting, scoring))

    def save_text(self, path):
        self._L.ref_voc_save_text(self._h, os.fsencode(path))

    def stop_words(self, min_weight):
        return self._L.ref_voc_stop_words(self._h, float(min_weight))

    def transform(self, desc, levelsup=4):
        return _transform(self._L.ref_voc_transform, self._h, desc, levelsup)

    def score(self, b1, b2) -> float:
        w1 = np.ascontiguousarray(b1[0], np.int32); v1 = np.ascontiguousarray(b1[1], np.float64)
        w2 = np.ascontiguousarray(b2[0], np.int32); v2 = np.ascontiguousarray(b2[1], np.float64)
        return float(self._L.ref_voc_score(self._h, w1.ctypes.data, v1.ctypes.data, len(w1), w2.ctypes.data,
                                           v2.ctypes.data, len(w2)))

    def __del__(self):
        try:
            self._L.ref_voc_free(self._h)
        except Exception:
            pass


def ref_forb_distance(a, b):
    a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32); b = np.ascontiguousarray(b, np.uint8).reshape(-1, 32)
    out = np.zeros(len(a), np.int32)
    _ref_lib().ref_forb_distance(a.ctypes.data, b.ctypes.data, len(a), out.ctypes.data)
    return out
