"""Second, independent restatement (plain Python, numpy float32 scalars) of the matcher routines, written from the
reference source separately from oracle/match_oracle.cc.  The reference ships no tests and cannot be
compiled here (§8c); agreement of two independent restatements on random inputs is what pins the C++ oracle."""
import math

import numpy as np

F = np.float32
TH_HIGH, TH_LOW, HISTO = 100, 50, 30
INT_MAX = 2**31 - 1


def dist(a, b):  # src/ORBmatcher.cc:1794-1810
    return int(np.unpackbits(np.bitwise_xor(a, b)).sum())


def c_round(v):  # C round(): half away from zero
    v = float(v)
    return int(math.floor(v + 0.5)) if v >= 0 else int(math.ceil(v - 0.5))


def inv_wh(f):
    return F(64) / F(F(f.max_x) - F(f.min_x)), F(48) / F(F(f.max_y) - F(f.min_y))


def build_grid(f):  # src/Frame.cc:108-122, 266-276
    iw, ih = inv_wh(f)
    grid = [[[] for _ in range(48)] for _ in range(64)]
    for i in range(len(f.x)):
        px = c_round(F(F(f.x[i]) - F(f.min_x)) * iw)
        py = c_round(F(F(f.y[i]) - F(f.min_y)) * ih)
        if 0 <= px < 64 and 0 <= py < 48:
            grid[px][py].append(i)
    return grid


def features_in_area(f, grid, x, y, r, minl, maxl):  # src/Frame.cc:199-264
    iw, ih = inv_wh(f)
    x, y, r = F(x), F(y), F(r)
    a = max(0, math.floor(F(F(F(x - F(f.min_x)) - r) * iw)))
    if a >= 64:
        return []
    b = min(63, math.ceil(F(F(F(x - F(f.min_x)) + r) * iw)))
    if b < 0:
        return []
    c = max(0, math.floor(F(F(F(y - F(f.min_y)) - r) * ih)))
    if c >= 48:
        return []
    d = min(47, math.ceil(F(F(F(y - F(f.min_y)) + r) * ih)))
    if d < 0:
        return []
    check = not (minl == -1 and maxl == -1)
    same = check and minl == maxl
    out = []
    for ix in range(a, b + 1):
        for iy in range(c, d + 1):
            for k in grid[ix][iy]:
                o = int(f.octave[k])
                if check and not same:
                    if o < minl or o > maxl:
                        continue
                elif same and o != minl:
                    continue
                if abs(F(F(f.x[k]) - x)) > r or abs(F(F(f.y[k]) - y)) > r:
                    continue
                out.append(k)
    return out


def three_maxima(sizes):  # src/ORBmatcher.cc:1748-1789
    m1 = m2 = m3 = 0
    i1 = i2 = i3 = -1
    for i, s in enumerate(sizes):
        if s > m1:
            m3, m2, m1, i3, i2, i1 = m2, m1, s, i2, i1, i
        elif s > m2:
            m3, m2, i3, i2 = m2, s, i2, i
        elif s > m3:
            m3, i3 = s, i
    if F(m2) < F(0.1) * F(m1):
        i2 = i3 = -1
    elif F(m3) < F(0.1) * F(m1):
        i3 = -1
    return i1, i2, i3


def rot_bin(a1, a2):
    rot = F(F(a1) - F(a2))
    if rot < 0:
        rot = F(rot + F(360))
    b = c_round(F(rot * F(F(1) / F(HISTO))))
    return 0 if b == HISTO else b


def search_for_initialization(f1, f2, prev, window, ratio, ori):  # src/ORBmatcher.cc:598-713
    n1, n2 = len(f1.x), len(f2.x)
    m12 = [-1] * n1
    m21 = [-1] * n2
    md = [INT_MAX] * n2
    hist = [[] for _ in range(HISTO)]
    g2 = build_grid(f2)
    nm = 0
    for i1 in range(n1):
        if f1.octave[i1] > 0:
            continue
        cand = features_in_area(f2, g2, prev[i1, 0], prev[i1, 1], window, 0, 0)
        best = best2 = INT_MAX
        bi = -1
        for i2 in cand:
            d = dist(f1.desc[i1], f2.desc[i2])
            if md[i2] <= d:
                continue
            if d < best:
                best2, best, bi = best, d, i2
            elif d < best2:
                best2 = d
        if best <= TH_LOW and F(best) < F(F(best2) * F(ratio)):
            if m21[bi] >= 0:
                m12[m21[bi]] = -1
                nm -= 1
            m12[i1], m21[bi], md[bi] = bi, i1, best
            nm += 1
            if ori:
                hist[rot_bin(f1.angle[i1], f2.angle[bi])].append(i1)
    if ori:
        keep = three_maxima([len(h) for h in hist])
        for b in range(HISTO):
            if b in keep:
                continue
            for i1 in hist[b]:
                if m12[i1] >= 0:
                    m12[i1] = -1
                    nm -= 1
    for i1 in range(n1):
        if m12[i1] >= 0:
            prev[i1, 0], prev[i1, 1] = f2.x[m12[i1]], f2.y[m12[i1]]
    return nm, np.asarray(m12, np.int32)


def shared_nodes(fv1, fv2):
    ids2 = {int(v): k for k, v in enumerate(fv2[0])}
    return [(k1, ids2[int(v)]) for k1, v in enumerate(fv1[0]) if int(v) in ids2]


def search_by_bow_kf_f(kf, f, ratio, ori):  # src/ORBmatcher.cc:155-284
    mF = [-1] * len(f.x)
    hist = [[] for _ in range(HISTO)]
    nm = 0
    for k1, k2 in shared_nodes(kf.featvec, f.featvec):
        l1 = kf.featvec[2][kf.featvec[1][k1]:kf.featvec[1][k1 + 1]]
        l2 = f.featvec[2][f.featvec[1][k2]:f.featvec[1][k2 + 1]]
        for a in l1:
            if kf.valid is None or not kf.valid[a]:
                continue
            best = best2 = INT_MAX
            bi = -1
            for b in l2:
                if mF[b] >= 0:
                    continue
                d = dist(kf.desc[a], f.desc[b])
                if d < best:
                    best2, best, bi = best, d, int(b)
                elif d < best2:
                    best2 = d
            if best <= TH_LOW and F(best) < F(F(ratio) * F(best2)):
                mF[bi] = int(a)
                if ori:
                    hist[rot_bin(kf.angle[a], f.angle[bi])].append(bi)
                nm += 1
    if ori:
        keep = three_maxima([len(h) for h in hist])
        for b in range(HISTO):
            if b in keep:
                continue
            for i in hist[b]:
                mF[i] = -1
                nm -= 1
    return nm, np.asarray(mF, np.int32)


def knn2(q, db):
    out = np.empty((len(q), 2), np.uint32)
    for i in range(len(q)):
        d = np.unpackbits(np.bitwise_xor(db, q[i][None, :]), axis=1).sum(1).astype(np.int64)
        key = np.sort((d << 23) + np.arange(len(db)))
        out[i, 0] = key[0]
        out[i, 1] = key[1] if len(key) > 1 else 0xFFFFFFFF
    return out
