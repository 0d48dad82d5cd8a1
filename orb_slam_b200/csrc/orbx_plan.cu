// Host-side planning of the extractor: per-level scale tables, feature quotas, level sizes, resize weight tables
// and the per-cell detection grid, replayed with the reference's exact float/double arithmetic
// (reference src/ORBextractor.cc:457-511 constructor, :527-599 grid geometry, :781-786 level sizes).
#include <cfloat>
#include <cmath>
#include <cstring>

#include "orbx_internal.cuh"

namespace orbx {

static inline int cv_round_f(float v) { return (int)nearbyintf(v); }
static inline int cv_round_d(double v) { return (int)nearbyint(v); }
static inline int cv_floor_d(double v) { int i = (int)v; return i - (i > v); }
static inline int cv_ceil_d(double v) { int i = (int)v; return i + (i < v); }
static inline int align_up(int v, int a) { return (v + a - 1) / a * a; }

// cv::resize(INTER_LINEAR, 8U) coefficient tables: 11-bit fixed point, round-half-even.
static void resize_tables(int sw, int sh, int dw, int dh, std::vector<ResizeX>* rx, std::vector<ResizeY>* ry) {
  const double scale_x = (double)sw / dw, scale_y = (double)sh / dh;
  for (int dx = 0; dx < dw; dx++) {
    float fx = (float)((dx + 0.5) * scale_x - 0.5);
    int sx = cv_floor_d(fx);
    fx -= sx;
    if (sx < 0) { fx = 0; sx = 0; }
    if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
    ResizeX e;
    e.sx = (int16_t)sx;
    e.a0 = (int16_t)cv_round_f((1.f - fx) * 2048.f);
    e.a1 = (int16_t)cv_round_f(fx * 2048.f);
    e.pad = 0;
    rx->push_back(e);
  }
  for (int dy = 0; dy < dh; dy++) {
    float fy = (float)((dy + 0.5) * scale_y - 0.5);
    int sy = cv_floor_d(fy);
    fy -= sy;
    ResizeY e;
    e.sy0 = (int16_t)std::min(std::max(sy, 0), sh - 1);
    e.sy1 = (int16_t)std::min(std::max(sy + 1, 0), sh - 1);
    e.b0 = (int16_t)cv_round_f((1.f - fy) * 2048.f);
    e.b1 = (int16_t)cv_round_f(fy * 2048.f);
    ry->push_back(e);
  }
}

int build_plan(const orbx_params& p, int width, int height, HostPlan* out, std::string* err) {
  HostPlan& hp = *out;
  hp.cells.clear(); hp.rx.clear(); hp.ry.clear();
  DevicePlan& dp = hp.dp;
  memset(&dp, 0, sizeof(dp));
  if (p.nlevels < 1 || p.nlevels > kMaxLevels || p.nfeatures < 1 || width < 1 || height < 1 ||
      width > 32767 || height > 32767) {
    *err = "bad nlevels / nfeatures / image size";
    return ORBX_ERR_ARG;
  }
  const int nlevels = p.nlevels, nfeatures = p.nfeatures;
  const double scaleFactor = p.scale_factor;  // double member initialised from a float (ORBextractor.h:62)
  dp.nlevels = nlevels; dp.nfeatures = nfeatures; dp.fast_th = p.fast_th; dp.score_type = p.score_type;
  dp.blur_mode = p.blur_mode; dp.retain_mode = p.retain_mode; dp.width = width; dp.height = height;

  // :462-471
  std::vector<float> mvScaleFactor(nlevels), mvInvScaleFactor(nlevels);
  mvScaleFactor[0] = 1;
  for (int i = 1; i < nlevels; i++) mvScaleFactor[i] = mvScaleFactor[i - 1] * scaleFactor;
  float invScaleFactor = 1.0f / scaleFactor;
  mvInvScaleFactor[0] = 1;
  for (int i = 1; i < nlevels; i++) mvInvScaleFactor[i] = mvInvScaleFactor[i - 1] * invScaleFactor;
  // :476-487
  std::vector<int> mnFeaturesPerLevel(nlevels);
  float factor = (float)(1.0 / scaleFactor);
  float nDesiredFeaturesPerScale = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
  int sumFeatures = 0;
  for (int level = 0; level < nlevels - 1; level++) {
    mnFeaturesPerLevel[level] = cv_round_f(nDesiredFeaturesPerScale);
    sumFeatures += mnFeaturesPerLevel[level];
    nDesiredFeaturesPerScale *= factor;
  }
  mnFeaturesPerLevel[nlevels - 1] = std::max(nfeatures - sumFeatures, 0);
  // :495-510
  {
    int umax[kHalfPatch + 2] = {0};
    int v, v0, vmax = cv_floor_d(kHalfPatch * sqrtf(2.f) / 2 + 1);
    int vmin = cv_ceil_d(kHalfPatch * sqrtf(2.f) / 2);
    const double hp2 = kHalfPatch * kHalfPatch;
    for (v = 0; v <= vmax; ++v) umax[v] = cv_round_d(sqrt(hp2 - v * v));
    for (v = kHalfPatch, v0 = 0; v >= vmin; --v) {
      while (umax[v0] == umax[v0 + 1]) ++v0;
      umax[v] = v0;
      ++v0;
    }
    for (int i = 0; i < 16; i++) dp.umax[i] = umax[i];
    static const int kExpect[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};  // describe kernel's table
    for (int i = 0; i < 16; i++)
      if (umax[i] != kExpect[i]) { *err = "umax table differs from the kernel's constant"; return ORBX_ERR_UNSUPPORTED; }
  }

  const float imageRatio = (float)width / height;  // :527
  int64_t pyr_off = 0, blur_off = 0, cand_off = 0;
  int cell_base = 0, kp_base = 0, list_base = 0;
  hp.fast_smem_bytes = 0;
  hp.big_cells.clear();
  hp.big_bytes_per_frame = 0;
  hp.fast_tma_smem_bytes = 0;
  hp.fast_tma_ok = true;
  for (int l = 0; l < kMaxLevels; l++) hp.fast_box_w[l] = hp.fast_box_h[l] = 0;
  for (int level = 0; level < nlevels; level++) {
    LevelPlan& L = dp.lv[level];
    float scale = mvInvScaleFactor[level];
    L.w = cv_round_f((float)width * scale);   // :785-786
    L.h = cv_round_f((float)height * scale);
    if (L.w < 2 * kEdge + 1 || L.h < 2 * kEdge + 1) {
      *err = "pyramid level smaller than the 16-px border";
      return ORBX_ERR_UNSUPPORTED;
    }
    L.pitch = align_up(L.w, 16);
    L.off = blur_off;                       // blurred slab holds all levels ...
    blur_off += (int64_t)L.pitch * L.h;
    if (level > 0) pyr_off += (int64_t)L.pitch * L.h;  // ... the unblurred one levels >= 1 (same offsets minus lv0)
    L.scale = mvScaleFactor[level];
    L.size = (float)(int)(kPatch * mvScaleFactor[level]);  // :675
    L.nDesired = mnFeaturesPerLevel[level];
    L.kp_base = kp_base;
    kp_base += L.nDesired;
    L.tab_x = (int)hp.rx.size();
    L.tab_y = (int)hp.ry.size();
    hp.resize_box_w[level] = hp.resize_box_h[level] = 0;
    if (level > 0) {
      resize_tables(dp.lv[level - 1].w, dp.lv[level - 1].h, L.w, L.h, &hp.rx, &hp.ry);
      // source box of a 128 x 64 destination tile: columns from the 16-byte boundary at or below sx(first column) to
      // sx(last column) + 1, rows sy0(first row) .. sy1(last row); 4 consecutive columns must span <= 8 source bytes
      const ResizeX* rx = hp.rx.data() + L.tab_x;
      const ResizeY* ry = hp.ry.data() + L.tab_y;
      int bw = 0, bh = 0;
      bool ok = true;
      for (int x0 = 0; x0 < L.w; x0 += 128) {
        const int x1 = std::min(x0 + 127, L.w - 1);
        bw = std::max(bw, rx[x1].sx + 1 - (rx[x0].sx & ~15) + 1);
      }
      for (int x0 = 0; x0 < L.w; x0 += 4) ok = ok && rx[std::min(x0 + 3, L.w - 1)].sx + 1 - rx[x0].sx <= 7;
      for (int y0 = 0; y0 < L.h; y0 += 64) bh = std::max(bh, ry[std::min(y0 + 63, L.h - 1)].sy1 - ry[y0].sy0 + 1);
      bw = (bw + 12 + 15) & ~15;                    // + the aligned 12-byte read of the last column group
      if (ok && bw <= 256 && bh <= 256) { hp.resize_box_w[level] = bw; hp.resize_box_h[level] = bh; }
    }

    // grid geometry :531-547
    const int nDesiredFeatures = L.nDesired;
    const int levelCols = sqrtf((float)nDesiredFeatures / (5 * imageRatio));
    const int levelRows = imageRatio * levelCols;
    if (levelCols <= 0 || levelRows <= 0) {
      *err = "cell grid with zero rows/columns (the reference divides by zero here)";
      return ORBX_ERR_UNSUPPORTED;
    }
    const int minBorderX = kEdge, minBorderY = kEdge;
    const int maxBorderX = L.w - kEdge, maxBorderY = L.h - kEdge;
    const int W = maxBorderX - minBorderX, H = maxBorderY - minBorderY;
    const int cellW = ceilf((float)W / levelCols);
    const int cellH = ceilf((float)H / levelRows);
    const int nCells = levelRows * levelCols;
    L.cols = levelCols; L.rows = levelRows; L.cellW = cellW; L.cellH = cellH; L.nCells = nCells;
    L.quota = ceilf((float)nDesiredFeatures / nCells);
    L.cell_base = cell_base;
    cell_base += nCells;
    L.list_base = list_base;
    L.list_cap = 2 * nCells * L.quota + 2 * nCells + 8;
    list_base += L.list_cap;

    int max_rw = 0, max_rh = 0;   // this level's largest shifted scored region (TMA kernel)
    // cell loop :558-599
    std::vector<int> iniXCol(levelCols);
    float hY = cellH + 6;
    for (int i = 0; i < levelRows; i++) {
      const float iniY = minBorderY + i * cellH - 3;
      bool skip_row = false;
      if (i == levelRows - 1) {
        hY = maxBorderY + 3 - iniY;
        if (hY <= 0) skip_row = true;
      }
      float hX = cellW + 6;
      for (int j = 0; j < levelCols; j++) {
        float iniX;
        if (i == 0) { iniX = minBorderX + j * cellW - 3; iniXCol[j] = iniX; }
        else iniX = iniXCol[j];
        bool skip = skip_row;
        if (!skip_row && j == levelCols - 1) {
          hX = maxBorderX + 3 - iniX;
          if (hX <= 0) skip = true;
        }
        CellPlan c;
        memset(&c, 0, sizeof(c));
        c.level = (int16_t)level;
        c.cand_off = (int32_t)cand_off;
        c.skipped = skip ? 1 : 0;
        if (!skip) {
          const int cx0 = (int)iniX, cy0 = (int)iniY;
          const int cw = (int)(iniX + hX) - cx0, ch = (int)(iniY + hY) - cy0;
          if (cx0 < 0 || cy0 < 0 || cx0 + cw > L.w || cy0 + ch > L.h) {
            *err = "cell sub-image outside the level (the reference's cv::Mat range assertion fails here)";
            return ORBX_ERR_UNSUPPORTED;
          }
          if (fast_ngrp(cw) > 256) {
            *err = "cell wider than 2040 px (fast_cells_kernel stages one 8-pixel group per thread and row)";
            return ORBX_ERR_UNSUPPORTED;
          }
          if (cw >= 7 && ch >= 7) {  // FAST scores nothing on a sub-image narrower than 7 px
            c.x0 = (int16_t)(cx0 + 3); c.y0 = (int16_t)(cy0 + 3);
            c.rw = (int16_t)(cw - 6); c.rh = (int16_t)(ch - 6);
            c.cand_cap = ((c.rw + 1) / 2) * ((c.rh + 1) / 2);
            c.m_grp = fast_magic(fast_ngrp(cw));
            c.m_g8 = fast_magic((c.rw + 7) / 8);
            c.m_wpr = fast_magic((c.rw + 31) / 32);
            if (fast_cell_smem(cw, ch) > 200 * 1024) {
              // too large for one CTA's shared memory: fast_cells_big_kernel works on it in global memory
              c.skipped |= 2;
              c.big_off = (int32_t)hp.big_bytes_per_frame;
              hp.big_bytes_per_frame += (fast_sc_bytes(c.rw, c.rh) + 8 * fast_bm_words(c.rw, c.rh) + 255) & ~255;
              hp.big_cells.push_back((int32_t)hp.cells.size());
              hp.fast_tma_ok = false;
            } else {
              hp.fast_smem_bytes = std::max(hp.fast_smem_bytes, fast_cell_smem(cw, ch));
            }
            c.lead = (int16_t)(c.x0 & 15);
            c.rwt = (int16_t)(c.lead + c.rw);
            c.g16 = (int16_t)((c.rwt + 15) / 16);
            c.scp = (int16_t)fast_scp(c.rwt);
            c.wpr = (int16_t)((c.rwt + 31) / 32);
            c.nbw = (int16_t)(c.wpr * c.rh);
            c.ngrp = c.g16 * c.rh;
            c.m_g16 = fast_magic(c.g16);
            c.m_wprt = fast_magic(c.wpr);
            max_rw = std::max(max_rw, (int)c.rwt); max_rh = std::max(max_rh, (int)c.rh);
          }
        }
        cand_off += c.cand_cap;
        hp.cells.push_back(c);
      }
    }
    if (max_rw > 0) {
      const int bw = fast_tma_box_w(max_rw), bh = max_rh + 6;
      hp.fast_box_w[level] = bw; hp.fast_box_h[level] = bh;
      if (bw > 256 || bh > 256) hp.fast_tma_ok = false;          // the tensor-copy box is limited to 256 per dimension
      for (size_t k = hp.cells.size() - (size_t)nCells; k < hp.cells.size(); k++) {
        CellPlan& c = hp.cells[k];
        if (c.rw <= 0) continue;
        const int sc = fast_sc_bytes(c.rwt, c.rh);
        if (sc > bw * bh) hp.fast_tma_ok = false;                // the score map takes the image's place
        c.off_sc = 0; c.off_bm = bw * bh;
        c.off_q = (c.off_bm + 8 * c.nbw + 15) & ~15;
        c.nz16 = (sc + 15) / 16;
        if (c.ngrp + 4 > 4 * 512) hp.fast_tma_ok = false;         // the flag bytes share list 2's bytes (kFastQ2 entries)
        if ((1024 + 512) * 4 / c.scp < 3) hp.fast_tma_ok = false; // the dense walk needs a band of at least one row
        hp.fast_tma_smem_bytes = std::max(hp.fast_tma_smem_bytes, c.off_q + kFastTmaListBytes);
      }
    }
  }
  dp.total_cells = cell_base;
  dp.pyr_frame_bytes = blur_off;   // both slabs use the same per-frame layout (level 0 slot unused when unblurred)
  dp.blur_frame_bytes = blur_off;
  dp.cand_per_frame = cand_off;
  dp.list_per_frame = list_base;
  (void)pyr_off;
  return ORBX_OK;
}

}  // namespace orbx
