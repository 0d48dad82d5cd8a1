// Internal structures shared by the extractor's host planner (orbx_plan.cu), kernels (orbx_kernels.cu) and
// C-ABI layer (orbx_api.cu).  Not installed; the public surface is include/orbx.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/orbx.h"

namespace orbx {

constexpr int kMaxLevels = 16;
constexpr int kEdge = 16;        // EDGE_THRESHOLD, reference src/ORBextractor.cc:77
constexpr int kHalfPatch = 15;   // HALF_PATCH_SIZE :76
constexpr int kPatch = 31;       // PATCH_SIZE :75
constexpr int kRetryTh = 7;      // FAST retry threshold :613

// One FAST candidate / keypoint record while it moves through cell + level selection (8 bytes, swapped as a unit
// by the std::nth_element port).  xy = (y << 16) | x in level coordinates.
struct __align__(8) Cand {
  float resp;
  uint32_t xy;
};

struct LevelPlan {
  int w, h;              // level size: cvRound((float)cols * mvInvScaleFactor[l])            (:785-786)
  int pitch;             // bytes between rows of this level inside the pyramid slabs (multiple of 16)
  int64_t off;           // byte offset of the level inside one frame's pyramid slab (level 0 of the unblurred
                         // pyramid lives in the caller's frame buffer instead)
  int cols, rows;        // cell grid                                                          (:533-534)
  int cellW, cellH;      //                                                                    (:543-544)
  int nCells, quota;     // nCells = rows*cols, quota = nfeaturesCell                          (:546-547)
  int nDesired;          // mnFeaturesPerLevel[l]                                              (:476-487)
  int cell_base;         // index of the level's first cell in the per-frame cell arrays
  int kp_base;           // first keypoint slot of the level = sum of nDesired of the levels below
  int list_base, list_cap;  // per-frame level-list scratch (records) used between cell and level retention
  float scale;           // mvScaleFactor[l]                                                   (:462-465)
  float size;            // (float)(int)(PATCH_SIZE * mvScaleFactor[l])                        (:675)
  int tab_x, tab_y;      // offsets of this level's resize tables (xofs/alpha, yofs/beta)
};

struct alignas(16) CellPlan {   // 16-byte aligned: a kernel reads its record with five 128-bit loads
  int16_t level;
  int16_t x0, y0;        // first scored pixel of the cell (= iniX + 3, iniY + 3) in level coordinates
  int16_t rw, rh;        // scored region size; 0 when the reference skips the cell (hX <= 0 / hY <= 0 / < 7 px)
  int16_t skipped;       // bit 0: the reference `continue`s over this cell (hX <= 0 or hY <= 0); bit 1: "big" cell - its
                         // sub-image does not fit the shared-memory kernels, fast_cells_big_kernel serves it
  int32_t cand_off;      // offset of the cell's candidate list in the per-frame candidate array
  int32_t cand_cap;      // ceil(rw/2)*ceil(rh/2): NMS survivors are pairwise non-adjacent, so this cannot overflow
  // reciprocal multipliers (fast_magic) of the three row lengths fast_cells_kernel splits flat indices by
  uint32_t m_grp;        // staging groups (8 px) per sub-image row      = fast_ngrp(rw + 6)
  uint32_t m_g8;         // reject groups (8 px) per scored row          = (rw + 7) / 8
  uint32_t m_wpr;        // bitmap words (32 px) per scored row          = (rw + 31) / 32
  // fast_cells_tma_kernel: a tensor copy must start on a 16-byte boundary of the image row, so the box starts `lead`
  // columns left of x0 - 8 and the kernel works in shifted coordinates xs = x + lead over rows of lead + rw columns
  // Everything below is computed once on the host so that the kernel's prologue is one record load.
  int16_t lead;          // x0 & 15
  int16_t rwt;           // lead + rw: columns per shifted row
  int16_t g16;           // 16-pixel reject groups per shifted row = (rwt + 15) / 16
  int16_t scp;           // score-map pitch  = fast_scp(rwt)
  int16_t wpr;           // bitmap words per shifted row = (rwt + 31) / 32
  int16_t nbw;           // bitmap words = wpr * rh
  int32_t ngrp;          // reject groups = g16 * rh
  int32_t nz16;          // uint4 stores that clear the score map
  int32_t off_sc, off_bm, off_q;   // shared-memory byte offsets: score map (0: it overlays the raw image), bitmaps A + B, lists
  uint32_t m_g16;        // fast_magic(g16)
  uint32_t m_wprt;       // fast_magic(wpr)
  int32_t big_off;       // big cells: byte offset of score map + bitmaps in the per-frame global scratch
};

struct ResizeX {  // per destination column: source column and the two 11-bit weights (cv::resize INTER_LINEAR)
  int16_t sx, pad, a0, a1;  // (a0, a1) form one 32-bit word: the s16x2 operand of DP2A
};
struct ResizeY {
  int16_t sy0, sy1, b0, b1;
};

// Everything the kernels need to know about one (image size, knob set); lives in device global memory.
struct DevicePlan {
  int nlevels, nfeatures, fast_th, score_type, blur_mode, retain_mode;
  int width, height;
  int total_cells;           // cells per frame over all levels
  int64_t pyr_frame_bytes;   // bytes of one frame's pyramid slab (levels 1.. for unblurred, 0.. for blurred)
  int64_t blur_frame_bytes;
  int64_t cand_per_frame;    // candidate records per frame
  int list_per_frame;        // level-list scratch records per frame
  int umax[16];              // :495-510
  LevelPlan lv[kMaxLevels];
};

struct HostPlan {
  DevicePlan dp;
  std::vector<CellPlan> cells;
  std::vector<ResizeX> rx;
  std::vector<ResizeY> ry;
  int fast_smem_bytes;       // dynamic shared memory of the FAST kernel: max over cells of fast_cell_smem()
  // TMA-staged FAST kernel (fast_cells_tma_kernel): per level the tensor-copy box (the level's largest cell incl. the
  // 8-byte lead-in and the reject's over-read), its shared-memory footprint, and whether every level fits a box
  int fast_box_w[kMaxLevels], fast_box_h[kMaxLevels];
  int fast_tma_smem_bytes;
  bool fast_tma_ok;
  // TMA-staged resize (resize_tma_kernel): source box of a 128 x 64 destination tile of level l, 0 = not available
  int resize_box_w[kMaxLevels], resize_box_h[kMaxLevels];
  // cells beyond the shared-memory kernels (few features on a large image): indices and global scratch per frame
  std::vector<int32_t> big_cells;
  int64_t big_bytes_per_frame;
};

// Shared-memory geometry of fast_cells_kernel for a cell sub-image of cw x ch pixels (scored region (cw-6) x (ch-6)):
//   u16 image   zero-extended pixels, two per 32-bit word, for the packed VIMNMX3.U16x2 ladder; sub-image column c at
//               halfword c + 5 so that scored column 0 sits on a 16-byte boundary
//   u8 image    the same pixels >> 1 (7 bits) at byte c + 5, for the byte-SIMD (VABSDIFF4) reject
//   score map   u8, one zero ring
//   4 bitmaps   one bit per scored pixel, rows padded to 32-pixel words (row-major word order = row-major pixel order)
inline __host__ __device__ int fast_ngrp(int cw) { return (cw + 5 + 7) >> 3; }           // 8-pixel staging groups/row
inline __host__ __device__ int fast_sp16(int cw) { return (cw + 12 + 7) & ~7; }          // halfwords per row
inline __host__ __device__ int fast_bp(int rw) { return 8 * ((rw + 7) >> 3) + 16; }      // u8-image bytes per row
inline __host__ __device__ int fast_scp(int rw) { return ((rw + 3) & ~3) + 8; }          // score-map bytes per row
inline __host__ __device__ int fast_img_bytes(int cw, int ch) { return (ch * fast_sp16(cw) * 2 + 15) & ~15; }
inline __host__ __device__ int fast_h8_bytes(int cw, int ch) { return (ch * fast_bp(cw - 6) + 15) & ~15; }
inline __host__ __device__ int fast_sc_bytes(int rw, int rh) { return ((rh + 2) * fast_scp(rw) + 15) & ~15; }
inline __host__ __device__ int fast_bm_words(int rw, int rh) { return rh * ((rw + 31) >> 5); }
inline __host__ __device__ int fast_cell_smem(int cw, int ch) {
  return fast_img_bytes(cw, ch) + fast_h8_bytes(cw, ch) + fast_sc_bytes(cw - 6, ch - 6) +
         16 * fast_bm_words(cw - 6, ch - 6) + 16;
}
// fast_cells_tma_kernel: box width for a level whose widest scored region is rw (scored pixel x at byte 8 + x; the last
// 8-pixel reject group reads 4 bytes beyond its 8), multiple of 16 as the tensor map wants
inline __host__ __device__ int fast_tma_box_w(int rwt) { return 16 * ((rwt + 15) >> 4) + 32; }   // rwt = lead + rw; position xs at byte 16 + xs
constexpr int kFastTmaListBytes = (1024 + 512) * 4;   // kFastQ1 + kFastQ2 entries
// floor(i / d) == umulhi(i, fast_magic(d)) for i * d < 2^32; d == 1 gives 0, which magic_div() reads as "i itself"
inline uint32_t fast_magic(uint32_t d) { return 0xffffffffu / d + 1u; }

// Builds the plan exactly as the reference's constructor and ComputeKeyPoints do (float/double mix preserved).
// Returns ORBX_OK, ORBX_ERR_UNSUPPORTED (grid with 0 columns/rows, cell outside the level) or ORBX_ERR_ARG.
int build_plan(const orbx_params& p, int width, int height, HostPlan* out, std::string* err);

}  // namespace orbx
