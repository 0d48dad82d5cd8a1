// Kernel launch interface between orbx_api.cu and orbx_kernels.cu.
#pragma once
#include <cuda.h>   // CUtensorMap (type only; the encoder is resolved at run time, no -lcuda)

#include "orbx_internal.cuh"

namespace orbx {

// Device buffers of one chunk of frames (all sized at orbx_create for max_batch frames of max_w x max_h).
struct ChunkBuffers {
  const uint8_t* lv0;        // level 0 of the unblurred pyramid = the frames themselves
  int64_t lv0_frame_stride;  // bytes between frames
  int lv0_pitch;             // bytes between rows
  uint8_t* pyr;              // unblurred levels >= 1, [frame][pyr_frame_bytes]
  uint8_t* blur;             // blurred levels 0.., [frame][blur_frame_bytes]
  Cand* cand;                // [frame][cand_per_frame]   per-cell FAST candidates, row-major inside a cell
  int32_t* n_total;          // [frame][total_cells]      candidates per cell (after the th=7 retry rule)
  int32_t* n_retain;         // [frame][total_cells]      quota after redistribution
  Cand* list;                // [frame][list_per_frame]   level lists between cell and level retention
  Cand* kp;                  // [frame][nfeatures]        selected keypoints, slot = kp_base[level] + i
  int32_t* kp_count;         // [frame][kMaxLevels]
  int32_t* err_flag;         // [1] sticky device-side overflow flag
  const DevicePlan* plan;    // device copy
  const CellPlan* cells;     // device copy [total_cells]
  const ResizeX* rx;
  const ResizeY* ry;
};

// Tensor maps of fast_cells_tma_kernel, passed by value (__grid_constant__): one 3-D map (x, y, frame) of bytes per
// pyramid level, box = the level's largest cell sub-image.
struct FastTma {
  CUtensorMap map[kMaxLevels];
  int box_w[kMaxLevels], box_h[kMaxLevels];
  int smem_bytes;
};

// Tensor maps of resize_tma_kernel: map[l] describes the SOURCE of level l (= level l - 1; level 0 = the caller's frames)
// as a 3-D byte tensor (x, y, frame) with the box one CTA tile needs.
struct ResizeTma {
  CUtensorMap map[kMaxLevels];
  int box_w[kMaxLevels], box_h[kMaxLevels];
};
constexpr int kResizeTileW = 128, kResizeTileH = 64;   // destination tile of one resize_tma_kernel CTA (4 warps x 16 rows)

struct BlurTiles {            // everything blur7_kernel needs per level, passed by value (constant bank, no global loads)
  int nlevels, blur_mode;
  int base[kMaxLevels + 1];   // prefix of tiles per level
  int tiles_x[kMaxLevels];
  uint32_t magic_x[kMaxLevels];  // fast_magic(tiles_x): tile row = magic_div(tile, magic_x)
  int w[kMaxLevels], h[kMaxLevels], pitch[kMaxLevels];
  int64_t off[kMaxLevels];    // LevelPlan::off
  int64_t frame_bytes;        // DevicePlan::pyr_frame_bytes == blur_frame_bytes
};

// Programmatic dependent launch (griddepcontrol): when on, every extractor kernel is launched with
// cudaLaunchAttributeProgrammaticStreamSerialization, signals `launch_dependents` on entry and waits for its
// predecessor with `griddepcontrol.wait` before touching memory, so the NEXT kernel of the chain is already resident
// when this one drains.  Pays off when the chain is launch-latency bound (a chunk of a few frames); per host thread.
void set_launch_pdl(bool on);

void launch_resize_level(const ChunkBuffers& b, const DevicePlan& hp, int level, int nframes, cudaStream_t s);
void launch_resize_level_tma(const ChunkBuffers& b, const DevicePlan& hp, const ResizeTma& tm, int level, int nframes,
                             cudaStream_t s);
void launch_fast_cells(const ChunkBuffers& b, const DevicePlan& hp, int smem_bytes, int nframes, cudaStream_t s);
void launch_fast_cells_tma(const ChunkBuffers& b, const DevicePlan& hp, const FastTma& tm, int nframes, cudaStream_t s);
void launch_fast_cells_big(const ChunkBuffers& b, const DevicePlan& hp, const int32_t* d_big_cells, int nbig, uint8_t* d_scratch,
                           int64_t scratch_per_frame, int nframes, cudaStream_t s);
void launch_select(const ChunkBuffers& b, const DevicePlan& hp, int nframes, cudaStream_t s, int force);
void launch_retain_best_debug(Cand* v, const int32_t* start, const int32_t* keep, int nlists, int retain_mode, int batch,
                              cudaStream_t s);
void launch_blur(const ChunkBuffers& b, const DevicePlan& hp, const BlurTiles& t, int nframes, cudaStream_t s);
void launch_describe(const ChunkBuffers& b, const DevicePlan& hp, int nframes, orbx_keypoint* d_kps, uint8_t* d_desc,
                     int32_t* d_n_out, cudaStream_t s);
void launch_synth(uint8_t* d_out, uint8_t* d_tmp, int nframes, int w, int h, size_t stride, size_t frame_stride,
                  uint64_t seed0, cudaStream_t s);

}  // namespace orbx
