// Internal structures of the matcher (orbm_*.cu).  Public surface: include/orbm.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/orbm.h"

namespace orbm {

constexpr int kGridCols = ORBM_GRID_COLS, kGridRows = ORBM_GRID_ROWS, kGridCells = kGridCols * kGridRows;
constexpr uint32_t kNone = 0xFFFFFFFFu;
// candidate list entry: (distance << 22) | target index;  per-query list position < 2^22 as well
constexpr int kCandShift = 22;
constexpr uint32_t kCandMask = (1u << kCandShift) - 1;

// 256-bit Hamming distance (ORBmatcher::DescriptorDistance, reference src/ORBmatcher.cc:1794-1810: SWAR popcount of
// eight 32-bit XORs == sum of __popc).
__device__ __forceinline__ int hamming256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
  return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
         __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// A growable device buffer owned by the handle (grown outside the timed path; never shrunk).
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = bytes + bytes / 4 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

}  // namespace orbm

struct orbm_handle {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  std::string err;
  int64_t launches = 0;
  float last_ms = 0.f;
  int num_sms = 148;
  int commit_mode = 0;       // 0 by job size, 1 in-order replay, 2 fixed point (ORBM_COMMIT, read at orbm_create)
  int knn_variant = 2;       // distance arithmetic of knn2_kernel (ORBM_KNN_VARIANT, read at orbm_create)
  int knn_tc = -1;           // tensor-core 2-NN: -1 by job size, 0 never, 1 always (ORBM_KNN_TC, read at orbm_create)
  bool knn_tc_ready = false; // knn2_tc_kernel's dynamic shared-memory limit has been raised on this device
  // workspaces
  orbm::DevBuf knn_part;     // per-chunk partial top-2 keys
  orbm::DevBuf stage_a, stage_b, stage_c;  // staging for host-pointer entry points
  orbm::DevBuf ws_problem, ws_query, ws_target, ws_cand, ws_misc, ws_io;
  orbm::DevBuf ws_bow;       // per-keypoint descent results + uncompacted per-frame vectors (orbm_bow.cu)
  void* pinned = nullptr;    // pinned host staging (1 MB): small inputs go up / small results come back through it
  size_t pinned_cap = 0;
  void* pinned_big = nullptr;       // pinned gather block of large projected-point batches (grown on demand)
  size_t pinned_big_cap = 0;
  cudaMemPool_t pool = nullptr;     // private stream-ordered pool of the frame-set arrays
  void* last_comm = nullptr;        // communicator of the last orbm_knn2_sharded (asynchronous NCCL errors, orbm_synchronize)
  cudaEvent_t ev_order = nullptr;   // orders the handle's stream behind a caller's stream (created on first use)
};

// Device-resident frames: concatenated SoA + per-frame offsets + grids (+ optional feature vectors).
struct orbm_frameset {
  orbm_handle* h = nullptr;
  int nframes = 0, max_n = 0;
  int64_t total_kp = 0, total_nodes = 0, total_fv_items = 0;
  std::vector<int32_t> frame_off;   // host copy [nframes + 1]
  std::vector<int32_t> node_off;    // host copy [nframes + 1] (empty when no feature vectors)
  std::vector<int32_t> fvitem_off;  // host copy [nframes + 1]
  // device arrays
  int32_t* d_frame_off = nullptr;
  float *d_x = nullptr, *d_y = nullptr, *d_angle = nullptr;
  int32_t* d_octave = nullptr;
  uint8_t *d_desc = nullptr, *d_valid = nullptr;
  float4* d_gridp = nullptr;        // (min_x, min_y, inv_w, inv_h) per frame
  int32_t* d_cell_start = nullptr;  // [nframes][kGridCells + 1]
  int32_t* d_items = nullptr;       // [total_kp] grid CSR items (frame-local keypoint indices) at frame_off
  int32_t *d_node_off = nullptr, *d_fvitem_off = nullptr, *d_node_id = nullptr, *d_node_start = nullptr,
          *d_fv_items = nullptr;    // node_start has nnodes + 1 entries per frame, stored at node_off[f] + f
  bool has_fv = false;
  // BowVectors (orbm_frameset_compute_bow): words of frame f at frame_off[f] .. + bow_cnt[f], ascending word id
  int32_t* d_bow_word = nullptr;
  double* d_bow_val = nullptr;
  std::vector<int32_t> bow_cnt;     // host copy [nframes] (empty until computed)
  bool bow_sized = false;           // the feature / bow vector arrays have their upper-bound sizes (reusable)
};

// Frame-set arrays come from a stream-ordered pool PRIVATE to the handle (cudaMemPoolCreate at orbm_create, release
// threshold lifted so freed blocks are reused instead of returned to the driver, destroyed with the handle; the
// device's default pool, which other cudaMallocAsync users of the process share, is left alone): creating and
// destroying a frame set per tracked batch then costs microseconds, not a dozen cudaMalloc/cudaFree round trips.
inline cudaError_t orbm_alloc(orbm_handle* h, void** p, size_t bytes) {
  return cudaMallocFromPoolAsync(p, bytes, h->pool, h->stream);
}
template <class T> inline cudaError_t orbm_alloc(orbm_handle* h, T** p, size_t bytes) {
  return cudaMallocFromPoolAsync(reinterpret_cast<void**>(p), bytes, h->pool, h->stream);
}
inline void orbm_free(orbm_handle* h, void* p) { if (p) cudaFreeAsync(p, h->stream); }
constexpr size_t kOrbmPinnedBytes = 1u << 20;
inline cudaError_t orbm_ensure_pinned(orbm_handle* h) {
  if (h->pinned_cap >= kOrbmPinnedBytes) return cudaSuccess;
  if (h->pinned) cudaFreeHost(h->pinned);
  h->pinned = nullptr; h->pinned_cap = 0;
  const cudaError_t e = cudaHostAlloc(&h->pinned, kOrbmPinnedBytes, cudaHostAllocDefault);
  if (e == cudaSuccess) h->pinned_cap = kOrbmPinnedBytes;
  return e;
}

#define ORBM_CK(call)                                                          \
  do {                                                                         \
    cudaError_t e_ = (call);                                                   \
    if (e_ != cudaSuccess) {                                                   \
      h->err = std::string(#call) + ": " + cudaGetErrorString(e_);             \
      return ORBM_ERR_CUDA;                                                    \
    }                                                                          \
  } while (0)
