// C-ABI layer of the extractor (include/orbx.h): handle, device memory, chunked / pipelined execution.
// Host glue only — all arithmetic is in orbx_kernels.cu.  No CPU fallback: without a CUDA device orbx_create fails.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>

#include "orbx_kernels.cuh"

using namespace orbx;

// Independent chunk lanes (scratch set + streams) and host staging slots.  Consecutive chunks have no data dependence;
// more lanes let the issue-bound stages of one chunk fill the barrier / tail bubbles of another.
constexpr int kLanes = 4;          // upper bound; a handle uses h->nlanes of them (ORBX_LANES in the environment, default below)
constexpr int kDefaultLanes = 2;
constexpr int kPdlMaxFrames = 16;

struct orbx_handle {
  orbx_params p;
  HostPlan hp;
  int plan_w = 0, plan_h = 0;
  BlurTiles tiles;
  std::string err;
  // device memory (sized once for max_batch frames of max_width x max_height)
  // Two complete scratch sets + streams: consecutive chunks have no data dependence, so chunk k+1's kernels run beside
  // chunk k's and fill the SMs during its latency-bound stages (select) and kernel tails.
  struct Lane {
    uint8_t *d_pyr = nullptr, *d_blur = nullptr;
    Cand *d_cand = nullptr, *d_list = nullptr, *d_kp = nullptr;
    int32_t *d_n_total = nullptr, *d_n_retain = nullptr, *d_kp_count = nullptr;
    cudaStream_t s_main = nullptr, s_aux = nullptr;                       // blur runs beside select on s_aux
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_done = nullptr;
    cudaEvent_t ev_last = nullptr;      // end of the last chunk that used this scratch set, on whatever stream it ran
    cudaStream_t last_s = nullptr;
    bool used = false;
    // tensor maps of the TMA-staged FAST kernel: levels >= 1 point into d_pyr (valid for one plan), level 0 into the
    // caller's frames (re-encoded when pointer / strides change)
    uint8_t* d_big = nullptr;           // scratch of fast_cells_big_kernel (plans with oversized cells only)
    size_t big_cap = 0;
    FastTma tma;
    ResizeTma rtma;                     // source boxes of resize_tma_kernel; rtma_ok[l]: level l has a valid map
    bool rtma_ok[kMaxLevels] = {};
    int tma_plan = -1;
    const uint8_t* tma_lv0 = nullptr;
    size_t tma_stride = 0, tma_frame_stride = 0;
    int rtma_plan = -1;
    const uint8_t* rtma_lv0 = nullptr;
    size_t rtma_stride = 0, rtma_frame_stride = 0;
    bool rtma_lv0_valid = false;
  } lane[kLanes];
  int32_t* d_big_cells = nullptr;   // indices of the oversized cells of the current plan
  size_t big_cells_cap = 0;
  int plan_version = 0;
  int fast_mode = 0;                // 0 TMA kernel when the buffers allow it, 1 legacy kernel only (ORBX_FAST=legacy)
  int resize_mode = 0;              // pyramid kernels: 0 TMA for chunks of > 16 frames, 1 legacy only, 2 TMA always (ORBX_RESIZE)
  int nlanes = kDefaultLanes;
  int select_mode = 0;              // 0 by chunk size, 1 global, 2 shared memory (ORBX_SELECT)
  int pdl_mode = 1;                 // 0 never, 1 chunks of <= kPdlMaxFrames frames, 2 always (ORBX_PDL)
  uint8_t* d_in[kLanes] = {};
  uint8_t* d_raw[kLanes] = {};   // tightly packed host rows land here first (one contiguous H2D), then D2D repitch
  int32_t* d_err = nullptr;
  int32_t* h_err = nullptr;                 // pinned mirror of d_err, filled by an async copy behind the outputs
  DevicePlan* d_plan = nullptr;
  CellPlan* d_cells = nullptr;
  ResizeX* d_rx = nullptr;
  ResizeY* d_ry = nullptr;
  orbx_keypoint* d_out_kps[kLanes] = {};
  uint8_t* d_out_desc[kLanes] = {};
  int32_t* d_out_n[kLanes] = {};
  size_t cap_cells = 0, cap_rx = 0, cap_ry = 0;
  int64_t cap_pyr = 0, cap_cand = 0, cap_list = 0, cap_cells_frame = 0;
  int in_pitch = 0;
  int64_t in_frame_bytes = 0;
  cudaStream_t s_in = nullptr, s_out = nullptr;
  cudaEvent_t ev_user = nullptr;                      // fork point on a caller's stream (device-resident API)
  cudaEvent_t ev_in[kLanes], ev_compute[kLanes], ev_out[kLanes], ev_stage[ORBX_STAGE_COUNT + 1];
  bool profiling = false;
  float prof_ms[ORBX_STAGE_COUNT] = {0};
  int64_t prof_launch[ORBX_STAGE_COUNT] = {0};
  int64_t launches = 0;
  // last chunk (for the debug getters)
  ChunkBuffers last{};
  int last_frames = 0;
};

static thread_local std::string g_create_err;

#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      h->err = std::string(#call) + ": " + cudaGetErrorString(e_);                                 \
      return ORBX_ERR_CUDA;                                                                        \
    }                                                                                              \
  } while (0)

extern "C" int orbx_default_params(orbx_params* p) {
  if (!p) return ORBX_ERR_ARG;
  p->nfeatures = 1000; p->scale_factor = 1.2f; p->nlevels = 8; p->score_type = ORBX_FAST_SCORE; p->fast_th = 20;
  p->max_width = 640; p->max_height = 480; p->max_batch = 256; p->device = 0;
  p->blur_mode = ORBX_BLUR_CV24; p->retain_mode = ORBX_RETAIN_CV24;
  return ORBX_OK;
}

// Minimal `%YAML:1.0` "key: value" reader for the five ORBextractor.* keys (the reference uses cv::FileStorage,
// src/Tracking.cc:52,105-109).  Note the shipped file writes "Camera.k1:-0.4157" without a space; accepted.
extern "C" int orbx_params_from_yaml(const char* path, orbx_params* p) {
  if (!path || !p) return ORBX_ERR_ARG;
  std::ifstream f(path);
  if (!f) { g_create_err = std::string("cannot open ") + path; return ORBX_ERR_IO; }
  std::string line;
  int found = 0;
  while (std::getline(f, line)) {
    size_t hash = line.find('#');
    if (hash != std::string::npos) line.resize(hash);
    size_t colon = line.find(':');
    if (colon == std::string::npos) continue;
    std::string key = line.substr(0, colon), val = line.substr(colon + 1);
    while (!key.empty() && isspace((unsigned char)key.back())) key.pop_back();
    size_t k0 = 0;
    while (k0 < key.size() && isspace((unsigned char)key[k0])) k0++;
    key = key.substr(k0);
    double v = atof(val.c_str());
    if (key == "ORBextractor.nFeatures") { p->nfeatures = (int)v; found |= 1; }
    else if (key == "ORBextractor.scaleFactor") { p->scale_factor = (float)v; found |= 2; }
    else if (key == "ORBextractor.nLevels") { p->nlevels = (int)v; found |= 4; }
    else if (key == "ORBextractor.fastTh") { p->fast_th = (int)v; found |= 8; }
    else if (key == "ORBextractor.nScoreType") { p->score_type = (int)v; found |= 16; }
  }
  if (found != 31) { g_create_err = "Settings file lacks one of the ORBextractor.* keys"; return ORBX_ERR_IO; }
  if (p->score_type != 0 && p->score_type != 1) {  // assert(Score==1 || Score==0), src/Tracking.cc:111
    g_create_err = "ORBextractor.nScoreType must be 0 or 1";
    return ORBX_ERR_ARG;
  }
  return ORBX_OK;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time libcuda dependency)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}

// 3-D byte tensor (x, y, frame) with rows `pitch` and frames `frame_bytes` apart; box = bw x bh x 1, zero fill outside
static bool encode_level_map(CUtensorMap* m, const void* base, int w, int hgt, int nframes, size_t pitch, size_t frame_bytes,
                             int bw, int bh) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn || (reinterpret_cast<uintptr_t>(base) & 15) || (pitch & 15) || (frame_bytes & 15) || bw > 256 || bh > 256) return false;
  const cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)hgt, (cuuint64_t)nframes};
  const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)frame_bytes};
  const cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, 1};
  const cuuint32_t es[3] = {1, 1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// Tensor maps of one chunk; false = this chunk takes the legacy kernel (misaligned caller buffer, oversized cells, ...)
static bool prepare_fast_tma(orbx_handle* h, orbx_handle::Lane& ln, const uint8_t* lv0, size_t stride, size_t frame_stride) {
  const HostPlan& hp = h->hp;
  if (h->fast_mode == 1 || !hp.fast_tma_ok || hp.fast_tma_smem_bytes > 200 * 1024) return false;
  if (h->p.fast_th > 120) return false;          // its byte-SIMD reject assumes threshold + 1 < 128
  const DevicePlan& dp = hp.dp;
  const int B = h->p.max_batch;
  if (ln.tma_plan != h->plan_version) {
    for (int l = 1; l < dp.nlevels; l++) {
      if (!hp.fast_box_w[l]) continue;                               // level without a scored cell: never addressed
      if (!encode_level_map(&ln.tma.map[l], ln.d_pyr + dp.lv[l].off, dp.lv[l].w, dp.lv[l].h, B, dp.lv[l].pitch,
                            (size_t)dp.pyr_frame_bytes, hp.fast_box_w[l], hp.fast_box_h[l]))
        return false;
    }
    for (int l = 0; l < kMaxLevels; l++) { ln.tma.box_w[l] = hp.fast_box_w[l]; ln.tma.box_h[l] = hp.fast_box_h[l]; }
    ln.tma.smem_bytes = hp.fast_tma_smem_bytes;
    ln.tma_plan = h->plan_version;
    ln.tma_lv0 = nullptr;
  }
  if (ln.tma_lv0 != lv0 || ln.tma_stride != stride || ln.tma_frame_stride != frame_stride) {
    ln.tma_lv0 = nullptr;
    if (hp.fast_box_w[0] &&
        !encode_level_map(&ln.tma.map[0], lv0, dp.lv[0].w, dp.lv[0].h, B, stride, frame_stride, hp.fast_box_w[0], hp.fast_box_h[0]))
      return false;
    ln.tma_lv0 = lv0; ln.tma_stride = stride; ln.tma_frame_stride = frame_stride;
  }
  return true;
}

// Source maps of resize_tma_kernel for this chunk (ln.rtma_ok[l] says which levels have one; the others take
// resize_level_kernel).  Levels >= 2 read the lane's own pyramid slab, level 1 the caller's frames.
static void prepare_resize_tma(orbx_handle* h, orbx_handle::Lane& ln, const uint8_t* lv0, size_t stride, size_t frame_stride) {
  const HostPlan& hp = h->hp;
  const DevicePlan& dp = hp.dp;
  const int B = h->p.max_batch;
  if (ln.rtma_plan != h->plan_version) {
    for (int l = 0; l < kMaxLevels; l++) { ln.rtma_ok[l] = false; ln.rtma.box_w[l] = hp.resize_box_w[l]; ln.rtma.box_h[l] = hp.resize_box_h[l]; }
    if (h->resize_mode != 1)
      for (int l = 2; l < dp.nlevels; l++)
        ln.rtma_ok[l] = hp.resize_box_w[l] > 0 &&
                        encode_level_map(&ln.rtma.map[l], ln.d_pyr + dp.lv[l - 1].off, dp.lv[l - 1].w, dp.lv[l - 1].h, B,
                                         dp.lv[l - 1].pitch, (size_t)dp.pyr_frame_bytes, hp.resize_box_w[l], hp.resize_box_h[l]);
    ln.rtma_plan = h->plan_version;
    ln.rtma_lv0 = nullptr;
    ln.rtma_lv0_valid = false;
  }
  if (dp.nlevels > 1 && (!ln.rtma_lv0_valid || ln.rtma_lv0 != lv0 || ln.rtma_stride != stride || ln.rtma_frame_stride != frame_stride)) {
    ln.rtma_ok[1] = h->resize_mode != 1 && hp.resize_box_w[1] > 0 &&
                    encode_level_map(&ln.rtma.map[1], lv0, dp.lv[0].w, dp.lv[0].h, B, stride, frame_stride, hp.resize_box_w[1],
                                     hp.resize_box_h[1]);
    ln.rtma_lv0 = lv0; ln.rtma_stride = stride; ln.rtma_frame_stride = frame_stride; ln.rtma_lv0_valid = true;
  }
}

static int upload_plan(orbx_handle* h, int w, int hgt) {
  if (h->plan_w == w && h->plan_h == hgt) return ORBX_OK;
  int rc = build_plan(h->p, w, hgt, &h->hp, &h->err);
  if (rc != ORBX_OK) return rc;
  const DevicePlan& dp = h->hp.dp;
  if (h->hp.cells.size() > h->cap_cells || h->hp.rx.size() > h->cap_rx || h->hp.ry.size() > h->cap_ry ||
      dp.pyr_frame_bytes > h->cap_pyr || dp.cand_per_frame > h->cap_cand || dp.list_per_frame > h->cap_list ||
      dp.total_cells > h->cap_cells_frame) {
    h->err = "image larger than the max_width x max_height the handle was created for";
    return ORBX_ERR_ARG;
  }
  // (cells too large for the shared-memory FAST kernels are served by fast_cells_big_kernel: scratch below)
  // the tables below are read by kernels of earlier calls, which may still run on the lanes' streams or - single-chunk
  // device-API calls - on a caller's stream (ev_last): wait for all of them before overwriting
  for (int li = 0; li < h->nlanes; li++) {
    CK(cudaStreamSynchronize(h->lane[li].s_main));
    if (h->lane[li].used) CK(cudaEventSynchronize(h->lane[li].ev_last));
  }
  h->plan_version++;
  if (!h->hp.big_cells.empty()) {
    // global scratch of the big-cell kernel (score map + bitmaps per big cell and frame), grown on demand: such plans
    // are rare (few features on a large image), so the memory is not reserved at orbx_create
    const size_t need = (size_t)h->hp.big_bytes_per_frame * h->p.max_batch;
    for (int li = 0; li < h->nlanes; li++) {
      orbx_handle::Lane& ln = h->lane[li];
      if (ln.big_cap < need) {
        if (ln.d_big) cudaFree(ln.d_big);
        ln.d_big = nullptr; ln.big_cap = 0;
        CK(cudaMalloc(&ln.d_big, need));
        ln.big_cap = need;
      }
    }
    if (h->big_cells_cap < h->hp.big_cells.size()) {
      if (h->d_big_cells) cudaFree(h->d_big_cells);
      h->d_big_cells = nullptr; h->big_cells_cap = 0;
      CK(cudaMalloc(&h->d_big_cells, h->hp.big_cells.size() * sizeof(int32_t)));
      h->big_cells_cap = h->hp.big_cells.size();
    }
    CK(cudaMemcpy(h->d_big_cells, h->hp.big_cells.data(), h->hp.big_cells.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
  }
  CK(cudaMemcpy(h->d_plan, &dp, sizeof(dp), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(h->d_cells, h->hp.cells.data(), h->hp.cells.size() * sizeof(CellPlan), cudaMemcpyHostToDevice));
  if (!h->hp.rx.empty()) {
    CK(cudaMemcpy(h->d_rx, h->hp.rx.data(), h->hp.rx.size() * sizeof(ResizeX), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(h->d_ry, h->hp.ry.data(), h->hp.ry.size() * sizeof(ResizeY), cudaMemcpyHostToDevice));
  }
  int base = 0;
  memset(&h->tiles, 0, sizeof(h->tiles));
  h->tiles.nlevels = dp.nlevels; h->tiles.blur_mode = dp.blur_mode; h->tiles.frame_bytes = dp.pyr_frame_bytes;
  for (int l = 0; l < dp.nlevels; l++) {
    h->tiles.w[l] = dp.lv[l].w; h->tiles.h[l] = dp.lv[l].h; h->tiles.pitch[l] = dp.lv[l].pitch;
    h->tiles.off[l] = dp.lv[l].off;
    h->tiles.base[l] = base;
    h->tiles.tiles_x[l] = (dp.lv[l].w + 63) / 64;          // blur7_kernel tiles: 64 columns x 58 rows
    h->tiles.magic_x[l] = fast_magic(h->tiles.tiles_x[l]);
    base += h->tiles.tiles_x[l] * ((dp.lv[l].h + 57) / 58);
  }
  h->tiles.base[dp.nlevels] = base;
  h->plan_w = w; h->plan_h = hgt;
  return ORBX_OK;
}

extern "C" int orbx_create(const orbx_params* pp, orbx_handle** out) {
  if (!pp || !out) return ORBX_ERR_ARG;
  *out = nullptr;
  orbx_handle* h = new orbx_handle();
  h->p = *pp;
  auto fail = [&](int rc) { g_create_err = h->err; delete h; return rc; };
  if (pp->max_batch < 1 || pp->max_width < 1 || pp->max_height < 1 || pp->fast_th < 1 || pp->fast_th > 254 ||
      (pp->score_type != 0 && pp->score_type != 1) || pp->blur_mode < 0 || pp->blur_mode > 1 ||
      pp->retain_mode < 0 || pp->retain_mode > 1 || !(pp->scale_factor > 1.0f)) {
    h->err = "bad parameters";
    return fail(ORBX_ERR_ARG);
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || pp->device < 0 || pp->device >= ndev) {
    h->err = "no usable CUDA device (this library has no CPU fallback)";
    return fail(ORBX_ERR_CUDA);
  }
  auto body = [&]() -> int {
    CK(cudaSetDevice(pp->device));
    HostPlan mp;  // plan of the largest frame sizes every buffer
    int rc = build_plan(h->p, pp->max_width, pp->max_height, &mp, &h->err);
    if (rc != ORBX_OK) return rc;
    const DevicePlan& dp = mp.dp;
    const int64_t B = pp->max_batch;
    // smaller frames have fewer pixels but may have (slightly) different grids: keep generous slack
    h->cap_cells = mp.cells.size() * 2 + 64; h->cap_rx = mp.rx.size() + 64; h->cap_ry = mp.ry.size() + 64;
    h->cap_pyr = dp.pyr_frame_bytes; h->cap_cand = dp.cand_per_frame + 1024; h->cap_list = dp.list_per_frame * 2;
    h->cap_cells_frame = dp.total_cells * 2 + 64;
    h->in_pitch = (pp->max_width + 15) / 16 * 16;
    h->in_frame_bytes = (int64_t)h->in_pitch * pp->max_height;
    if (const char* e = getenv("ORBX_PDL")) h->pdl_mode = std::min(2, std::max(0, atoi(e)));
    if (const char* e = getenv("ORBX_SELECT")) h->select_mode = e[0] == 'g' ? 1 : (e[0] == 's' ? 2 : 0);
    if (const char* e = getenv("ORBX_FAST")) h->fast_mode = e[0] == 'l' ? 1 : 0;
    if (const char* e = getenv("ORBX_RESIZE")) h->resize_mode = e[0] == 'l' ? 1 : (e[0] == 't' ? 2 : 0);
    if (const char* e = getenv("ORBX_LANES")) h->nlanes = std::min(kLanes, std::max(1, atoi(e)));
    for (int li = 0; li < h->nlanes; li++) {
      orbx_handle::Lane& ln = h->lane[li];
      CK(cudaMalloc(&ln.d_pyr, B * h->cap_pyr));
      CK(cudaMalloc(&ln.d_blur, B * h->cap_pyr));
      // 4 pad records in front (cap_cand already leaves slack behind): the batched partition scans of select_kernel may
      // read up to 3 records beyond either end of a cell's list
      CK(cudaMalloc(&ln.d_cand, (B * h->cap_cand + 8) * sizeof(Cand)));
      CK(cudaMemset(ln.d_cand, 0, (B * h->cap_cand + 8) * sizeof(Cand)));
      CK(cudaMalloc(&ln.d_list, B * h->cap_list * sizeof(Cand)));
      CK(cudaMalloc(&ln.d_kp, B * pp->nfeatures * sizeof(Cand)));
      CK(cudaMalloc(&ln.d_n_total, B * h->cap_cells_frame * sizeof(int32_t)));
      CK(cudaMalloc(&ln.d_n_retain, B * h->cap_cells_frame * sizeof(int32_t)));
      CK(cudaMalloc(&ln.d_kp_count, B * kMaxLevels * sizeof(int32_t)));
      CK(cudaStreamCreateWithFlags(&ln.s_main, cudaStreamNonBlocking));
      CK(cudaStreamCreateWithFlags(&ln.s_aux, cudaStreamNonBlocking));
      CK(cudaEventCreateWithFlags(&ln.ev_fork, cudaEventDisableTiming));
      CK(cudaEventCreateWithFlags(&ln.ev_join, cudaEventDisableTiming));
      CK(cudaEventCreateWithFlags(&ln.ev_done, cudaEventDisableTiming));
      CK(cudaEventCreateWithFlags(&ln.ev_last, cudaEventDisableTiming));
    }
    CK(cudaMalloc(&h->d_err, sizeof(int32_t)));
    CK(cudaMemset(h->d_err, 0, sizeof(int32_t)));
    CK(cudaHostAlloc((void**)&h->h_err, sizeof(int32_t), cudaHostAllocDefault));
    *h->h_err = 0;
    CK(cudaMalloc(&h->d_plan, sizeof(DevicePlan)));
    CK(cudaMalloc(&h->d_cells, h->cap_cells * sizeof(CellPlan)));
    CK(cudaMalloc(&h->d_rx, h->cap_rx * sizeof(ResizeX)));
    CK(cudaMalloc(&h->d_ry, h->cap_ry * sizeof(ResizeY)));
    for (int i = 0; i < h->nlanes; i++) {
      CK(cudaMalloc(&h->d_in[i], B * h->in_frame_bytes));
      CK(cudaMalloc(&h->d_raw[i], B * h->in_frame_bytes));
      CK(cudaMalloc(&h->d_out_kps[i], B * pp->nfeatures * sizeof(orbx_keypoint)));
      CK(cudaMalloc(&h->d_out_desc[i], B * pp->nfeatures * 32));
      CK(cudaMalloc(&h->d_out_n[i], B * sizeof(int32_t)));
      CK(cudaEventCreateWithFlags(&h->ev_in[i], cudaEventDisableTiming));
      CK(cudaEventCreateWithFlags(&h->ev_compute[i], cudaEventDisableTiming));
      CK(cudaEventCreateWithFlags(&h->ev_out[i], cudaEventDisableTiming));
    }
    for (int i = 0; i <= ORBX_STAGE_COUNT; i++) CK(cudaEventCreate(&h->ev_stage[i]));
    CK(cudaStreamCreateWithFlags(&h->s_in, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&h->s_out, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&h->ev_user, cudaEventDisableTiming));
    return upload_plan(h, pp->max_width, pp->max_height);
  };
  int rc = body();
  if (rc != ORBX_OK) { std::string m = h->err; orbx_destroy(h); g_create_err = m; return rc; }
  *out = h;
  return ORBX_OK;
}

extern "C" void orbx_destroy(orbx_handle* h) {
  if (!h) return;
  cudaSetDevice(h->p.device);
  cudaDeviceSynchronize();
  void* ptrs[] = {h->d_err, h->d_plan, h->d_cells, h->d_rx, h->d_ry, h->d_big_cells};
  for (void* p : ptrs) if (p) cudaFree(p);
  if (h->h_err) cudaFreeHost(h->h_err);
  for (int i = 0; i < kLanes; i++) {
    void* sp[] = {h->d_in[i], h->d_raw[i], h->d_out_kps[i], h->d_out_desc[i], h->d_out_n[i]};
    for (void* p : sp) if (p) cudaFree(p);
  }
  for (auto& ln : h->lane) {
    void* lp[] = {ln.d_pyr, ln.d_blur, ln.d_cand, ln.d_list, ln.d_kp, ln.d_n_total, ln.d_n_retain, ln.d_kp_count, ln.d_big};
    for (void* p : lp) if (p) cudaFree(p);
    if (ln.s_main) cudaStreamDestroy(ln.s_main);
    if (ln.s_aux) cudaStreamDestroy(ln.s_aux);
    if (ln.ev_fork) cudaEventDestroy(ln.ev_fork);
    if (ln.ev_join) cudaEventDestroy(ln.ev_join);
    if (ln.ev_done) cudaEventDestroy(ln.ev_done);
    if (ln.ev_last) cudaEventDestroy(ln.ev_last);
  }
  if (h->s_in) {
    cudaStreamDestroy(h->s_in); cudaStreamDestroy(h->s_out);
    if (h->ev_user) cudaEventDestroy(h->ev_user);
    for (int i = 0; i < h->nlanes; i++) { cudaEventDestroy(h->ev_in[i]); cudaEventDestroy(h->ev_compute[i]); cudaEventDestroy(h->ev_out[i]); }
    for (int i = 0; i <= ORBX_STAGE_COUNT; i++) cudaEventDestroy(h->ev_stage[i]);
  }
  delete h;
}

extern "C" int orbx_get_levels(const orbx_handle* h) { return h ? h->p.nlevels : 0; }
extern "C" float orbx_get_scale_factor(const orbx_handle* h) { return h ? (float)(double)h->p.scale_factor : 0.f; }
extern "C" const char* orbx_last_error(const orbx_handle* h) { return h ? h->err.c_str() : g_create_err.c_str(); }
extern "C" int64_t orbx_kernel_launches(const orbx_handle* h) { return h ? h->launches : 0; }

// One chunk (<= max_batch frames) on stream s: 7 resizes, FAST cells, select, blur, describe = nlevels + 3 launches.
// `li` selects the scratch set; s is the stream the chunk is ordered on (the lane's own s_main or a caller's stream).
static int run_chunk(orbx_handle* h, int li, const uint8_t* d_imgs, int n, size_t stride, size_t frame_stride,
                     orbx_keypoint* d_kps, uint8_t* d_desc, int32_t* d_n, cudaStream_t s) {
  const DevicePlan& dp = h->hp.dp;
  orbx_handle::Lane& ln = h->lane[li];
  ChunkBuffers b{};
  b.lv0 = d_imgs; b.lv0_frame_stride = (int64_t)frame_stride; b.lv0_pitch = (int)stride;
  b.pyr = ln.d_pyr; b.blur = ln.d_blur; b.cand = ln.d_cand + 4; b.n_total = ln.d_n_total; b.n_retain = ln.d_n_retain;
  b.list = ln.d_list; b.kp = ln.d_kp; b.kp_count = ln.d_kp_count; b.err_flag = h->d_err;
  b.plan = h->d_plan; b.cells = h->d_cells; b.rx = h->d_rx; b.ry = h->d_ry;
  // this scratch set was last used on another stream (a caller's, or the lane's own): order behind that chunk
  if (ln.used && ln.last_s != s) CK(cudaStreamWaitEvent(s, ln.ev_last, 0));
  const bool fast_tma = prepare_fast_tma(h, ln, d_imgs, stride, frame_stride);
  prepare_resize_tma(h, ln, d_imgs, stride, frame_stride);
  const bool prof = h->profiling;
  // launch-latency-bound chains (a few frames per chunk) use programmatic dependent launch; full chunks keep the
  // plain stream order, where the other lane's kernels already fill the tails (ORBX_PDL=0 never / 2 always)
  set_launch_pdl(h->pdl_mode == 2 || (h->pdl_mode == 1 && n <= kPdlMaxFrames));
  int st = 0;
  if (prof) cudaEventRecord(h->ev_stage[st++], s);
  // a chunk of a few frames is a chain of 7 latency-bound launches: there the register kernel, which starts computing
  // with its first loads, beats waiting for a tensor copy per CTA (28 vs 42 us for one frame); ORBX_RESIZE=tma forces it
  const bool resize_tma = h->resize_mode == 2 || n > kPdlMaxFrames;
  for (int l = 1; l < dp.nlevels; l++) {
    if (resize_tma && ln.rtma_ok[l]) launch_resize_level_tma(b, dp, ln.rtma, l, n, s);
    else launch_resize_level(b, dp, l, n, s);
  }
  if (prof) cudaEventRecord(h->ev_stage[st++], s);
  if (fast_tma) launch_fast_cells_tma(b, dp, ln.tma, n, s);
  else launch_fast_cells(b, dp, h->hp.fast_smem_bytes, n, s);
  if (!h->hp.big_cells.empty()) {
    launch_fast_cells_big(b, dp, h->d_big_cells, (int)h->hp.big_cells.size(), ln.d_big, h->hp.big_bytes_per_frame, n, s);
    h->launches++;
  }
  if (prof) cudaEventRecord(h->ev_stage[st++], s);
  if (prof) {                     // serial, so that the per-stage events mean something
    launch_select(b, dp, n, s, h->select_mode);
    cudaEventRecord(h->ev_stage[st++], s);
    launch_blur(b, dp, h->tiles, n, s);
    cudaEventRecord(h->ev_stage[st++], s);
  } else {
    // select is a latency-bound kernel of a few thousand warps and blur does not depend on it: blur runs beside it
    // on s_aux (fork after FAST, whose CTAs fill the SMs' shared memory; join before describe)
    CK(cudaEventRecord(ln.ev_fork, s));
    launch_select(b, dp, n, s, h->select_mode);
    CK(cudaStreamWaitEvent(ln.s_aux, ln.ev_fork, 0));
    launch_blur(b, dp, h->tiles, n, ln.s_aux);
    CK(cudaEventRecord(ln.ev_join, ln.s_aux));
    CK(cudaStreamWaitEvent(s, ln.ev_join, 0));
  }
  launch_describe(b, dp, n, d_kps, d_desc, d_n, s);
  if (prof) cudaEventRecord(h->ev_stage[st++], s);
  h->launches += dp.nlevels + 3;
  CK(cudaGetLastError());
  if (prof) {
    CK(cudaEventSynchronize(h->ev_stage[ORBX_STAGE_COUNT]));
    const int nl[ORBX_STAGE_COUNT] = {dp.nlevels - 1, 1, 1, 1, 1};
    for (int i = 0; i < ORBX_STAGE_COUNT; i++) {
      float ms = 0;
      cudaEventElapsedTime(&ms, h->ev_stage[i], h->ev_stage[i + 1]);
      h->prof_ms[i] += ms;
      h->prof_launch[i] += nl[i];
    }
  }
  CK(cudaEventRecord(ln.ev_last, s));
  ln.last_s = s; ln.used = true;
  h->last = b; h->last_frames = n;
  return ORBX_OK;
}

static int check_args(orbx_handle* h, const void* img, int n, int w, int hgt, size_t stride, const void* kps,
                      const void* desc, const void* n_out) {
  if (!h) return ORBX_ERR_ARG;
  if (!kps || !desc || !n_out || n < 0 || (img && (w < 0 || hgt < 0))) { h->err = "null output or negative size"; return ORBX_ERR_ARG; }
  if (img && w > 0 && hgt > 0 && stride < (size_t)w) { h->err = "stride < width"; return ORBX_ERR_ARG; }
  if (w > h->p.max_width || hgt > h->p.max_height) { h->err = "image larger than max_width x max_height"; return ORBX_ERR_ARG; }
  return ORBX_OK;
}

static int check_err_flag(orbx_handle* h);

extern "C" int orbx_synchronize(orbx_handle* h) {
  if (!h) return ORBX_ERR_ARG;
  CK(cudaSetDevice(h->p.device));
  CK(cudaStreamSynchronize(h->s_in));
  for (int li = 0; li < h->nlanes; li++) CK(cudaStreamSynchronize(h->lane[li].s_main));
  CK(cudaStreamSynchronize(h->s_out));
  CK(cudaMemcpyAsync(h->h_err, h->d_err, sizeof(int32_t), cudaMemcpyDeviceToHost, h->s_out));
  CK(cudaStreamSynchronize(h->s_out));
  return check_err_flag(h);
}

// the sticky device-side overflow flag, after its pinned mirror has been refreshed
static int check_err_flag(orbx_handle* h) {
  const int32_t flag = *h->h_err;
  if (flag) {
    cudaMemset(h->d_err, 0, sizeof(flag));
    h->err = flag == 1 ? "cell candidate list overflow" : (flag == 3 ? "FAST tile never arrived (TMA wait timed out)" : "level list overflow");
    return ORBX_ERR_OVERFLOW;
  }
  return ORBX_OK;
}

extern "C" int orbx_extract_batch_device(orbx_handle* h, const uint8_t* d_imgs, int nframes, int w, int hgt,
                                         size_t stride, size_t frame_stride, orbx_keypoint* d_kps, uint8_t* d_desc,
                                         int* d_n_out, void* stream) {
  int rc = check_args(h, d_imgs, nframes, w, hgt, stride, d_kps, d_desc, d_n_out);
  if (rc != ORBX_OK) return rc;
  CK(cudaSetDevice(h->p.device));
  cudaStream_t s = stream ? (cudaStream_t)stream : h->lane[0].s_main;
  if (!d_imgs || w == 0 || hgt == 0 || nframes == 0) {  // empty image: nothing extracted (:721-722)
    if (nframes > 0) CK(cudaMemsetAsync(d_n_out, 0, sizeof(int) * nframes, s));
    return ORBX_OK;
  }
  rc = upload_plan(h, w, hgt);
  if (rc != ORBX_OK) return rc;
  const int B = h->p.max_batch, nf = h->p.nfeatures;
  if (nframes <= B || h->profiling) {   // one chunk (or serial, timed stages): everything on the caller's stream
    for (int f0 = 0; f0 < nframes; f0 += B) {
      const int n = std::min(B, nframes - f0);
      rc = run_chunk(h, 0, d_imgs + (size_t)f0 * frame_stride, n, stride, frame_stride, d_kps + (size_t)f0 * nf,
                     d_desc + (size_t)f0 * nf * 32, d_n_out + f0, s);
      if (rc != ORBX_OK) return rc;
    }
    return ORBX_OK;
  }
  // several chunks: fork the caller's stream into the lanes, deal chunks round-robin, join at the end
  CK(cudaEventRecord(h->ev_user, s));
  for (int li = 0; li < h->nlanes; li++) CK(cudaStreamWaitEvent(h->lane[li].s_main, h->ev_user, 0));
  int k = 0;
  for (int f0 = 0; f0 < nframes; f0 += B, k++) {
    const int n = std::min(B, nframes - f0);
    rc = run_chunk(h, k % h->nlanes, d_imgs + (size_t)f0 * frame_stride, n, stride, frame_stride, d_kps + (size_t)f0 * nf,
                   d_desc + (size_t)f0 * nf * 32, d_n_out + f0, h->lane[k % h->nlanes].s_main);
    if (rc != ORBX_OK) return rc;
  }
  for (int li = 0; li < h->nlanes; li++) {
    orbx_handle::Lane& ln = h->lane[li];
    CK(cudaEventRecord(ln.ev_done, ln.s_main));
    CK(cudaStreamWaitEvent(s, ln.ev_done, 0));
  }
  return ORBX_OK;
}

// HOST buffers: chunk k's H2D (s_in), kernels (lane k % h->nlanes) and D2H (s_out) overlap with chunks k-1 / k+1 through kLanes
// staging slots.  Pinned caller memory makes the copies truly asynchronous; pageable memory still works.
extern "C" int orbx_extract_batch(orbx_handle* h, const uint8_t* imgs, int nframes, int w, int hgt, size_t stride,
                                  size_t frame_stride, orbx_keypoint* kps, uint8_t* desc, int* n_out) {
  int rc = check_args(h, imgs, nframes, w, hgt, stride, kps, desc, n_out);
  if (rc != ORBX_OK) return rc;
  if (!imgs || w == 0 || hgt == 0 || nframes == 0) {
    for (int f = 0; f < nframes; f++) n_out[f] = 0;
    return ORBX_OK;
  }
  CK(cudaSetDevice(h->p.device));
  rc = upload_plan(h, w, hgt);
  if (rc != ORBX_OK) return rc;
  const int B = h->p.max_batch, nf = h->p.nfeatures;
  const int pitch = (w + 15) / 16 * 16;
  const size_t fbytes = (size_t)pitch * hgt;
  int k = 0;
  for (int f0 = 0; f0 < nframes; f0 += B, k++) {
    const int n = std::min(B, nframes - f0), slot = k % h->nlanes;
    if (k >= h->nlanes) CK(cudaStreamWaitEvent(h->s_in, h->ev_compute[slot], 0));  // slot's previous input consumed
    const uint8_t* src = imgs + (size_t)f0 * frame_stride;
    if (frame_stride == stride * (size_t)hgt && stride == (size_t)pitch) {
      // rows already have the device pitch: one contiguous copy (a strided 2-D H2D costs a DMA descriptor per row)
      CK(cudaMemcpyAsync(h->d_in[slot], src, fbytes * n, cudaMemcpyHostToDevice, h->s_in));
    } else if (frame_stride == stride * (size_t)hgt && frame_stride <= (size_t)h->in_frame_bytes) {
      // tightly packed rows: contiguous H2D into the raw slot, then re-pitch on the device at HBM speed
      CK(cudaMemcpyAsync(h->d_raw[slot], src, frame_stride * n, cudaMemcpyHostToDevice, h->s_in));
      CK(cudaMemcpy2DAsync(h->d_in[slot], pitch, h->d_raw[slot], stride, w, (size_t)hgt * n, cudaMemcpyDeviceToDevice,
                           h->s_in));
    } else if (frame_stride == stride * (size_t)hgt) {
      CK(cudaMemcpy2DAsync(h->d_in[slot], pitch, src, stride, w, (size_t)hgt * n, cudaMemcpyHostToDevice, h->s_in));
    } else {
      for (int f = 0; f < n; f++)
        CK(cudaMemcpy2DAsync(h->d_in[slot] + f * fbytes, pitch, src + (size_t)f * frame_stride, stride, w, hgt,
                             cudaMemcpyHostToDevice, h->s_in));
    }
    CK(cudaEventRecord(h->ev_in[slot], h->s_in));
    cudaStream_t sc = h->lane[h->profiling ? 0 : slot].s_main;
    CK(cudaStreamWaitEvent(sc, h->ev_in[slot], 0));
    if (k >= h->nlanes) CK(cudaStreamWaitEvent(sc, h->ev_out[slot], 0));  // slot's previous output drained
    rc = run_chunk(h, h->profiling ? 0 : slot, h->d_in[slot], n, pitch, fbytes, h->d_out_kps[slot], h->d_out_desc[slot],
                   h->d_out_n[slot], sc);
    if (rc != ORBX_OK) return rc;
    CK(cudaEventRecord(h->ev_compute[slot], sc));
    CK(cudaStreamWaitEvent(h->s_out, h->ev_compute[slot], 0));
    CK(cudaMemcpyAsync(kps + (size_t)f0 * nf, h->d_out_kps[slot], sizeof(orbx_keypoint) * nf * n,
                       cudaMemcpyDeviceToHost, h->s_out));
    CK(cudaMemcpyAsync(desc + (size_t)f0 * nf * 32, h->d_out_desc[slot], (size_t)32 * nf * n, cudaMemcpyDeviceToHost,
                       h->s_out));
    CK(cudaMemcpyAsync(n_out + f0, h->d_out_n[slot], sizeof(int) * n, cudaMemcpyDeviceToHost, h->s_out));
    CK(cudaEventRecord(h->ev_out[slot], h->s_out));
  }
  // s_out is ordered behind every chunk's kernels (which are ordered behind their H2D): one copy of the error flag
  // behind the last outputs and ONE synchronisation end the call
  CK(cudaMemcpyAsync(h->h_err, h->d_err, sizeof(int32_t), cudaMemcpyDeviceToHost, h->s_out));
  CK(cudaStreamSynchronize(h->s_out));
  return check_err_flag(h);
}

extern "C" int orbx_extract(orbx_handle* h, const uint8_t* img, int w, int hgt, size_t stride, orbx_keypoint* kps,
                            uint8_t* desc, int* n_out) {
  return orbx_extract_batch(h, img, 1, w, hgt, stride, stride * (size_t)(hgt > 0 ? hgt : 0), kps, desc, n_out);
}

extern "C" int orbx_host_alloc(size_t bytes, int write_combined, void** out) {
  if (!out || !bytes) return ORBX_ERR_ARG;
  *out = nullptr;
  const cudaError_t e = cudaHostAlloc(out, bytes, write_combined ? cudaHostAllocWriteCombined : cudaHostAllocDefault);
  if (e != cudaSuccess) { g_create_err = std::string("cudaHostAlloc: ") + cudaGetErrorString(e); return ORBX_ERR_CUDA; }
  return ORBX_OK;
}
extern "C" void orbx_host_free(void* p) { if (p) cudaFreeHost(p); }

extern "C" int orbx_profile_enable(orbx_handle* h, int on) {
  if (!h) return ORBX_ERR_ARG;
  h->profiling = on != 0;
  return ORBX_OK;
}
extern "C" int orbx_profile_read(orbx_handle* h, float ms[ORBX_STAGE_COUNT], int64_t launches[ORBX_STAGE_COUNT]) {
  if (!h) return ORBX_ERR_ARG;
  for (int i = 0; i < ORBX_STAGE_COUNT; i++) {
    if (ms) ms[i] = h->prof_ms[i];
    if (launches) launches[i] = h->prof_launch[i];
    h->prof_ms[i] = 0; h->prof_launch[i] = 0;
  }
  return ORBX_OK;
}

#include "orbx_api_debug.inl"
