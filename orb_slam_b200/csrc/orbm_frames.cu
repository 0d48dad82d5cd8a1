// Device-resident frame sets and the 64x48 keypoint grid (M3).  include/orbm.h.
// Reference: grid binning src/Frame.cc:108-122 + PosInGrid :266-276; window query Frame::GetFeaturesInArea :199-264
// and KeyFrame::GetFeaturesInArea src/KeyFrame.cc:612-652.
#include <cstring>

#include "orbm_match.cuh"
#include "orbm_undistort.cuh"

using namespace orbm;

// ------------------------------------------------------------------------------------------------------------------
// grid_build: one warp per frame.  cell = ix * 48 + iy with ix = round((x - minX) * invW) (half away from zero, the
// C `round` the reference calls), keypoints outside [0,64) x [0,48) are dropped (:272-273).  CSR items keep the
// insertion order (ascending keypoint index) inside every cell; the CSR order over cells (ix outer, iy inner) is the
// visiting order of GetFeaturesInArea, so "position in the CSR" reproduces the reference's candidate order.
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32) grid_build_kernel(FrameSetView fs) {
  __shared__ int s_cnt[kGridCells + 1];
  const int f = blockIdx.x, lane = threadIdx.x;
  const int base = fs.frame_off[f], n = fs.frame_off[f + 1] - base;
  const float4 gp = fs.gridp[f];
  for (int c = lane; c <= kGridCells; c += 32) s_cnt[c] = 0;
  __syncwarp();
  for (int i = lane; i < n; i += 32) {
    const int c = grid_cell_of(fs.x[base + i], fs.y[base + i], gp);
    if (c >= 0) atomicAdd(&s_cnt[c], 1);
  }
  __syncwarp();
  // exclusive scan over 3072 cells: each lane owns 96 consecutive cells
  constexpr int per = kGridCells / 32;
  int sum = 0;
  for (int k = 0; k < per; k++) sum += s_cnt[lane * per + k];
  int incl = sum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += o;
  }
  int run = incl - sum;
  for (int k = 0; k < per; k++) {
    const int c = lane * per + k, v = s_cnt[c];
    s_cnt[c] = run;
    run += v;
  }
  if (lane == 31) s_cnt[kGridCells] = run;
  __syncwarp();
  int32_t* cs = fs.cell_start + (int64_t)f * (kGridCells + 1);
  for (int c = lane; c <= kGridCells; c += 32) cs[c] = s_cnt[c];
  __syncwarp();
  // stable placement: rounds of 32 keypoints in index order; inside a round, lanes of the same cell are ranked
  // by lane (= index) with __match_any_sync, then the cell cursor advances by the group size
  int32_t* items = fs.items + base;
  for (int i0 = 0; i0 < n; i0 += 32) {
    const int i = i0 + lane;
    const int c = i < n ? grid_cell_of(fs.x[base + i], fs.y[base + i], gp) : -1;
    const unsigned peers = __match_any_sync(0xffffffffu, c);
    if (c >= 0) {
      const int rank = __popc(peers & ((1u << lane) - 1));
      items[s_cnt[c] + rank] = i;
    }
    __syncwarp();
    if (c >= 0 && (peers & ((1u << lane) - 1)) == 0) s_cnt[c] += __popc(peers);  // group leader
    __syncwarp();
  }
}

// Frame::GetFeaturesInArea evaluated by one warp (test / introspection entry point)
__global__ void __launch_bounds__(32) features_in_area_kernel(FrameSetView fs, int f, float x, float y, float r,
                                                              int minl, int maxl, int32_t* out, int cap, int* n_out) {
  const int lane = threadIdx.x;
  const int base = fs.frame_off[f];
  const float4 gp = fs.gridp[f];
  const int32_t* cs = fs.cell_start + (int64_t)f * (kGridCells + 1);
  int cx0, cx1, cy0, cy1;
  int n = 0;
  if (area_cells(x, y, r, gp, &cx0, &cx1, &cy0, &cy1)) {
    for (int ix = cx0; ix <= cx1; ix++) {
      const int s0 = cs[ix * kGridRows + cy0], s1 = cs[ix * kGridRows + cy1 + 1];
      for (int sb = s0; sb < s1; sb += 32) {
        const int s = sb + lane;
        int k = -1;
        bool on = false;
        if (s < s1) {
          k = fs.items[base + s];
          on = area_accept(fs.x[base + k], fs.y[base + k], fs.octave[base + k], x, y, r, minl, maxl);
        }
        const unsigned bal = __ballot_sync(0xffffffffu, on);
        if (on) {
          const int pos = n + __popc(bal & ((1u << lane) - 1));
          if (pos < cap) out[pos] = k;
        }
        n += __popc(bal);
      }
    }
  }
  if (lane == 0) *n_out = n;
}

// ------------------------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------------------------
FrameSetView orbm::view_of(const orbm_frameset* s) {
  FrameSetView v;
  v.frame_off = s->d_frame_off; v.x = s->d_x; v.y = s->d_y; v.angle = s->d_angle; v.octave = s->d_octave;
  v.desc = reinterpret_cast<const uint4*>(s->d_desc); v.valid = s->d_valid; v.gridp = s->d_gridp;
  v.cell_start = s->d_cell_start; v.items = s->d_items;
  v.node_off = s->d_node_off; v.fvitem_off = s->d_fvitem_off; v.node_id = s->d_node_id;
  v.node_start = s->d_node_start; v.fv_items = s->d_fv_items;
  return v;
}

// (Gathering the host vectors in pinned staging first was tried: 134 instead of 111 us for two 1000-keypoint frames -
// the driver's own path for small pageable copies is faster than an extra host memcpy.)
template <class T>
static cudaError_t upload(orbm_handle* h, T** dptr, const std::vector<T>& v, cudaStream_t s) {
  const size_t bytes = std::max<size_t>(v.size(), 1) * sizeof(T);
  cudaError_t e = orbm_alloc(h, dptr, bytes + 16);
  if (e != cudaSuccess) return e;
  if (!v.empty()) e = cudaMemcpyAsync(*dptr, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, s);
  return e;
}

extern "C" int orbm_frameset_create(orbm_handle* h, const orbm_frame* frames, const orbm_featvec* fvs, int nframes,
                                    orbm_frameset** out) {
  if (!h) return ORBM_ERR_ARG;
  if (!frames || !out || nframes < 1) { h->err = "null frames / nframes < 1"; return ORBM_ERR_ARG; }
  *out = nullptr;
  ORBM_CK(cudaSetDevice(h->device));
  orbm_frameset* s = new orbm_frameset();
  s->h = h; s->nframes = nframes; s->has_fv = fvs != nullptr;
  s->frame_off.assign(nframes + 1, 0);
  for (int f = 0; f < nframes; f++) {
    const orbm_frame& F = frames[f];
    if (F.n < 0 || F.n > 65535 || (F.n > 0 && (!F.x || !F.y || !F.angle || !F.octave || !F.desc))) {
      h->err = "frame with null arrays or more than 65535 keypoints";
      delete s;
      return ORBM_ERR_ARG;
    }
    s->frame_off[f + 1] = s->frame_off[f] + F.n;
    s->max_n = std::max(s->max_n, F.n);
  }
  const int64_t N = s->total_kp = s->frame_off[nframes];
  std::vector<float> x(N), y(N), ang(N);
  std::vector<int32_t> oct(N);
  std::vector<uint8_t> desc((size_t)N * 32), valid(N, 0);
  std::vector<float4> gp(nframes);
  for (int f = 0; f < nframes; f++) {
    const orbm_frame& F = frames[f];
    const size_t o = s->frame_off[f], n = F.n;
    if (n) {
      memcpy(&x[o], F.x, n * 4); memcpy(&y[o], F.y, n * 4); memcpy(&ang[o], F.angle, n * 4);
      memcpy(&oct[o], F.octave, n * 4); memcpy(&desc[o * 32], F.desc, n * 32);
      if (F.valid) memcpy(&valid[o], F.valid, n);
    }
    gp[f] = make_float4(F.min_x, F.min_y, F.inv_w, F.inv_h);
  }
  std::vector<int32_t> node_id, node_start, fv_items;
  if (fvs) {
    s->node_off.assign(nframes + 1, 0);
    s->fvitem_off.assign(nframes + 1, 0);
    for (int f = 0; f < nframes; f++) {
      const orbm_featvec& V = fvs[f];
      const int nn = V.nnodes;
      if (nn < 0 || (nn > 0 && (!V.node_id || !V.start || !V.items)) || (nn > 0 && V.start[0] != 0)) {
        h->err = "malformed feature vector";
        delete s;
        return ORBM_ERR_ARG;
      }
      const int ni = nn ? V.start[nn] : 0;
      for (int k = 0; k < nn; k++) {
        if ((k && V.node_id[k] <= V.node_id[k - 1]) || V.start[k + 1] < V.start[k]) {
          h->err = "feature vector node ids must be strictly ascending with monotone starts";
          delete s;
          return ORBM_ERR_ARG;
        }
        node_id.push_back(V.node_id[k]);
      }
      for (int k = 0; k <= nn; k++) node_start.push_back(nn ? V.start[k] : 0);
      for (int k = 0; k < ni; k++) {
        if (V.items[k] < 0 || V.items[k] >= frames[f].n) { h->err = "feature vector item out of range"; delete s; return ORBM_ERR_ARG; }
        fv_items.push_back(V.items[k]);
      }
      s->node_off[f + 1] = s->node_off[f] + nn;
      s->fvitem_off[f + 1] = s->fvitem_off[f] + ni;
    }
    s->total_nodes = s->node_off[nframes];
    s->total_fv_items = s->fvitem_off[nframes];
  }
  auto fail = [&](cudaError_t e) {
    h->err = std::string("frameset upload: ") + cudaGetErrorString(e);
    orbm_frameset_destroy(s);
    return ORBM_ERR_CUDA;
  };
  cudaError_t e;
  cudaStream_t st = h->stream;
  if ((e = upload(h, &s->d_frame_off, s->frame_off, st)) != cudaSuccess) return fail(e);
  if ((e = upload(h, &s->d_x, x, st)) != cudaSuccess) return fail(e);
  if ((e = upload(h, &s->d_y, y, st)) != cudaSuccess) return fail(e);
  if ((e = upload(h, &s->d_angle, ang, st)) != cudaSuccess) return fail(e);
  if ((e = upload(h, &s->d_octave, oct, st)) != cudaSuccess) return fail(e);
  if ((e = upload(h, &s->d_desc, desc, st)) != cudaSuccess) return fail(e);
  if ((e = upload(h, &s->d_valid, valid, st)) != cudaSuccess) return fail(e);
  if ((e = upload(h, &s->d_gridp, gp, st)) != cudaSuccess) return fail(e);
  if ((e = orbm_alloc(h, &s->d_cell_start, (size_t)nframes * (kGridCells + 1) * 4)) != cudaSuccess) return fail(e);
  if ((e = orbm_alloc(h, &s->d_items, std::max<size_t>(N, 1) * 4)) != cudaSuccess) return fail(e);
  if (fvs) {
    if ((e = upload(h, &s->d_node_off, s->node_off, st)) != cudaSuccess) return fail(e);
    if ((e = upload(h, &s->d_fvitem_off, s->fvitem_off, st)) != cudaSuccess) return fail(e);
    if ((e = upload(h, &s->d_node_id, node_id, st)) != cudaSuccess) return fail(e);
    if ((e = upload(h, &s->d_node_start, node_start, st)) != cudaSuccess) return fail(e);
    if ((e = upload(h, &s->d_fv_items, fv_items, st)) != cudaSuccess) return fail(e);
  }
  grid_build_kernel<<<nframes, 32, 0, st>>>(view_of(s));
  h->launches++;
  if ((e = cudaGetLastError()) != cudaSuccess) return fail(e);
  if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return fail(e);  // host vectors die here
  *out = s;
  return ORBM_OK;
}

// ------------------------------------------------------------------------------------------------------------------
// N1  Frame glue on the device (reference src/Frame.cc:59-127): extractor output -> frame set, no host round trip of
// the features.  One CTA per frame: AoS cv::KeyPoint records -> SoA, UndistortKeyPoints (:288-318) in double with
// individually rounded operations (orbm_undistort.cuh), descriptors gathered to the packed layout.
// ------------------------------------------------------------------------------------------------------------------
struct KeyPoint28 { float x, y, size, angle, response; int32_t octave, class_id; };  // == cv::KeyPoint == orbx_keypoint

__global__ void __launch_bounds__(256) frame_glue_kernel(const KeyPoint28* kps, const uint8_t* desc, int kp_stride,
                                                         const int32_t* frame_off, Camera cam, int undistort,
                                                         float* ox, float* oy, float* oangle, int32_t* ooct,
                                                         uint4* odesc, uint8_t* ovalid) {
  const int f = blockIdx.x;
  const int base = frame_off[f], n = frame_off[f + 1] - base;
  const KeyPoint28* kp = kps + (int64_t)f * kp_stride;
  const uint4* d = reinterpret_cast<const uint4*>(desc + (int64_t)f * kp_stride * 32);
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const KeyPoint28 k = kp[i];
    float x = k.x, y = k.y;
    if (undistort) undistort_point(cam, k.x, k.y, &x, &y);
    ox[base + i] = x; oy[base + i] = y; oangle[base + i] = k.angle; ooct[base + i] = k.octave;
    ovalid[base + i] = 0;
    odesc[2 * (int64_t)(base + i)] = d[2 * i];
    odesc[2 * (int64_t)(base + i) + 1] = d[2 * i + 1];
  }
}

static Camera camera_of(const orbm_camera& c) {
  Camera k;
  k.fx = c.fx; k.fy = c.fy; k.cx = c.cx; k.cy = c.cy; k.k1 = c.k1; k.k2 = c.k2; k.p1 = c.p1; k.p2 = c.p2; k.k3 = c.k3;
  return k;
}

// Frame::ComputeImageBounds (:320-349) + grid element sizes (:75-76); four points, on the host
extern "C" int orbm_image_bounds(const orbm_camera* cam, int32_t bounds[4], float inv[2]) {
  if (!cam || !bounds || !inv || cam->width <= 0 || cam->height <= 0) return ORBM_ERR_ARG;
  if (cam->k1 != 0.0f) {
    const Camera k = camera_of(*cam);
    const float cx[4] = {0.f, (float)cam->width, 0.f, (float)cam->width};
    const float cy[4] = {0.f, 0.f, (float)cam->height, (float)cam->height};
    float ux[4], uy[4];
    for (int i = 0; i < 4; i++) undistort_point(k, cx[i], cy[i], &ux[i], &uy[i]);
    bounds[0] = (int32_t)fminf(floorf(ux[0]), floorf(ux[2]));
    bounds[1] = (int32_t)fmaxf(ceilf(ux[1]), ceilf(ux[3]));
    bounds[2] = (int32_t)fminf(floorf(uy[0]), floorf(uy[1]));
    bounds[3] = (int32_t)fmaxf(ceilf(uy[2]), ceilf(uy[3]));
  } else {
    bounds[0] = 0; bounds[1] = cam->width; bounds[2] = 0; bounds[3] = cam->height;
  }
  if (bounds[1] <= bounds[0] || bounds[3] <= bounds[2]) return ORBM_ERR_ARG;
  inv[0] = (float)kGridCols / (float)(bounds[1] - bounds[0]);
  inv[1] = (float)kGridRows / (float)(bounds[3] - bounds[2]);
  return ORBM_OK;
}

extern "C" int orbm_frameset_from_extractor(orbm_handle* h, const void* d_kps, const uint8_t* d_desc,
                                            const int32_t* d_n, int nframes, int kp_stride, const orbm_camera* cam,
                                            void* stream, orbm_frameset** out) {
  if (!h) return ORBM_ERR_ARG;
  if (!d_kps || !d_desc || !d_n || !cam || !out || nframes < 1 || kp_stride < 1 || kp_stride > 65535) {
    h->err = "null device arrays / bad sizes";
    return ORBM_ERR_ARG;
  }
  *out = nullptr;
  int32_t bounds[4];
  float inv[2];
  if (orbm_image_bounds(cam, bounds, inv) != ORBM_OK) { h->err = "bad camera (image size / degenerate bounds)"; return ORBM_ERR_ARG; }
  ORBM_CK(cudaSetDevice(h->device));
  cudaStream_t st = h->stream;
  // the extractor wrote its outputs on the caller's stream: order this handle's stream after it
  if (!h->ev_order) ORBM_CK(cudaEventCreateWithFlags(&h->ev_order, cudaEventDisableTiming));
  ORBM_CK(cudaEventRecord(h->ev_order, (cudaStream_t)stream));
  ORBM_CK(cudaStreamWaitEvent(st, h->ev_order, 0));
  // small frame sets (a tracker's one frame per call): counts come back and the two host-built arrays go up through
  // pinned staging - a pageable copy costs a staging round trip of its own
  ORBM_CK(orbm_ensure_pinned(h));
  const bool staged = (size_t)(nframes + 1) * 4 + (size_t)nframes * 20 + 64 <= kOrbmPinnedBytes;
  uint8_t* const stg = static_cast<uint8_t*>(h->pinned);
  std::vector<int32_t> cnt(nframes);
  if (staged) {
    ORBM_CK(cudaMemcpyAsync(stg, d_n, (size_t)nframes * 4, cudaMemcpyDeviceToHost, st));          // 4 bytes per frame
    ORBM_CK(cudaStreamSynchronize(st));
    memcpy(cnt.data(), stg, (size_t)nframes * 4);
  } else {
    ORBM_CK(cudaMemcpyAsync(cnt.data(), d_n, (size_t)nframes * 4, cudaMemcpyDeviceToHost, st));
    ORBM_CK(cudaStreamSynchronize(st));
  }
  orbm_frameset* s = new orbm_frameset();
  s->h = h; s->nframes = nframes; s->has_fv = false;
  s->frame_off.assign(nframes + 1, 0);
  for (int f = 0; f < nframes; f++) {
    if (cnt[f] < 0 || cnt[f] > kp_stride) { h->err = "keypoint count outside [0, kp_stride]"; delete s; return ORBM_ERR_ARG; }
    s->frame_off[f + 1] = s->frame_off[f] + cnt[f];
    s->max_n = std::max(s->max_n, cnt[f]);
  }
  const int64_t N = s->total_kp = s->frame_off[nframes];
  auto fail = [&](cudaError_t e) {
    h->err = std::string("frameset from extractor: ") + cudaGetErrorString(e);
    orbm_frameset_destroy(s);
    return ORBM_ERR_CUDA;
  };
  cudaError_t e;
  const size_t n1 = std::max<size_t>(N, 1);
  std::vector<float4> gp(nframes, make_float4((float)bounds[0], (float)bounds[2], inv[0], inv[1]));
  if (staged) {   // the staging area stays untouched until the synchronisation that ends this call
    const size_t ob = (size_t)(nframes + 1) * 4, gb = (size_t)nframes * sizeof(float4), go = (ob + 15) & ~(size_t)15;
    memcpy(stg, s->frame_off.data(), ob);
    memcpy(stg + go, gp.data(), gb);
    if ((e = orbm_alloc(h, &s->d_frame_off, ob + 16)) != cudaSuccess) return fail(e);
    if ((e = orbm_alloc(h, &s->d_gridp, gb + 16)) != cudaSuccess) return fail(e);
    if ((e = cudaMemcpyAsync(s->d_frame_off, stg, ob, cudaMemcpyHostToDevice, st)) != cudaSuccess) return fail(e);
    if ((e = cudaMemcpyAsync(s->d_gridp, stg + go, gb, cudaMemcpyHostToDevice, st)) != cudaSuccess) return fail(e);
  } else {
    if ((e = upload(h, &s->d_frame_off, s->frame_off, st)) != cudaSuccess) return fail(e);
    if ((e = upload(h, &s->d_gridp, gp, st)) != cudaSuccess) return fail(e);
  }
  if ((e = orbm_alloc(h, &s->d_x, n1 * 4 + 16)) != cudaSuccess) return fail(e);
  if ((e = orbm_alloc(h, &s->d_y, n1 * 4 + 16)) != cudaSuccess) return fail(e);
  if ((e = orbm_alloc(h, &s->d_angle, n1 * 4 + 16)) != cudaSuccess) return fail(e);
  if ((e = orbm_alloc(h, &s->d_octave, n1 * 4 + 16)) != cudaSuccess) return fail(e);
  if ((e = orbm_alloc(h, &s->d_desc, n1 * 32 + 16)) != cudaSuccess) return fail(e);
  if ((e = orbm_alloc(h, &s->d_valid, n1 + 16)) != cudaSuccess) return fail(e);
  if ((e = orbm_alloc(h, &s->d_cell_start, (size_t)nframes * (kGridCells + 1) * 4)) != cudaSuccess) return fail(e);
  if ((e = orbm_alloc(h, &s->d_items, n1 * 4)) != cudaSuccess) return fail(e);
  frame_glue_kernel<<<nframes, 256, 0, st>>>(static_cast<const KeyPoint28*>(d_kps), d_desc, kp_stride, s->d_frame_off,
                                             camera_of(*cam), cam->k1 != 0.0f ? 1 : 0, s->d_x, s->d_y, s->d_angle,
                                             s->d_octave, reinterpret_cast<uint4*>(s->d_desc), s->d_valid);
  grid_build_kernel<<<nframes, 32, 0, st>>>(view_of(s));
  h->launches += 2;
  if ((e = cudaGetLastError()) != cudaSuccess) return fail(e);
  if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return fail(e);  // host vectors die here
  *out = s;
  return ORBM_OK;
}

// introspection: per-frame keypoint counts and the SoA columns of one frame (any of the outputs may be NULL)
extern "C" int orbm_frameset_counts(orbm_frameset* s, int32_t* n_out) {
  if (!s || !n_out) return ORBM_ERR_ARG;
  for (int f = 0; f < s->nframes; f++) n_out[f] = s->frame_off[f + 1] - s->frame_off[f];
  return ORBM_OK;
}

extern "C" int orbm_frameset_keypoints(orbm_frameset* s, int frame, float* x, float* y, float* angle, int32_t* octave,
                                       uint8_t* desc) {
  if (!s) return ORBM_ERR_ARG;
  orbm_handle* h = s->h;
  if (frame < 0 || frame >= s->nframes) { h->err = "frame index out of range"; return ORBM_ERR_ARG; }
  ORBM_CK(cudaSetDevice(h->device));
  const size_t o = s->frame_off[frame], n = s->frame_off[frame + 1] - o;
  if (!n) return ORBM_OK;
  if (x) ORBM_CK(cudaMemcpy(x, s->d_x + o, n * 4, cudaMemcpyDeviceToHost));
  if (y) ORBM_CK(cudaMemcpy(y, s->d_y + o, n * 4, cudaMemcpyDeviceToHost));
  if (angle) ORBM_CK(cudaMemcpy(angle, s->d_angle + o, n * 4, cudaMemcpyDeviceToHost));
  if (octave) ORBM_CK(cudaMemcpy(octave, s->d_octave + o, n * 4, cudaMemcpyDeviceToHost));
  if (desc) ORBM_CK(cudaMemcpy(desc, s->d_desc + o * 32, n * 32, cudaMemcpyDeviceToHost));
  return ORBM_OK;
}

extern "C" void orbm_frameset_destroy(orbm_frameset* s) {
  if (!s) return;
  cudaSetDevice(s->h->device);
  cudaStreamSynchronize(s->h->stream);
  void* ptrs[] = {s->d_frame_off, s->d_x, s->d_y, s->d_angle, s->d_octave, s->d_desc, s->d_valid, s->d_gridp,
                  s->d_cell_start, s->d_items, s->d_node_off, s->d_fvitem_off, s->d_node_id, s->d_node_start,
                  s->d_fv_items, s->d_bow_word, s->d_bow_val};
  for (void* p : ptrs) orbm_free(s->h, p);
  delete s;
}

extern "C" int orbm_frameset_set_valid(orbm_frameset* s, int frame, const uint8_t* valid) {
  if (!s) return ORBM_ERR_ARG;
  orbm_handle* h = s->h;
  if (frame < 0 || frame >= s->nframes) { h->err = "frame index out of range"; return ORBM_ERR_ARG; }
  ORBM_CK(cudaSetDevice(h->device));
  const int o = s->frame_off[frame], n = s->frame_off[frame + 1] - o;
  if (n == 0) return ORBM_OK;
  if (valid) ORBM_CK(cudaMemcpyAsync(s->d_valid + o, valid, n, cudaMemcpyHostToDevice, h->stream));
  else ORBM_CK(cudaMemsetAsync(s->d_valid + o, 0, n, h->stream));
  ORBM_CK(cudaStreamSynchronize(h->stream));
  return ORBM_OK;
}

extern "C" int orbm_frameset_fill_valid(orbm_frameset* s, int value) {
  if (!s) return ORBM_ERR_ARG;
  orbm_handle* h = s->h;
  ORBM_CK(cudaSetDevice(h->device));
  if (s->total_kp) ORBM_CK(cudaMemsetAsync(s->d_valid, value ? 1 : 0, (size_t)s->total_kp, h->stream));
  return ORBM_OK;
}

extern "C" int orbm_frameset_grid(orbm_frameset* s, int frame, int32_t* cell_start, int32_t* items) {
  if (!s) return ORBM_ERR_ARG;
  orbm_handle* h = s->h;
  if (frame < 0 || frame >= s->nframes || !cell_start || !items) { h->err = "bad grid query"; return ORBM_ERR_ARG; }
  ORBM_CK(cudaSetDevice(h->device));
  ORBM_CK(cudaMemcpy(cell_start, s->d_cell_start + (size_t)frame * (kGridCells + 1), (kGridCells + 1) * 4,
                     cudaMemcpyDeviceToHost));
  const int n = cell_start[kGridCells];
  if (n) ORBM_CK(cudaMemcpy(items, s->d_items + s->frame_off[frame], (size_t)n * 4, cudaMemcpyDeviceToHost));
  return ORBM_OK;
}

extern "C" int orbm_features_in_area(orbm_frameset* s, int frame, float x, float y, float r, int min_level,
                                     int max_level, int32_t* out, int cap, int* n_out) {
  if (!s) return ORBM_ERR_ARG;
  orbm_handle* h = s->h;
  if (frame < 0 || frame >= s->nframes || !out || !n_out || cap < 0) { h->err = "bad area query"; return ORBM_ERR_ARG; }
  ORBM_CK(cudaSetDevice(h->device));
  ORBM_CK(h->ws_io.reserve((size_t)cap * 4 + 64));
  int32_t* d_out = h->ws_io.as<int32_t>() + 16;
  int* d_n = h->ws_io.as<int>();
  features_in_area_kernel<<<1, 32, 0, h->stream>>>(view_of(s), frame, x, y, r, min_level, max_level, d_out, cap, d_n);
  h->launches++;
  ORBM_CK(cudaGetLastError());
  ORBM_CK(cudaMemcpyAsync(n_out, d_n, 4, cudaMemcpyDeviceToHost, h->stream));
  ORBM_CK(cudaStreamSynchronize(h->stream));
  if (*n_out > cap) { h->err = "output capacity too small"; return ORBM_ERR_OVERFLOW; }
  if (*n_out) ORBM_CK(cudaMemcpy(out, d_out, (size_t)*n_out * 4, cudaMemcpyDeviceToHost));
  return ORBM_OK;
}
