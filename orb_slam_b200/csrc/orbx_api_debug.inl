// Included by orbx_api.cu: parity introspection getters (tests only) and the synthetic-frame utility.
extern "C" int orbx_debug_level_info(orbx_handle* h, int level, int32_t out8[8]) {
  if (!h || level < 0 || level >= h->hp.dp.nlevels) return ORBX_ERR_ARG;
  const LevelPlan& L = h->hp.dp.lv[level];
  const int32_t v[8] = {L.w, L.h, L.cols, L.rows, L.cellW, L.cellH, L.quota, L.nDesired};
  memcpy(out8, v, sizeof(v));
  return ORBX_OK;
}

extern "C" int orbx_debug_level_image(orbx_handle* h, int frame, int level, int which, uint8_t* out) {
  if (!h || !out || level < 0 || level >= h->hp.dp.nlevels || frame < 0 || frame >= h->last_frames) return ORBX_ERR_ARG;
  CK(cudaSetDevice(h->p.device));
  CK(cudaDeviceSynchronize());
  const DevicePlan& dp = h->hp.dp;
  const LevelPlan& L = dp.lv[level];
  const uint8_t* src;
  size_t pitch;
  if (which == 1) { src = h->last.blur + (int64_t)frame * dp.blur_frame_bytes + L.off; pitch = L.pitch; }
  else if (level == 0) { src = h->last.lv0 + (int64_t)frame * h->last.lv0_frame_stride; pitch = h->last.lv0_pitch; }
  else { src = h->last.pyr + (int64_t)frame * dp.pyr_frame_bytes + L.off; pitch = L.pitch; }
  CK(cudaMemcpy2D(out, L.w, src, pitch, L.w, L.h, cudaMemcpyDeviceToHost));
  return ORBX_OK;
}

extern "C" int orbx_debug_level_cells(orbx_handle* h, int frame, int level, int32_t* n_total, int32_t* n_retain) {
  if (!h || level < 0 || level >= h->hp.dp.nlevels || frame < 0 || frame >= h->last_frames) return ORBX_ERR_ARG;
  CK(cudaSetDevice(h->p.device));
  CK(cudaDeviceSynchronize());
  const DevicePlan& dp = h->hp.dp;
  const LevelPlan& L = dp.lv[level];
  const int64_t off = (int64_t)frame * dp.total_cells + L.cell_base;
  if (n_total) CK(cudaMemcpy(n_total, h->last.n_total + off, sizeof(int32_t) * L.nCells, cudaMemcpyDeviceToHost));
  if (n_retain) CK(cudaMemcpy(n_retain, h->last.n_retain + off, sizeof(int32_t) * L.nCells, cudaMemcpyDeviceToHost));
  return ORBX_OK;
}

// NOTE: after the select kernel the first nToRetain entries of a cell list have been reordered by nth_element; the
// SET of candidates is unchanged, so callers compare sorted lists (or call this on a handle whose quota is huge).
extern "C" int orbx_debug_cell_candidates(orbx_handle* h, int frame, int level, int cell, int32_t* xs, int32_t* ys,
                                          float* resp, int cap, int* n) {
  if (!h || !n || level < 0 || level >= h->hp.dp.nlevels || frame < 0 || frame >= h->last_frames) return ORBX_ERR_ARG;
  const DevicePlan& dp = h->hp.dp;
  const LevelPlan& L = dp.lv[level];
  if (cell < 0 || cell >= L.nCells) return ORBX_ERR_ARG;
  CK(cudaSetDevice(h->p.device));
  CK(cudaDeviceSynchronize());
  const CellPlan& c = h->hp.cells[L.cell_base + cell];
  int32_t cnt = 0;
  CK(cudaMemcpy(&cnt, h->last.n_total + (int64_t)frame * dp.total_cells + L.cell_base + cell, sizeof(cnt),
                cudaMemcpyDeviceToHost));
  *n = cnt;
  const int m = std::min(cnt, cap);
  std::vector<Cand> tmp(m);
  if (m) CK(cudaMemcpy(tmp.data(), h->last.cand + (int64_t)frame * dp.cand_per_frame + c.cand_off, sizeof(Cand) * m,
                       cudaMemcpyDeviceToHost));
  for (int i = 0; i < m; i++) { xs[i] = tmp[i].xy & 0xffff; ys[i] = tmp[i].xy >> 16; resp[i] = tmp[i].resp; }
  return ORBX_OK;
}

extern "C" int orbx_synth_frames_device(uint8_t* d_out, int nframes, int width, int height, size_t stride,
                                        size_t frame_stride, uint64_t seed0, void* stream) {
  if (!d_out || nframes < 0 || width < 8 || height < 8 || stride < (size_t)width) return ORBX_ERR_ARG;
  if ((size_t)(width * height / 768) * 5 * sizeof(int) > 48 * 1024) return ORBX_ERR_UNSUPPORTED;
  if (nframes == 0) return ORBX_OK;
  uint8_t* tmp = nullptr;
  const int chunk = std::min(nframes, 1024);
  if (cudaMalloc(&tmp, (size_t)chunk * width * height) != cudaSuccess) return ORBX_ERR_CUDA;
  cudaStream_t s = (cudaStream_t)stream;
  for (int f0 = 0; f0 < nframes; f0 += chunk) {
    const int n = std::min(chunk, nframes - f0);
    launch_synth(d_out + (size_t)f0 * frame_stride, tmp, n, width, height, stride, frame_stride, seed0 + f0, s);
  }
  cudaError_t e = cudaStreamSynchronize(s);
  cudaFree(tmp);
  return e == cudaSuccess && cudaGetLastError() == cudaSuccess ? ORBX_OK : ORBX_ERR_CUDA;
}

// KeyPointsFilter::retainBest(list, keep) + resize(keep) through the device port of std::nth_element, for nlists
// caller-made lists (list i = entries start[i] .. start[i+1]).  order[]: the original position (inside its list) of the
// record now at each place; only the first min(keep, size) of a list are the reference's output.  batch != 0 selects
// the four-keys-per-trip partition scans the shared-memory selection kernel uses.
extern "C" int orbx_debug_retain_best(orbx_handle* h, const float* resp, const int32_t* start, const int32_t* keep,
                                      int nlists, int retain_mode, int batch, int32_t* order) {
  if (!h || !resp || !start || !keep || !order || nlists < 0) return ORBX_ERR_ARG;
  if (nlists == 0) return ORBX_OK;
  CK(cudaSetDevice(h->p.device));
  const int total = start[nlists];
  constexpr int kPad = 4;
  std::vector<Cand> host((size_t)total + 2 * kPad);
  for (auto& c : host) { c.resp = 0.f; c.xy = 0xFFFFFFFFu; }
  std::vector<int32_t> st(nlists + 1);
  for (int i = 0; i <= nlists; i++) st[i] = start[i] + kPad;
  for (int i = 0; i < nlists; i++)
    for (int k = start[i]; k < start[i + 1]; k++) { host[kPad + k].resp = resp[k]; host[kPad + k].xy = (uint32_t)(k - start[i]); }
  Cand* d_v = nullptr; int32_t *d_st = nullptr, *d_keep = nullptr;
  auto cleanup = [&] { cudaFree(d_v); cudaFree(d_st); cudaFree(d_keep); };
  cudaError_t e = cudaMalloc(&d_v, host.size() * sizeof(Cand));
  if (e == cudaSuccess) e = cudaMalloc(&d_st, st.size() * 4);
  if (e == cudaSuccess) e = cudaMalloc(&d_keep, (size_t)nlists * 4);
  if (e == cudaSuccess) e = cudaMemcpy(d_v, host.data(), host.size() * sizeof(Cand), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(d_st, st.data(), st.size() * 4, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(d_keep, keep, (size_t)nlists * 4, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    launch_retain_best_debug(d_v, d_st, d_keep, nlists, retain_mode, batch, nullptr);
    h->launches++;
    e = cudaDeviceSynchronize();
  }
  if (e == cudaSuccess) e = cudaMemcpy(host.data(), d_v, host.size() * sizeof(Cand), cudaMemcpyDeviceToHost);
  cleanup();
  if (e != cudaSuccess) { h->err = std::string("orbx_debug_retain_best: ") + cudaGetErrorString(e); return ORBX_ERR_CUDA; }
  for (int k = 0; k < total; k++) order[k] = (int32_t)host[kPad + k].xy;
  return ORBX_OK;
}
