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
