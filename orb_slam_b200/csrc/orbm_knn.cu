// Matcher handle + M1 (DescriptorDistance) + M8 (all-pairs 2-NN) kernels and entry points.  include/orbm.h.
#include <dlfcn.h>

#include <algorithm>
#include <climits>
#include <cuda_fp16.h>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>

#include "orbm_internal.cuh"

using namespace orbm;

static thread_local std::string g_orbm_create_err;

// ------------------------------------------------------------------------------------------------------------------
// handle
// ------------------------------------------------------------------------------------------------------------------
extern "C" int orbm_create(int device, orbm_handle** out) {
  if (!out) return ORBM_ERR_ARG;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
    g_orbm_create_err = "no usable CUDA device (this library has no CPU fallback)";
    return ORBM_ERR_CUDA;
  }
  orbm_handle* h = new orbm_handle();
  h->device = device;
  auto body = [&]() -> int {
    ORBM_CK(cudaSetDevice(device));
    ORBM_CK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    ORBM_CK(cudaEventCreate(&h->ev0));
    ORBM_CK(cudaEventCreate(&h->ev1));
    cudaDeviceProp prop;
    ORBM_CK(cudaGetDeviceProperties(&prop, device));
    h->num_sms = prop.multiProcessorCount;
    // run-time switches are per handle (a test can create one handle per setting inside one process)
    ORBM_CK(h->ws_misc.reserve(256));
    ORBM_CK(cudaMemset(h->ws_misc.p, 0, 256));          // [0] sticky knn2 error flag, [2..3] search arena cursor
    if (const char* e = getenv("ORBM_COMMIT")) h->commit_mode = e[0] == 's' ? 1 : (e[0] == 'p' ? 2 : 0);
    if (const char* e = getenv("ORBM_KNN_VARIANT")) h->knn_variant = std::min(2, std::max(0, atoi(e)));
    if (const char* e = getenv("ORBM_KNN_TC")) h->knn_tc = atoi(e) ? 1 : 0;
    // frame-set arrays are pool allocations (orbm_alloc) from a pool of the handle's own: freed blocks stay cached in it
    cudaMemPoolProps props = {};
    props.allocType = cudaMemAllocationTypePinned;
    props.handleTypes = cudaMemHandleTypeNone;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = device;
    ORBM_CK(cudaMemPoolCreate(&h->pool, &props));
    unsigned long long keep = ~0ull;
    ORBM_CK(cudaMemPoolSetAttribute(h->pool, cudaMemPoolAttrReleaseThreshold, &keep));
    return ORBM_OK;
  };
  int rc = body();
  if (rc != ORBM_OK) { g_orbm_create_err = h->err; orbm_destroy(h); return rc; }
  *out = h;
  return ORBM_OK;
}

extern "C" void orbm_destroy(orbm_handle* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  if (h->stream) cudaStreamSynchronize(h->stream);
  DevBuf* bufs[] = {&h->knn_part, &h->stage_a, &h->stage_b, &h->stage_c, &h->ws_problem, &h->ws_query,
                    &h->ws_target, &h->ws_cand, &h->ws_misc, &h->ws_io, &h->ws_bow};
  for (DevBuf* b : bufs) b->release();
  if (h->pinned) cudaFreeHost(h->pinned);
  if (h->pinned_big) cudaFreeHost(h->pinned_big);
  if (h->ev_order) cudaEventDestroy(h->ev_order);
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  if (h->stream) cudaStreamDestroy(h->stream);
  if (h->pool) cudaMemPoolDestroy(h->pool);       // frame sets still alive lose their memory with their handle
  delete h;
}

extern "C" const char* orbm_last_error(const orbm_handle* h) { return h ? h->err.c_str() : g_orbm_create_err.c_str(); }
extern "C" int64_t orbm_kernel_launches(const orbm_handle* h) { return h ? h->launches : 0; }
extern "C" float orbm_last_kernel_ms(const orbm_handle* h) { return h ? h->last_ms : 0.f; }

// ------------------------------------------------------------------------------------------------------------------
// M1  batched DescriptorDistance: one thread per pair, two 16-byte loads per descriptor
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) distance_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, int n,
                                                       int32_t* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[i] = hamming256(a[2 * i], a[2 * i + 1], b[2 * i], b[2 * i + 1]);
}

extern "C" int orbm_descriptor_distance(orbm_handle* h, const uint8_t* a, const uint8_t* b, int n, int32_t* dist) {
  if (!h) return ORBM_ERR_ARG;
  if (!a || !b || !dist || n < 0) { h->err = "null pointer or negative n"; return ORBM_ERR_ARG; }
  if (n == 0) return ORBM_OK;
  ORBM_CK(cudaSetDevice(h->device));
  const size_t bytes = (size_t)n * 32;
  ORBM_CK(h->stage_a.reserve(bytes));
  ORBM_CK(h->stage_b.reserve(bytes));
  ORBM_CK(h->stage_c.reserve((size_t)n * 4));
  ORBM_CK(cudaMemcpyAsync(h->stage_a.p, a, bytes, cudaMemcpyHostToDevice, h->stream));
  ORBM_CK(cudaMemcpyAsync(h->stage_b.p, b, bytes, cudaMemcpyHostToDevice, h->stream));
  distance_kernel<<<(n + 255) / 256, 256, 0, h->stream>>>(h->stage_a.as<uint4>(), h->stage_b.as<uint4>(), n,
                                                         h->stage_c.as<int32_t>());
  h->launches++;
  ORBM_CK(cudaGetLastError());
  ORBM_CK(cudaMemcpyAsync(dist, h->stage_c.p, (size_t)n * 4, cudaMemcpyDeviceToHost, h->stream));
  ORBM_CK(cudaStreamSynchronize(h->stream));
  return ORBM_OK;
}

#include "orbm_knn_kernels.inl"
#include "orbm_knn_tc.inl"

namespace {

// Work split of knn2_tc_kernel: (query-tile pairs) x (database tiles) tile steps, linearised pair-major, `per` consecutive
// steps per cluster, one cluster per SM pair; `pieces` bounds how many clusters can share one query-tile pair (rows of the
// partial-result buffer).  Host logic, covered without a GPU by tests/test_knn_tc_plan.py through orbm_debug_knn2_tc_plan.
void knn_tc_plan(int nq, int64_t ndb, int num_sms, int* npairs, int* ntiles, int* clusters, int64_t* per, int* pieces) {
  *npairs = ((nq + kTcQ - 1) / kTcQ + 1) / 2;
  *ntiles = (int)((ndb + kTcN - 1) / kTcN);
  const int64_t total = (int64_t)*npairs * *ntiles;          // tile steps, shared out evenly
  int ncl = (int)std::max<int64_t>(1, std::min<int64_t>(num_sms / 2, total));
  *per = (total + ncl - 1) / ncl;
  *clusters = (int)((total + *per - 1) / *per);
  *pieces = (int)((*ntiles + *per - 1) / *per) + 1;
}

int knn2_launch(orbm_handle* h, const uint8_t* d_q, int nq, const uint8_t* d_db, int64_t ndb, int64_t idx_base,
                uint32_t* d_top2, cudaStream_t s) {
  if (!d_q || !d_db || !d_top2 || nq < 0 || ndb < 0 || idx_base < 0) {
    h->err = "null pointer or negative size";
    return ORBM_ERR_ARG;
  }
  if (idx_base + ndb > (1 << 23)) {
    h->err = "idx_base + ndb exceeds the 2^23 index range of the packed keys";
    return ORBM_ERR_UNSUPPORTED;
  }
  if (((uintptr_t)d_q | (uintptr_t)d_db) & 15) { h->err = "descriptor arrays must be 16-byte aligned"; return ORBM_ERR_ARG; }
  // sticky: set by a CTA whose mbarrier wait timed out (knn2_kernel: a TMA tile never arrived; knn2_tc_kernel: a stage /
  // accumulator signal), read and cleared by orbm_knn2 / orbm_synchronize
  int* d_err = h->ws_misc.as<int>();
  if (nq == 0) return ORBM_OK;
  if (ndb == 0) { ORBM_CK(cudaMemsetAsync(d_top2, 0xFF, (size_t)nq * 8, s)); return ORBM_OK; }
  // Jobs of a few million pairs and more go to the tensor cores (knn2_tc_kernel: 5.5 T pairs/s against 0.73 T of the
  // POPC kernels at 8,000 x 5 M; level at 1,000 x 1,000, where both are launch latency).  ORBM_KNN_TC=0|1 forces one.
  const bool tc = h->knn_tc < 0 ? (int64_t)nq * ndb >= (4ll << 20) : h->knn_tc == 1;
  if (tc) {
    int npairs, ntiles, ncl, pieces;
    int64_t per;
    knn_tc_plan(nq, ndb, h->num_sms, &npairs, &ntiles, &ncl, &per, &pieces);
    const int64_t total = (int64_t)npairs * ntiles;
    ORBM_CK(h->knn_part.reserve((size_t)pieces * nq * 8 + 16));
    if (!h->knn_tc_ready) {
      ORBM_CK(cudaFuncSetAttribute(knn2_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes));
      h->knn_tc_ready = true;
    }
    uint32_t* part = h->knn_part.as<uint32_t>();
    ORBM_CK(cudaMemsetAsync(part, 0xFF, (size_t)pieces * nq * 8, s));   // pieces a pair does not have stay 'none'
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * ncl);
    cfg.blockDim = dim3(kTcThreads);
    cfg.dynamicSmemBytes = kTcSmemBytes;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    const cudaError_t le = cudaLaunchKernelEx(&cfg, knn2_tc_kernel, reinterpret_cast<const uint4*>(d_q), nq,
                                              reinterpret_cast<const uint4*>(d_db), ndb, ntiles, per, total,
                                              (uint32_t)idx_base, part, d_err);
    if (le == cudaSuccess) {
      merge_top2_kernel<<<(nq + 255) / 256, 256, 0, s>>>(reinterpret_cast<const uint2*>(part), pieces, nq, 0, 0.f,
                                                        reinterpret_cast<uint2*>(d_top2), nullptr);
      h->launches += 2;
      ORBM_CK(cudaGetLastError());
      return ORBM_OK;
    }
    // A device or partition that cannot place the CTA pairs (cluster launch refused): when the tensor-core kernel was
    // this handle's own choice, the POPC kernels below take the job - and every later one - instead; when it was forced
    // (ORBM_KNN_TC=1) the refusal is the caller's to see.
    (void)cudaGetLastError();
    if (h->knn_tc == 1) { h->err = std::string("knn2_tc_kernel launch refused: ") + cudaGetErrorString(le); return ORBM_ERR_CUDA; }
    h->knn_tc = 0;
  }
  // chunks: about 2 CTAs per SM over the whole grid; each chunk a whole number of tiles and >= 8 tiles (the TMA ring
  // needs a few tiles to reach its steady state).  A database too small to give every SM a CTA that way (frame-to-frame
  // matching: 1000 x 1000) is latency bound instead: one query per thread and single-tile chunks, which spreads the
  // same pairs over up to 8 x more CTAs (1000 x 1000: 1 CTA -> 32, 246 -> ~40 us).
  int qpt = nq > 2 * kKnnThreads ? 4 : (nq > kKnnThreads ? 2 : 1);
  int qtiles = 0, nchunks = 0;
  int64_t chunk_len = 0;
  auto plan = [&](int min_tiles) {
    qtiles = (nq + kKnnThreads * qpt - 1) / (kKnnThreads * qpt);
    const int64_t want = std::max<int64_t>(1, (2 * (int64_t)h->num_sms + qtiles - 1) / qtiles);
    chunk_len = (ndb + want - 1) / want;
    chunk_len = std::max<int64_t>((chunk_len + kKnnTile - 1) / kKnnTile * kKnnTile, (int64_t)min_tiles * kKnnTile);
    nchunks = (int)((ndb + chunk_len - 1) / chunk_len);
  };
  plan(8);
  if ((int64_t)nchunks * qtiles < h->num_sms) { qpt = 1; plan(1); }
  ORBM_CK(h->knn_part.reserve((size_t)nchunks * nq * 8 + 16));
  dim3 grid(nchunks, qtiles);
  const uint4* q4 = reinterpret_cast<const uint4*>(d_q);
  const uint4* db4 = reinterpret_cast<const uint4*>(d_db);
  uint32_t* part = h->knn_part.as<uint32_t>();
  const int cl = (int)chunk_len;
  const uint32_t ib = (uint32_t)idx_base;
  const int dv = h->knn_variant;
#define ORBM_KNN_LAUNCH(Q, D) knn2_kernel<Q, D><<<grid, kKnnThreads, 0, s>>>(q4, nq, db4, ndb, cl, ib, part, d_err)
#define ORBM_KNN_DV(Q) (dv == 0 ? ORBM_KNN_LAUNCH(Q, 0) : dv == 1 ? ORBM_KNN_LAUNCH(Q, 1) : ORBM_KNN_LAUNCH(Q, 2))
  if (qpt == 4) ORBM_KNN_DV(4);
  else if (qpt == 2) ORBM_KNN_DV(2);
  else ORBM_KNN_DV(1);
  merge_top2_kernel<<<(nq + 255) / 256, 256, 0, s>>>(reinterpret_cast<const uint2*>(part), nchunks, nq, 0, 0.f,
                                                    reinterpret_cast<uint2*>(d_top2), nullptr);
  h->launches += 2;
  ORBM_CK(cudaGetLastError());
  return ORBM_OK;
}

int knn_check_err(orbm_handle* h) {  // call after the stream has been synchronised
  int flag = 0;
  ORBM_CK(cudaMemcpy(&flag, h->ws_misc.p, sizeof(int), cudaMemcpyDeviceToHost));
  if (flag) {
    cudaMemset(h->ws_misc.p, 0, sizeof(int));
    h->err = "knn2: an mbarrier wait timed out (a TMA tile or a tensor-pipe signal never arrived); the affected queries hold 'none' keys";
    return ORBM_ERR_CUDA;
  }
  return ORBM_OK;
}

int nccl_check_async(orbm_handle* h, void* comm);

}  // namespace

extern "C" int orbm_knn2_device(orbm_handle* h, const uint8_t* d_q, int nq, const uint8_t* d_db, int64_t ndb,
                                int64_t idx_base, uint32_t* d_top2, void* stream) {
  if (!h) return ORBM_ERR_ARG;
  ORBM_CK(cudaSetDevice(h->device));
  return knn2_launch(h, d_q, nq, d_db, ndb, idx_base, d_top2, stream ? (cudaStream_t)stream : h->stream);
}

extern "C" int orbm_debug_knn2_tc_plan(int nq, int64_t ndb, int num_sms, int64_t* out5) {
  if (nq < 1 || ndb < 1 || num_sms < 2 || !out5) return ORBM_ERR_ARG;
  int npairs, ntiles, ncl, pieces;
  int64_t per;
  knn_tc_plan(nq, ndb, num_sms, &npairs, &ntiles, &ncl, &per, &pieces);
  out5[0] = npairs; out5[1] = ntiles; out5[2] = ncl; out5[3] = per; out5[4] = pieces;
  return ORBM_OK;
}

// ---- databases beyond the 2^23 indices of the packed 32-bit keys: 64-bit keys (dist << 40) | index ----------------------
// The database is searched in segments of 2^23 descriptors with the 32-bit kernels (segment-local indices); one small
// kernel per segment widens the keys with the segment's base and folds them into the running (best, second) pair.
// Key order is still (distance, index) order, so the result equals one sequential scan with strict `<`.
namespace {
__global__ void __launch_bounds__(256) widen_merge_kernel(const uint2* __restrict__ seg, int nq, uint64_t base, int first,
                                                          ulonglong2* __restrict__ io) {
  const int qi = blockIdx.x * blockDim.x + threadIdx.x;
  if (qi >= nq) return;
  ulonglong2 r = first ? make_ulonglong2(~0ull, ~0ull) : io[qi];
  const uint2 v = seg[qi];
  const uint32_t ks[2] = {v.x, v.y};
#pragma unroll
  for (int k = 0; k < 2; k++) {
    if (ks[k] == kNone) continue;
    const uint64_t key = ((uint64_t)(ks[k] >> 23) << 40) | (base + (uint64_t)(ks[k] & 0x7FFFFFu));
    const uint64_t hi = r.x > key ? r.x : key;
    r.y = r.y < hi ? r.y : hi;
    r.x = r.x < key ? r.x : key;
  }
  io[qi] = r;
}
}  // namespace

extern "C" int orbm_knn2_device64(orbm_handle* h, const uint8_t* d_q, int nq, const uint8_t* d_db, int64_t ndb,
                                  int64_t idx_base, uint64_t* d_top2, void* stream) {
  if (!h) return ORBM_ERR_ARG;
  if (!d_q || !d_db || !d_top2 || nq < 0 || ndb < 0 || idx_base < 0 || idx_base + ndb >= (1ll << 40)) {
    h->err = "null pointer, negative size or index beyond 2^40";
    return ORBM_ERR_ARG;
  }
  if (nq == 0) return ORBM_OK;
  ORBM_CK(cudaSetDevice(h->device));
  cudaStream_t s = stream ? (cudaStream_t)stream : h->stream;
  if (ndb == 0) { ORBM_CK(cudaMemsetAsync(d_top2, 0xFF, (size_t)nq * 16, s)); return ORBM_OK; }
  ORBM_CK(h->stage_c.reserve((size_t)nq * 8 + 16));
  const int64_t kSeg = 1ll << 23;
  for (int64_t o = 0; o < ndb; o += kSeg) {
    const int64_t len = std::min(kSeg, ndb - o);
    const int rc = knn2_launch(h, d_q, nq, d_db + o * 32, len, 0, h->stage_c.as<uint32_t>(), s);
    if (rc != ORBM_OK) return rc;
    widen_merge_kernel<<<(nq + 255) / 256, 256, 0, s>>>(h->stage_c.as<uint2>(), nq, (uint64_t)(idx_base + o), o == 0,
                                                       reinterpret_cast<ulonglong2*>(d_top2));
    h->launches++;
  }
  ORBM_CK(cudaGetLastError());
  return ORBM_OK;
}

extern "C" int orbm_knn2(orbm_handle* h, const uint8_t* q, int nq, const uint8_t* db, int64_t ndb, int64_t idx_base,
                         uint32_t* top2) {
  if (!h) return ORBM_ERR_ARG;
  if (!q || !db || !top2 || nq < 0 || ndb < 0) { h->err = "null pointer or negative size"; return ORBM_ERR_ARG; }
  if (nq == 0) return ORBM_OK;
  ORBM_CK(cudaSetDevice(h->device));
  ORBM_CK(h->stage_a.reserve((size_t)nq * 32 + 16));
  ORBM_CK(h->stage_b.reserve((size_t)ndb * 32 + 16));
  ORBM_CK(h->stage_c.reserve((size_t)nq * 8 + 16));
  ORBM_CK(cudaMemcpyAsync(h->stage_a.p, q, (size_t)nq * 32, cudaMemcpyHostToDevice, h->stream));
  if (ndb) ORBM_CK(cudaMemcpyAsync(h->stage_b.p, db, (size_t)ndb * 32, cudaMemcpyHostToDevice, h->stream));
  ORBM_CK(cudaEventRecord(h->ev0, h->stream));
  int rc = knn2_launch(h, h->stage_a.as<uint8_t>(), nq, h->stage_b.as<uint8_t>(), ndb, idx_base,
                       h->stage_c.as<uint32_t>(), h->stream);
  if (rc != ORBM_OK) return rc;
  ORBM_CK(cudaEventRecord(h->ev1, h->stream));
  ORBM_CK(cudaMemcpyAsync(top2, h->stage_c.p, (size_t)nq * 8, cudaMemcpyDeviceToHost, h->stream));
  ORBM_CK(cudaStreamSynchronize(h->stream));
  cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
  return knn_check_err(h);
}

// Waits for the handle's stream and reports what the asynchronous entry points (orbm_knn2_device,
// orbm_knn2_device_static, orbm_knn2_sharded, orbm_merge_top2_device) could not return: the sticky "tile never arrived"
// flag of knn2_kernel and an asynchronous NCCL error of the communicator last used by orbm_knn2_sharded.  Work launched
// on a CALLER'S stream must have been synchronised by the caller first.
extern "C" int orbm_synchronize(orbm_handle* h) {
  if (!h) return ORBM_ERR_ARG;
  ORBM_CK(cudaSetDevice(h->device));
  ORBM_CK(cudaStreamSynchronize(h->stream));
  int rc = knn_check_err(h);
  if (rc != ORBM_OK) return rc;
  if (h->last_comm) return nccl_check_async(h, h->last_comm);
  return ORBM_OK;
}

extern "C" int orbm_merge_top2_device(orbm_handle* h, const uint32_t* d_gathered, int nshards, int nq, int th_dist,
                                      float ratio, uint32_t* d_top2, int32_t* d_match, void* stream) {
  if (!h) return ORBM_ERR_ARG;
  if (!d_gathered || nshards < 1 || nq < 0 || (!d_top2 && !d_match)) { h->err = "bad merge arguments"; return ORBM_ERR_ARG; }
  if (nq == 0) return ORBM_OK;
  ORBM_CK(cudaSetDevice(h->device));
  cudaStream_t s = stream ? (cudaStream_t)stream : h->stream;
  merge_top2_kernel<<<(nq + 255) / 256, 256, 0, s>>>(reinterpret_cast<const uint2*>(d_gathered), nshards, nq, th_dist,
                                                    ratio, reinterpret_cast<uint2*>(d_top2), d_match);
  h->launches++;
  ORBM_CK(cudaGetLastError());
  return ORBM_OK;
}

// NCCL is resolved from the process at run time so that liborbx.so has no link-time NCCL dependency and shares the
// NCCL instance already loaded in the process (e.g. the one bundled with PyTorch), else the system libnccl.so.2.
namespace {
struct NcclUid { char b[128]; };  // ncclUniqueId, passed by value
struct NcclApi {
  int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
  int (*GetUniqueId)(void*) = nullptr;
  int (*CommInitRank)(void**, int, NcclUid, int) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  int (*CommGetAsyncError)(void*, int*) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  bool ok = false;
};
NcclApi* nccl_api() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return api.ok ? &api : nullptr;
  tried = true;
  void* lib = nullptr;
  auto sym = [&](const char* name) -> void* {
    void* p = dlsym(RTLD_DEFAULT, name);
    if (!p) {
      if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
      if (lib) p = dlsym(lib, name);
    }
    return p;
  };
  *(void**)&api.AllGather = sym("ncclAllGather");
  *(void**)&api.GetUniqueId = sym("ncclGetUniqueId");
  *(void**)&api.CommInitRank = sym("ncclCommInitRank");
  *(void**)&api.CommDestroy = sym("ncclCommDestroy");
  *(void**)&api.GetErrorString = sym("ncclGetErrorString");
  *(void**)&api.CommGetAsyncError = sym("ncclCommGetAsyncError");
  api.ok = api.AllGather && api.GetUniqueId && api.CommInitRank && api.CommDestroy;
  return api.ok ? &api : nullptr;
}
}  // namespace

namespace {
// ncclCommGetAsyncError: errors of collectives already enqueued (a peer died, a transport failed) that no enqueue call
// returned.  ncclSuccess (0) and ncclInProgress (7) are not errors.
int nccl_check_async(orbm_handle* h, void* comm) {
  NcclApi* n = nccl_api();
  if (!n || !n->CommGetAsyncError || !comm) return ORBM_OK;
  int async = 0;
  const int rc = n->CommGetAsyncError(comm, &async);
  if (rc != 0 || (async != 0 && async != 7)) {
    const int code = rc != 0 ? rc : async;
    h->err = std::string("NCCL asynchronous error: ") + (n->GetErrorString ? n->GetErrorString(code) : std::to_string(code).c_str());
    return ORBM_ERR_COMM;
  }
  return ORBM_OK;
}
}  // namespace

extern "C" int orbm_comm_unique_id(uint8_t id[ORBM_COMM_ID_BYTES]) {
  NcclApi* n = nccl_api();
  if (!n || !id) return ORBM_ERR_COMM;
  return n->GetUniqueId(id) == 0 ? ORBM_OK : ORBM_ERR_COMM;
}

extern "C" int orbm_comm_init_rank(orbm_handle* h, int nranks, const uint8_t id[ORBM_COMM_ID_BYTES], int rank,
                                   void** comm) {
  if (!h) return ORBM_ERR_ARG;
  if (!id || !comm || nranks < 1 || rank < 0 || rank >= nranks) { h->err = "bad communicator arguments"; return ORBM_ERR_ARG; }
  NcclApi* n = nccl_api();
  if (!n) { h->err = "NCCL not found in the process and libnccl.so.2 not loadable"; return ORBM_ERR_COMM; }
  ORBM_CK(cudaSetDevice(h->device));
  NcclUid uid;
  memcpy(uid.b, id, sizeof(uid.b));
  const int rc = n->CommInitRank(comm, nranks, uid, rank);
  if (rc != 0) {
    h->err = std::string("ncclCommInitRank: ") + (n->GetErrorString ? n->GetErrorString(rc) : "error");
    return ORBM_ERR_COMM;
  }
  return ORBM_OK;
}

extern "C" void orbm_comm_release(orbm_handle* h, void* comm) {   // destroy + forget (orbm_synchronize queries last_comm)
  if (h && h->last_comm == comm) h->last_comm = nullptr;
  orbm_comm_destroy(comm);
}

extern "C" void orbm_comm_destroy(void* comm) {
  NcclApi* n = nccl_api();
  if (n && comm) n->CommDestroy(comm);
}

extern "C" int orbm_knn2_sharded(orbm_handle* h, const uint8_t* d_q, int nq, const uint8_t* d_db_shard,
                                 int64_t ndb_shard, int64_t idx_base, int nranks, void* comm, uint32_t* d_gathered,
                                 int th_dist, float ratio, uint32_t* d_top2, int32_t* d_match, void* stream) {
  if (!h) return ORBM_ERR_ARG;
  if (!comm || !d_gathered || nranks < 1) { h->err = "bad communicator / gather buffer"; return ORBM_ERR_ARG; }
  NcclApi* nccl = nccl_api();
  if (!nccl) { h->err = "NCCL not found in the process and libnccl.so.2 not loadable"; return ORBM_ERR_COMM; }
  ORBM_CK(cudaSetDevice(h->device));
  cudaStream_t s = stream ? (cudaStream_t)stream : h->stream;
  ORBM_CK(h->stage_c.reserve((size_t)nq * 8 + 16));
  uint32_t* d_local = h->stage_c.as<uint32_t>();
  int rc = knn2_launch(h, d_q, nq, d_db_shard, ndb_shard, idx_base, d_local, s);
  if (rc != ORBM_OK) return rc;
  const int ncclUint32 = 3;
  const int nrc = nccl->AllGather(d_local, d_gathered, (size_t)nq * 2, ncclUint32, comm, s);
  if (nrc != 0) { h->err = "ncclAllGather failed with code " + std::to_string(nrc); return ORBM_ERR_COMM; }
  h->last_comm = comm;
  rc = nccl_check_async(h, comm);                 // errors of EARLIER collectives on this communicator
  if (rc != ORBM_OK) return rc;
  return orbm_merge_top2_device(h, d_gathered, nranks, nq, th_dist, ratio, d_top2, d_match, s);
}

// ------------------------------------------------------------------------------------------------------------------
// orbm_knn2_device_static: the 2-NN of orbm_knn2_device for callers that may not allocate (CUDA graph capture, tight
// serving loops).  No counterpart in the reference.  The per-chunk
// partial keys and the error flag live in a file-scope device array instead of the handle's grown-on-demand buffers,
// so the call issues no cudaMalloc / cudaFree / cudaMallocAsync.  Every GPU that loads the library holds the array
// until exit; every call on a device shares it (calls that overlap on the GPU race: serialise them on one stream).
// ------------------------------------------------------------------------------------------------------------------
namespace {

constexpr size_t kKnnStaticFlagBytes = 256;  // error flag (first int), keeps the key area 256-byte aligned
static __device__ __align__(256) unsigned char g_knn_static_scratch[ORBM_KNN2_STATIC_SCRATCH_BYTES];

constexpr int kKnnStaticMaxDev = 64;
std::mutex g_knn_static_mu;
unsigned char* g_knn_static_ptr[kKnnStaticMaxDev] = {};  // DEVICE address of the array on each device (set up once)

// The array's name is only a host-side stand-in here: the device address comes from cudaGetSymbolAddress, looked up
// once per device with that device current.
int knn_static_setup(orbm_handle* h, unsigned char** out) {
  if (h->device < 0 || h->device >= kKnnStaticMaxDev) { h->err = "device ordinal beyond the static scratch table"; return ORBM_ERR_UNSUPPORTED; }
  std::lock_guard<std::mutex> lk(g_knn_static_mu);
  if (!g_knn_static_ptr[h->device]) {
    void* p = nullptr;
    ORBM_CK(cudaGetSymbolAddress(&p, g_knn_static_scratch));
    if (!p) { h->err = "cudaGetSymbolAddress returned a null address"; return ORBM_ERR_CUDA; }
    g_knn_static_ptr[h->device] = static_cast<unsigned char*>(p);
  }
  *out = g_knn_static_ptr[h->device];
  return ORBM_OK;
}

}  // namespace

extern "C" int orbm_knn2_device_static(orbm_handle* h, const uint8_t* d_q, int nq, const uint8_t* d_db, int64_t ndb,
                                       int64_t idx_base, uint32_t* d_top2, void* stream) {
  if (!h) return ORBM_ERR_ARG;
  if (!d_q || !d_db || !d_top2 || nq < 0 || ndb < 0 || idx_base < 0) {
    h->err = "null pointer or negative size";
    return ORBM_ERR_ARG;
  }
  if (idx_base + ndb > (1 << 23)) {
    h->err = "idx_base + ndb exceeds the 2^23 index range of the packed keys";
    return ORBM_ERR_UNSUPPORTED;
  }
  if (((uintptr_t)d_q | (uintptr_t)d_db) & 15) { h->err = "descriptor arrays must be 16-byte aligned"; return ORBM_ERR_ARG; }
  ORBM_CK(cudaSetDevice(h->device));
  cudaStream_t s = stream ? (cudaStream_t)stream : h->stream;
  unsigned char* scratch = nullptr;
  const int rc = knn_static_setup(h, &scratch);
  if (rc != ORBM_OK) return rc;
  if (nq == 0) return ORBM_OK;
  if (ndb == 0) { ORBM_CK(cudaMemsetAsync(d_top2, 0xFF, (size_t)nq * 8, s)); return ORBM_OK; }
  // the chunk plan of knn2_launch (same grid => same partial keys => same result)
  int qpt = nq > 2 * kKnnThreads ? 4 : (nq > kKnnThreads ? 2 : 1);
  int qtiles = 0, nchunks = 0;
  int64_t chunk_len = 0;
  auto plan = [&](int min_tiles) {
    qtiles = (nq + kKnnThreads * qpt - 1) / (kKnnThreads * qpt);
    const int64_t want = std::max<int64_t>(1, (2 * (int64_t)h->num_sms + qtiles - 1) / qtiles);
    chunk_len = (ndb + want - 1) / want;
    chunk_len = std::max<int64_t>((chunk_len + kKnnTile - 1) / kKnnTile * kKnnTile, (int64_t)min_tiles * kKnnTile);
    nchunks = (int)((ndb + chunk_len - 1) / chunk_len);
  };
  plan(8);
  if ((int64_t)nchunks * qtiles < h->num_sms) { qpt = 1; plan(1); }
  const size_t need = kKnnStaticFlagBytes + (size_t)nchunks * (size_t)nq * 8;
  if (need > (size_t)ORBM_KNN2_STATIC_SCRATCH_BYTES) {
    h->err = "orbm_knn2_device_static: " + std::to_string(nchunks) + " chunks x " + std::to_string(nq) +
             " queries need " + std::to_string(need) + " bytes of scratch, the static array holds " +
             std::to_string((size_t)ORBM_KNN2_STATIC_SCRATCH_BYTES) + " (use orbm_knn2_device, or fewer queries per call)";
    return ORBM_ERR_UNSUPPORTED;
  }
  int* d_err = reinterpret_cast<int*>(scratch);
  uint32_t* part = reinterpret_cast<uint32_t*>(scratch + kKnnStaticFlagBytes);
  // the array keeps the previous call's contents: the flag must start at zero; every (chunk, query) key pair of the
  // part area read by the merge is written by this call's knn2 CTAs
  ORBM_CK(cudaMemsetAsync(d_err, 0, sizeof(int), s));
  dim3 grid(nchunks, qtiles);
  const uint4* q4 = reinterpret_cast<const uint4*>(d_q);
  const uint4* db4 = reinterpret_cast<const uint4*>(d_db);
  const int cl = (int)chunk_len;
  const uint32_t ib = (uint32_t)idx_base;
  if (qpt == 4) knn2_kernel<4, 2><<<grid, kKnnThreads, 0, s>>>(q4, nq, db4, ndb, cl, ib, part, d_err);
  else if (qpt == 2) knn2_kernel<2, 2><<<grid, kKnnThreads, 0, s>>>(q4, nq, db4, ndb, cl, ib, part, d_err);
  else knn2_kernel<1, 2><<<grid, kKnnThreads, 0, s>>>(q4, nq, db4, ndb, cl, ib, part, d_err);
  merge_top2_kernel<<<(nq + 255) / 256, 256, 0, s>>>(reinterpret_cast<const uint2*>(part), nchunks, nq, 0, 0.f,
                                                    reinterpret_cast<uint2*>(d_top2), nullptr);
  h->launches += 2;
  ORBM_CK(cudaGetLastError());
  return ORBM_OK;
}
