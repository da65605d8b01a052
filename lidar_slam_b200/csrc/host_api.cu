// Host-buffer entry of the ICP path: what a numpy caller of the reference binds directly.
//
//   b200icp_run_host   run_icp for a batch, HOST pointers in and out   modules/icp.py:123-189
//                      (the loops of modules/localization.py:112-130, main.py:161-172,
//                       code/icp_warm_up/test_icp.py:70-91 handed over as one pair list)
//
// The reference works on numpy arrays in host memory, so this entry owns the device side: a
// grow-only arena per device, a copy stream and two compute streams.  The scan pool is streamed
// to the device in slices while the pair list runs in chunks: before chunk c starts, the prefix
// of the pool its pairs reference is resident (consecutive pairs (i, i-1) need the pool in
// order, so copy and compute overlap; arbitrary pair lists simply wait for more of the pool).
// Planar batches take the GATED path instead of chunks: the pool is enqueued in slices on the copy
// stream, each slice followed by a 4-byte copy that raises a device-side counter of resident scans;
// ONE persistent kernel launch over the whole pair list is issued right behind them on a compute
// stream (it only depends on the small inputs) and waits (ld.acquire.sys) on that counter before it
// stages a pair.  No per-chunk
// launch tails, copy and compute overlap down to the slice.
// Results come back with one set of device-to-host copies at the end.  The call is synchronous.
// Pinned host memory makes the copies asynchronous and faster; pageable memory works too.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include <mutex>

#include "../../include/b200icp.h"

namespace {

constexpr int kMaxChunks = 16;
constexpr int kPipelineMinPairs = 2048;
constexpr int kMaxSlices = 64;      // gated launch: pool slices announced to the running kernel
constexpr int kMax2dPoints = 2048;  // the planar kernel's limit (api.cu)

struct HostArena {
  char* base = nullptr;
  size_t cap = 0;
  cudaStream_t copy = nullptr, comp[2] = {nullptr, nullptr};
  cudaEvent_t ready[kMaxChunks] = {}, done[2] = {nullptr, nullptr}, small = nullptr;
  unsigned int* progress_host = nullptr;   // pinned: cumulative scan counts, one per slice
  bool ok = false;
};

HostArena g_arena[64];
std::mutex g_mutex;

bool arena_init(HostArena& a) {
  if (a.ok) return true;
  if (cudaStreamCreateWithFlags(&a.copy, cudaStreamNonBlocking) != cudaSuccess) return false;
  for (int k = 0; k < 2; ++k) {
    if (cudaStreamCreateWithFlags(&a.comp[k], cudaStreamNonBlocking) != cudaSuccess) return false;
    if (cudaEventCreateWithFlags(&a.done[k], cudaEventDisableTiming) != cudaSuccess) return false;
  }
  for (int k = 0; k < kMaxChunks; ++k)
    if (cudaEventCreateWithFlags(&a.ready[k], cudaEventDisableTiming) != cudaSuccess) return false;
  if (cudaEventCreateWithFlags(&a.small, cudaEventDisableTiming) != cudaSuccess) return false;
  if (cudaHostAlloc(reinterpret_cast<void**>(&a.progress_host), kMaxSlices * sizeof(unsigned int),
                    cudaHostAllocDefault) != cudaSuccess) return false;
  a.ok = true;
  return true;
}

bool arena_reserve(HostArena& a, size_t bytes) {
  if (bytes <= a.cap) return true;
  if (a.base) { cudaFree(a.base); a.base = nullptr; a.cap = 0; }
  const size_t want = bytes + bytes / 4;       // head room: the next, slightly larger batch reuses it
  if (cudaMalloc(reinterpret_cast<void**>(&a.base), want) != cudaSuccess) { cudaGetLastError(); return false; }
  a.cap = want;
  return true;
}

struct Carver {
  size_t o = 0;
  size_t take(size_t n) { const size_t at = o; o += (n + 255) / 256 * 256; return at; }
};

// TSE3_from_TSE2 form (modules/utils.py:227-240): the z = 0 plane maps to itself and the
// homogeneous row is (0, 0, 0, 1) -- the same test the planar kernel applies per pair.
bool keeps_plane(const double* T) {
  return T[2] == 0.0 && T[6] == 0.0 && T[8] == 0.0 && T[9] == 0.0 && T[11] == 0.0 && T[12] == 0.0 &&
         T[13] == 0.0 && T[14] == 0.0 && T[15] == 1.0;
}

#define B200_TRY(expr) do { if ((expr) != cudaSuccess) { cudaGetLastError(); return B200ICP_ERR_CUDA; } } while (0)

}  // namespace

extern "C" {

int b200icp_run_host(const double* scans, const int32_t* counts, int n_scans, int n_max, int dim,
                     const int32_t* pairs, const double* init, const double* epsilon,
                     const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                     double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                     int32_t* corr_out) {
  if (!scans || !counts || !pairs || !init || !epsilon || !normalize || !T_out || n_scans <= 0 ||
      n_max <= 0 || n_pairs < 0 || max_iters < 0 || (dim != 2 && dim != 3) || ((size_t)n_max * dim) % 2 != 0)
    return B200ICP_ERR_BAD_ARG;
  if (n_pairs == 0) return B200ICP_OK;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { cudaGetLastError(); return B200ICP_ERR_NO_DEVICE; }
  std::lock_guard<std::mutex> lock(g_mutex);
  HostArena& a = g_arena[dev];
  if (!arena_init(a)) { cudaGetLastError(); return B200ICP_ERR_CUDA; }

  // An initial transform that leaves the plane: the reference lifts the scans to z = 0 and carries
  // on in 3-D (icp.py:153-156), so the whole batch goes through the general kernel on a z-padded pool.
  int run_dim = dim;
  if (dim == 2)
    for (int b = 0; b < n_pairs; ++b)
      if (!keeps_plane(init + (size_t)b * 16)) { run_dim = 3; break; }

  bool gated = run_dim == 2 && n_max <= kMax2dPoints && n_pairs >= kPipelineMinPairs;
  if (const char* e = getenv("B200ICP_HOST_GATED")) gated = gated && atoi(e) != 0;   // 0: chunked path (A/B)
  if (gated) {
    Carver cv;
    const size_t row_bytes = (size_t)n_max * 2 * sizeof(double);
    const size_t o_scans = cv.take((size_t)n_scans * row_bytes);
    const size_t o_counts = cv.take((size_t)n_scans * 4);
    const size_t o_pairs = cv.take((size_t)n_pairs * 8);
    const size_t o_init = cv.take((size_t)n_pairs * 128);
    const size_t o_eps = cv.take((size_t)n_pairs * 8);
    const size_t o_norm = cv.take((size_t)n_pairs);
    const size_t o_T = cv.take((size_t)n_pairs * 128);
    const size_t o_err = cv.take((size_t)n_pairs * 8);
    const size_t o_iters = cv.take((size_t)n_pairs * 4);
    const size_t o_status = cv.take((size_t)n_pairs * 4);
    const size_t o_corr = corr_out ? cv.take((size_t)n_pairs * n_max * 4) : 0;
    const size_t o_flag = cv.take(256);
    size_t ws_bytes = 0;
    const int rcw = b200icp_workspace_bytes(n_pairs, n_max, 2, &ws_bytes);
    if (rcw != B200ICP_OK) return rcw;
    const size_t o_ws = cv.take(ws_bytes);
    if (!arena_reserve(a, cv.o)) return B200ICP_ERR_CUDA;
    char* d = a.base;
    B200_TRY(cudaMemsetAsync(d + o_flag, 0, 256, a.copy));
    B200_TRY(cudaMemcpyAsync(d + o_counts, counts, (size_t)n_scans * 4, cudaMemcpyHostToDevice, a.copy));
    B200_TRY(cudaMemcpyAsync(d + o_pairs, pairs, (size_t)n_pairs * 8, cudaMemcpyHostToDevice, a.copy));
    B200_TRY(cudaMemcpyAsync(d + o_init, init, (size_t)n_pairs * 128, cudaMemcpyHostToDevice, a.copy));
    B200_TRY(cudaMemcpyAsync(d + o_eps, epsilon, (size_t)n_pairs * 8, cudaMemcpyHostToDevice, a.copy));
    B200_TRY(cudaMemcpyAsync(d + o_norm, normalize, (size_t)n_pairs, cudaMemcpyHostToDevice, a.copy));
    B200_TRY(cudaEventRecord(a.small, a.copy));
    B200_TRY(cudaStreamWaitEvent(a.comp[0], a.small, 0));
    // The slices are ENQUEUED before the launch: if something serialises launches (a profiler,
    // CUDA_LAUNCH_BLOCKING) the kernel must never wait for a copy the host has not issued yet.
    int n_slices = 24;
    if (const char* e = getenv("B200ICP_HOST_SLICES")) { const int v = atoi(e); if (v >= 1) n_slices = v; }
    if (n_slices > kMaxSlices - 1) n_slices = kMaxSlices - 1;   // the last pinned word receives the gave-up flag
    if (n_slices > n_scans) n_slices = n_scans;
    int copied = 0;
    bool copy_ok = true;
    for (int c = 0; c < n_slices && copy_ok; ++c) {
      const int upto = (int)((long long)n_scans * (c + 1) / n_slices);
      if (upto > copied)
        copy_ok = cudaMemcpyAsync(d + o_scans + (size_t)copied * row_bytes, scans + (size_t)copied * n_max * 2,
                                  (size_t)(upto - copied) * row_bytes, cudaMemcpyHostToDevice, a.copy) == cudaSuccess;
      copied = upto;
      a.progress_host[c] = (unsigned int)copied;
      copy_ok = copy_ok && cudaMemcpyAsync(d + o_flag, a.progress_host + c, sizeof(unsigned int),
                                           cudaMemcpyHostToDevice, a.copy) == cudaSuccess;
    }
    if (!copy_ok) {
      cudaGetLastError();
      cudaStreamSynchronize(a.copy);
      cudaGetLastError();
      return B200ICP_ERR_CUDA;
    }
    const int rc_run = b200icp_run_gated(
        reinterpret_cast<const double*>(d + o_scans), reinterpret_cast<const int32_t*>(d + o_counts), n_scans, n_max,
        2, reinterpret_cast<const int32_t*>(d + o_pairs), reinterpret_cast<const double*>(d + o_init),
        reinterpret_cast<const double*>(d + o_eps), reinterpret_cast<const uint8_t*>(d + o_norm), n_pairs, max_iters,
        stopping_thresh, reinterpret_cast<double*>(d + o_T), reinterpret_cast<double*>(d + o_err),
        reinterpret_cast<int32_t*>(d + o_iters), reinterpret_cast<int32_t*>(d + o_status),
        corr_out ? reinterpret_cast<int32_t*>(d + o_corr) : nullptr, d + o_ws, ws_bytes, a.comp[0],
        reinterpret_cast<const unsigned int*>(d + o_flag), nullptr, 0, 0);
    B200_TRY(cudaEventRecord(a.done[0], a.comp[0]));
    B200_TRY(cudaStreamWaitEvent(a.copy, a.done[0], 0));
    if (rc_run == B200ICP_OK) {
      B200_TRY(cudaMemcpyAsync(T_out, d + o_T, (size_t)n_pairs * 128, cudaMemcpyDeviceToHost, a.copy));
      if (err_out) B200_TRY(cudaMemcpyAsync(err_out, d + o_err, (size_t)n_pairs * 8, cudaMemcpyDeviceToHost, a.copy));
      if (iters_out) B200_TRY(cudaMemcpyAsync(iters_out, d + o_iters, (size_t)n_pairs * 4, cudaMemcpyDeviceToHost, a.copy));
      if (status_out) B200_TRY(cudaMemcpyAsync(status_out, d + o_status, (size_t)n_pairs * 4, cudaMemcpyDeviceToHost, a.copy));
      if (corr_out) B200_TRY(cudaMemcpyAsync(corr_out, d + o_corr, (size_t)n_pairs * n_max * 4, cudaMemcpyDeviceToHost, a.copy));
      // workspace word 5: the kernel gave up waiting for the pool (never expected; results void)
      B200_TRY(cudaMemcpyAsync(a.progress_host + kMaxSlices - 1, d + o_ws + 20, sizeof(unsigned int),
                               cudaMemcpyDeviceToHost, a.copy));
    }
    B200_TRY(cudaStreamSynchronize(a.copy));
    if (rc_run == B200ICP_OK && a.progress_host[kMaxSlices - 1] != 0u) return B200ICP_ERR_CUDA;
    return rc_run;
  }

  // chunks of the pair list and the pool prefix each one needs
  int n_chunks = 1;
  if (n_pairs >= kPipelineMinPairs && run_dim == dim) {
    n_chunks = 6;
    if (const char* e = getenv("B200ICP_PIPELINE_CHUNKS")) { const int v = atoi(e); if (v >= 1) n_chunks = v; }
    if (n_chunks > kMaxChunks) n_chunks = kMaxChunks;
  }
  int lo[kMaxChunks + 1], need[kMaxChunks];
  for (int c = 0; c <= n_chunks; ++c) lo[c] = (int)((long long)n_pairs * c / n_chunks);
  for (int c = 0; c < n_chunks; ++c) {
    int m = -1;
    for (int b = lo[c]; b < lo[c + 1]; ++b) {
      const int s = pairs[2 * b], t = pairs[2 * b + 1];
      if (s > m) m = s;
      if (t > m) m = t;
    }
    need[c] = (c == n_chunks - 1) ? n_scans : (m + 1 < 0 ? 0 : (m + 1 > n_scans ? n_scans : m + 1));
  }

  // arena layout
  size_t ws_bytes[kMaxChunks];
  Carver cv;
  const size_t row_bytes = (size_t)n_max * run_dim * sizeof(double);
  const size_t o_scans = cv.take((size_t)n_scans * row_bytes);
  const size_t o_counts = cv.take((size_t)n_scans * 4);
  const size_t o_pairs = cv.take((size_t)n_pairs * 8);
  const size_t o_init = cv.take((size_t)n_pairs * 128);
  const size_t o_eps = cv.take((size_t)n_pairs * 8);
  const size_t o_norm = cv.take((size_t)n_pairs);
  const size_t o_T = cv.take((size_t)n_pairs * 128);
  const size_t o_err = cv.take((size_t)n_pairs * 8);
  const size_t o_iters = cv.take((size_t)n_pairs * 4);
  const size_t o_status = cv.take((size_t)n_pairs * 4);
  const size_t o_corr = corr_out ? cv.take((size_t)n_pairs * n_max * 4) : 0;
  size_t o_ws[kMaxChunks];
  for (int c = 0; c < n_chunks; ++c) {
    const int rc = b200icp_workspace_bytes(lo[c + 1] - lo[c], n_max, run_dim, &ws_bytes[c]);
    if (rc != B200ICP_OK) return rc;
    o_ws[c] = cv.take(ws_bytes[c]);
  }
  if (!arena_reserve(a, cv.o)) return B200ICP_ERR_CUDA;
  char* d = a.base;

  // small inputs first
  B200_TRY(cudaMemcpyAsync(d + o_counts, counts, (size_t)n_scans * 4, cudaMemcpyHostToDevice, a.copy));
  B200_TRY(cudaMemcpyAsync(d + o_pairs, pairs, (size_t)n_pairs * 8, cudaMemcpyHostToDevice, a.copy));
  B200_TRY(cudaMemcpyAsync(d + o_init, init, (size_t)n_pairs * 128, cudaMemcpyHostToDevice, a.copy));
  B200_TRY(cudaMemcpyAsync(d + o_eps, epsilon, (size_t)n_pairs * 8, cudaMemcpyHostToDevice, a.copy));
  B200_TRY(cudaMemcpyAsync(d + o_norm, normalize, (size_t)n_pairs, cudaMemcpyHostToDevice, a.copy));
  if (run_dim != dim) B200_TRY(cudaMemsetAsync(d + o_scans, 0, (size_t)n_scans * row_bytes, a.copy));
  B200_TRY(cudaEventRecord(a.small, a.copy));
  B200_TRY(cudaStreamWaitEvent(a.comp[0], a.small, 0));
  B200_TRY(cudaStreamWaitEvent(a.comp[1], a.small, 0));

  int copied = 0, rc_run = B200ICP_OK;
  for (int c = 0; c < n_chunks; ++c) {
    if (need[c] > copied) {
      if (run_dim == dim) {
        B200_TRY(cudaMemcpyAsync(d + o_scans + (size_t)copied * row_bytes, scans + (size_t)copied * n_max * dim,
                                 (size_t)(need[c] - copied) * row_bytes, cudaMemcpyHostToDevice, a.copy));
      } else {   // (x, y) rows into (x, y, 0) rows
        B200_TRY(cudaMemcpy2DAsync(d + o_scans + (size_t)copied * row_bytes, 24, scans + (size_t)copied * n_max * 2, 16,
                                   16, (size_t)(need[c] - copied) * n_max, cudaMemcpyHostToDevice, a.copy));
      }
      copied = need[c];
    }
    B200_TRY(cudaEventRecord(a.ready[c], a.copy));
    const int nb = lo[c + 1] - lo[c];
    if (nb <= 0) continue;
    cudaStream_t st = a.comp[c & 1];
    B200_TRY(cudaStreamWaitEvent(st, a.ready[c], 0));
    const size_t b0 = (size_t)lo[c];
    rc_run = b200icp_run(reinterpret_cast<const double*>(d + o_scans), reinterpret_cast<const int32_t*>(d + o_counts),
                         n_scans, n_max, run_dim, reinterpret_cast<const int32_t*>(d + o_pairs) + 2 * b0,
                         reinterpret_cast<const double*>(d + o_init) + 16 * b0,
                         reinterpret_cast<const double*>(d + o_eps) + b0,
                         reinterpret_cast<const uint8_t*>(d + o_norm) + b0, nb, max_iters, stopping_thresh,
                         reinterpret_cast<double*>(d + o_T) + 16 * b0, reinterpret_cast<double*>(d + o_err) + b0,
                         reinterpret_cast<int32_t*>(d + o_iters) + b0, reinterpret_cast<int32_t*>(d + o_status) + b0,
                         corr_out ? reinterpret_cast<int32_t*>(d + o_corr) + b0 * n_max : nullptr,
                         d + o_ws[c], ws_bytes[c], st);
    if (rc_run != B200ICP_OK) break;
  }
  // results back once both compute streams are done
  for (int k = 0; k < 2; ++k) {
    B200_TRY(cudaEventRecord(a.done[k], a.comp[k]));
    B200_TRY(cudaStreamWaitEvent(a.copy, a.done[k], 0));
  }
  if (rc_run == B200ICP_OK) {
    B200_TRY(cudaMemcpyAsync(T_out, d + o_T, (size_t)n_pairs * 128, cudaMemcpyDeviceToHost, a.copy));
    if (err_out) B200_TRY(cudaMemcpyAsync(err_out, d + o_err, (size_t)n_pairs * 8, cudaMemcpyDeviceToHost, a.copy));
    if (iters_out) B200_TRY(cudaMemcpyAsync(iters_out, d + o_iters, (size_t)n_pairs * 4, cudaMemcpyDeviceToHost, a.copy));
    if (status_out) B200_TRY(cudaMemcpyAsync(status_out, d + o_status, (size_t)n_pairs * 4, cudaMemcpyDeviceToHost, a.copy));
    if (corr_out) B200_TRY(cudaMemcpyAsync(corr_out, d + o_corr, (size_t)n_pairs * n_max * 4, cudaMemcpyDeviceToHost, a.copy));
  }
  B200_TRY(cudaStreamSynchronize(a.copy));
  return rc_run;
}

int b200icp_host_release(void) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { cudaGetLastError(); return B200ICP_ERR_NO_DEVICE; }
  std::lock_guard<std::mutex> lock(g_mutex);
  HostArena& a = g_arena[dev];
  if (a.base) { cudaFree(a.base); a.base = nullptr; a.cap = 0; }
  return B200ICP_OK;
}

}  // extern "C"
