// Host-buffer entries of the ICP path: what a numpy caller of the reference binds directly.
//
//   b200icp_run_host[_ex]     run_icp for a batch, HOST pointers in and out   modules/icp.py:123-189
//                             (the loops of modules/localization.py:112-130, main.py:161-172,
//                              code/icp_warm_up/test_icp.py:70-91 handed over as one pair list)
//   b200icp_run_host_ranges   the same with the lidar RANGES as input: get_lidar_data
//                             (modules/localization.py:252-307) + run_icp in one launch, half the
//                             bytes over PCIe (a quarter with float32 ranges)
//
// The reference works on numpy arrays in host memory, so these entries own the device side: a
// grow-only arena per device, a copy stream and two compute streams.  Planar batches take the
// GATED path: the pool (points or range rows) is enqueued in slices on the copy stream, each slice
// followed by a 4-byte copy that raises a device-side counter of resident rows; ONE persistent
// kernel launch over the whole pair list is issued right behind them on a compute stream (it only
// depends on the small inputs) and waits (ld.acquire.sys) on that counter before it stages a pair:
// no per-chunk launch tails, copy and compute overlap down to the slice.  Other batches (3-D clouds,
// initial transforms that leave the plane, small batches) run in chunks of the pair list: before
// chunk c starts, the prefix of the pool its pairs reference is resident.
// Results come back with one set of device-to-host copies at the end.  The calls are synchronous.
// Pinned host memory makes the copies asynchronous and faster; pageable memory works too.
// With peer buffers (multi-GPU) the kernel's epilogue also stores every finished pose row into
// every rank's gather buffer (b200icp_run_exchange), so the same entry serves 1 and N GPUs.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "../../include/b200icp.h"

namespace b200icp { unsigned env_flags(); }

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
  unsigned int* progress_host = nullptr;   // pinned: cumulative row counts, one per slice (+ the gave-up word)
  char* pin = nullptr;                     // pinned staging of the single-pair entry (grow-only)
  size_t pin_cap = 0;
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

// The pool a host call works on: points [n_scans, n_max, dim] float64, or ranges [n_scans, n_beams].
struct HostPool {
  const void* data;
  int kind;                 // 0: points, 1: ranges float64, 2: ranges float32
  int n_scans, n_max, dim;  // ranges: n_max = n_beams rounded up to even, dim = 2
  int n_beams;
  const int32_t* counts;    // points only
  const double *cos_tab, *sin_tab;
  double range_min, range_max, off_x, off_y;
  int32_t* counts_out;      // ranges only, host, may be null
  size_t row_bytes() const {
    return kind == 0 ? (size_t)n_max * dim * sizeof(double) : (size_t)n_beams * (kind == 1 ? 8 : 4);
  }
};

struct HostIo {
  const int32_t* pairs; const double* init; const double* epsilon; const uint8_t* normalize;
  int n_pairs, max_iters; double stopping_thresh;
  double* T_out; double* err_out; int32_t* iters_out; int32_t* status_out; int32_t* corr_out;
  const b200icp_exchange* exchange;
  unsigned flags;
};

struct SmallOffsets { size_t pairs, init, eps, norm, T, err, iters, status, corr; };

SmallOffsets carve_small(Carver& cv, const HostIo& io, int n_max) {
  SmallOffsets o{};
  o.pairs = cv.take((size_t)io.n_pairs * 8);
  o.init = cv.take((size_t)io.n_pairs * 128);
  o.eps = cv.take((size_t)io.n_pairs * 8);
  o.norm = cv.take((size_t)io.n_pairs);
  o.T = cv.take((size_t)io.n_pairs * 128);
  o.err = cv.take((size_t)io.n_pairs * 8);
  o.iters = cv.take((size_t)io.n_pairs * 4);
  o.status = cv.take((size_t)io.n_pairs * 4);
  o.corr = io.corr_out ? cv.take((size_t)io.n_pairs * n_max * 4) : 0;
  return o;
}

// Page-locked host memory is mapped into the device's address space (unified addressing): a small result
// array there can be the kernel's output pointer itself, so that the results leave the device as pairs
// retire instead of in a train of small copies behind the launch (each costs 6-8 us of latency; T, error,
// passes, status and counts: ~45 us of a 2 ms call).  Returns nullptr for pageable memory.
void* device_view_of_pinned(const void* host_ptr) {
  if (!host_ptr) return nullptr;
  cudaPointerAttributes at{};
  if (cudaPointerGetAttributes(&at, host_ptr) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  if (at.type != cudaMemoryTypeHost || !at.devicePointer) return nullptr;
  return at.devicePointer;
}

struct DirectOut { double* T = nullptr; double* err = nullptr; int32_t* iters = nullptr; int32_t* status = nullptr;
                   int32_t* counts = nullptr; };

int upload_small(HostArena& a, char* d, const SmallOffsets& o, const HostIo& io) {
  B200_TRY(cudaMemcpyAsync(d + o.pairs, io.pairs, (size_t)io.n_pairs * 8, cudaMemcpyHostToDevice, a.copy));
  B200_TRY(cudaMemcpyAsync(d + o.init, io.init, (size_t)io.n_pairs * 128, cudaMemcpyHostToDevice, a.copy));
  B200_TRY(cudaMemcpyAsync(d + o.eps, io.epsilon, (size_t)io.n_pairs * 8, cudaMemcpyHostToDevice, a.copy));
  B200_TRY(cudaMemcpyAsync(d + o.norm, io.normalize, (size_t)io.n_pairs, cudaMemcpyHostToDevice, a.copy));
  return B200ICP_OK;
}

int download_results(HostArena& a, char* d, const SmallOffsets& o, const HostIo& io, int n_max,
                     const DirectOut& direct = DirectOut()) {
  const size_t n = (size_t)io.n_pairs;
  if (!direct.T) B200_TRY(cudaMemcpyAsync(io.T_out, d + o.T, n * 128, cudaMemcpyDeviceToHost, a.copy));
  if (io.err_out && !direct.err) B200_TRY(cudaMemcpyAsync(io.err_out, d + o.err, n * 8, cudaMemcpyDeviceToHost, a.copy));
  if (io.iters_out && !direct.iters) B200_TRY(cudaMemcpyAsync(io.iters_out, d + o.iters, n * 4, cudaMemcpyDeviceToHost, a.copy));
  if (io.status_out && !direct.status) B200_TRY(cudaMemcpyAsync(io.status_out, d + o.status, n * 4, cudaMemcpyDeviceToHost, a.copy));
  if (io.corr_out) B200_TRY(cudaMemcpyAsync(io.corr_out, d + o.corr, n * n_max * 4, cudaMemcpyDeviceToHost, a.copy));
  return B200ICP_OK;
}

// ONE gated launch over the whole pair list while the pool arrives in slices (planar kernel).
// gated_enqueue issues everything -- uploads, the launch, the copies back -- and returns without waiting;
// gated_finish waits for the copy stream and reports.  run_gated = both (the synchronous entries);
// b200icp_submit_host_ranges / b200icp_wait_host call them apart, on arenas of their own.
int gated_enqueue(HostArena& a, const HostPool& pool, const HostIo& io, int* rc_run_out) {
  *rc_run_out = B200ICP_OK;
  Carver cv;
  const size_t row_bytes = pool.row_bytes();
  const size_t o_pool = cv.take((size_t)pool.n_scans * row_bytes);
  const size_t o_counts = cv.take((size_t)pool.n_scans * 4);
  const size_t o_tabs = pool.kind ? cv.take((size_t)pool.n_beams * 16) : 0;
  const SmallOffsets o = carve_small(cv, io, pool.n_max);
  const size_t o_flag = cv.take(256);
  size_t ws_bytes = 0;
  const int rcw = b200icp_workspace_bytes(io.n_pairs, pool.n_max, 2, &ws_bytes);
  if (rcw != B200ICP_OK) return rcw;
  const size_t o_ws = cv.take(ws_bytes);
  if (!arena_reserve(a, cv.o)) return B200ICP_ERR_CUDA;
  char* d = a.base;
  a.progress_host[kMaxSlices - 1] = 0u;
  DirectOut direct;
  if (!(io.flags & B200ICP_FLAG_HOST_NO_DIRECT)) {
    direct.T = static_cast<double*>(device_view_of_pinned(io.T_out));
    direct.err = static_cast<double*>(device_view_of_pinned(io.err_out));
    direct.iters = static_cast<int32_t*>(device_view_of_pinned(io.iters_out));
    direct.status = static_cast<int32_t*>(device_view_of_pinned(io.status_out));
    if (pool.kind) direct.counts = static_cast<int32_t*>(device_view_of_pinned(pool.counts_out));
  }
  B200_TRY(cudaMemsetAsync(d + o_flag, 0, 256, a.copy));
  if (pool.kind == 0) {
    B200_TRY(cudaMemcpyAsync(d + o_counts, pool.counts, (size_t)pool.n_scans * 4, cudaMemcpyHostToDevice, a.copy));
  } else {
    B200_TRY(cudaMemcpyAsync(d + o_tabs, pool.cos_tab, (size_t)pool.n_beams * 8, cudaMemcpyHostToDevice, a.copy));
    B200_TRY(cudaMemcpyAsync(d + o_tabs + (size_t)pool.n_beams * 8, pool.sin_tab, (size_t)pool.n_beams * 8,
                             cudaMemcpyHostToDevice, a.copy));
    if (direct.counts) memset(pool.counts_out, 0xff, (size_t)pool.n_scans * 4);                          // -1: not referenced
    else if (pool.counts_out) B200_TRY(cudaMemsetAsync(d + o_counts, 0xff, (size_t)pool.n_scans * 4, a.copy));
  }
  const int rc_up = upload_small(a, d, o, io);
  if (rc_up != B200ICP_OK) return rc_up;
  B200_TRY(cudaEventRecord(a.small, a.copy));
  B200_TRY(cudaStreamWaitEvent(a.comp[0], a.small, 0));
  // The slices are ENQUEUED before the launch: if something serialises launches (a profiler,
  // CUDA_LAUNCH_BLOCKING) the kernel must never wait for a copy the host has not issued yet.
  int n_slices = (int)((io.flags >> 8) & 0xffu);
  // Default: 16 equal slices for range rows, 24 for points.  Measured (profiles/r2_host_entry_slices.txt):
  // 8..32 equal slices are within 1.5 % of each other; 12 slices growing by 1.45x (small first slice, fewer
  // enqueue calls ahead of the launch) were 3 % SLOWER than 16 equal ones although the upload alone runs at
  // 55 GB/s, twice the rate at which the kernel consumes it -- kept equal.
  if (n_slices == 0) n_slices = pool.kind != 0 ? 16 : 24;
  if (n_slices > kMaxSlices - 1) n_slices = kMaxSlices - 1;   // the last pinned word receives the gave-up flag
  if (n_slices > pool.n_scans) n_slices = pool.n_scans;
  int copied = 0;
  bool copy_ok = true;
  const char* src = static_cast<const char*>(pool.data);
  for (int c = 0; c < n_slices && copy_ok; ++c) {
    const int upto = (int)((long long)pool.n_scans * (c + 1) / n_slices);
    if (upto > copied)
      copy_ok = cudaMemcpyAsync(d + o_pool + (size_t)copied * row_bytes, src + (size_t)copied * row_bytes,
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
  int rc_run;
  if (pool.kind == 0) {
    rc_run = b200icp_run_ex(
        reinterpret_cast<const double*>(d + o_pool), reinterpret_cast<const int32_t*>(d + o_counts), pool.n_scans,
        pool.n_max, 2, reinterpret_cast<const int32_t*>(d + o.pairs), reinterpret_cast<const double*>(d + o.init),
        reinterpret_cast<const double*>(d + o.eps), reinterpret_cast<const uint8_t*>(d + o.norm), io.n_pairs,
        io.max_iters, io.stopping_thresh, direct.T ? direct.T : reinterpret_cast<double*>(d + o.T),
        direct.err ? direct.err : reinterpret_cast<double*>(d + o.err),
        direct.iters ? direct.iters : reinterpret_cast<int32_t*>(d + o.iters),
        direct.status ? direct.status : reinterpret_cast<int32_t*>(d + o.status),
        io.corr_out ? reinterpret_cast<int32_t*>(d + o.corr) : nullptr, d + o_ws, ws_bytes, a.comp[0],
        reinterpret_cast<const unsigned int*>(d + o_flag), io.exchange, io.flags);
  } else {
    rc_run = b200icp_run_ranges(
        d + o_pool, pool.kind == 2 ? 1 : 0, pool.n_scans, pool.n_beams, reinterpret_cast<const double*>(d + o_tabs),
        reinterpret_cast<const double*>(d + o_tabs) + pool.n_beams, pool.range_min, pool.range_max, pool.off_x,
        pool.off_y, reinterpret_cast<const int32_t*>(d + o.pairs), reinterpret_cast<const double*>(d + o.init),
        reinterpret_cast<const double*>(d + o.eps), reinterpret_cast<const uint8_t*>(d + o.norm), io.n_pairs,
        io.max_iters, io.stopping_thresh, direct.T ? direct.T : reinterpret_cast<double*>(d + o.T),
        direct.err ? direct.err : reinterpret_cast<double*>(d + o.err),
        direct.iters ? direct.iters : reinterpret_cast<int32_t*>(d + o.iters),
        direct.status ? direct.status : reinterpret_cast<int32_t*>(d + o.status),
        io.corr_out ? reinterpret_cast<int32_t*>(d + o.corr) : nullptr,
        direct.counts ? direct.counts : (pool.counts_out ? reinterpret_cast<int32_t*>(d + o_counts) : nullptr), d + o_ws, ws_bytes, a.comp[0],
        reinterpret_cast<const unsigned int*>(d + o_flag), io.exchange, io.flags);
  }
  B200_TRY(cudaEventRecord(a.done[0], a.comp[0]));
  B200_TRY(cudaStreamWaitEvent(a.copy, a.done[0], 0));
  if (rc_run == B200ICP_OK) {
    const int rc_dn = download_results(a, d, o, io, pool.n_max, direct);
    if (rc_dn != B200ICP_OK) return rc_dn;
    if (pool.kind && pool.counts_out && !direct.counts)
      B200_TRY(cudaMemcpyAsync(pool.counts_out, d + o_counts, (size_t)pool.n_scans * 4, cudaMemcpyDeviceToHost, a.copy));
    // workspace word 5: the kernel gave up waiting for the pool (never expected; those pairs were not run)
    B200_TRY(cudaMemcpyAsync(a.progress_host + kMaxSlices - 1, d + o_ws + 20, sizeof(unsigned int),
                             cudaMemcpyDeviceToHost, a.copy));
  }
  *rc_run_out = rc_run;
  return B200ICP_OK;
}

int gated_finish(HostArena& a, int rc_run) {
  B200_TRY(cudaStreamSynchronize(a.copy));
  if (rc_run == B200ICP_OK && a.progress_host[kMaxSlices - 1] != 0u) return B200ICP_ERR_CUDA;
  return rc_run;
}

int run_gated(HostArena& a, const HostPool& pool, const HostIo& io) {
  int rc_run = B200ICP_OK;
  const int rc = gated_enqueue(a, pool, io, &rc_run);
  if (rc != B200ICP_OK) { cudaStreamSynchronize(a.copy); cudaGetLastError(); return rc; }
  return gated_finish(a, rc_run);
}

// Chunks of the pair list on two compute streams; each chunk waits for the pool prefix it references.
int run_chunked(HostArena& a, const HostPool& pool, const HostIo& io, int run_dim) {
  int n_chunks = 1;
  if (io.n_pairs >= kPipelineMinPairs && run_dim == pool.dim) {
    n_chunks = (int)((io.flags >> 16) & 0xffu);
    if (n_chunks == 0) n_chunks = 6;
    if (n_chunks > kMaxChunks) n_chunks = kMaxChunks;
  }
  int lo[kMaxChunks + 1], need[kMaxChunks];
  for (int c = 0; c <= n_chunks; ++c) lo[c] = (int)((long long)io.n_pairs * c / n_chunks);
  for (int c = 0; c < n_chunks; ++c) {
    int m = -1;
    for (int b = lo[c]; b < lo[c + 1]; ++b) {
      const int s = io.pairs[2 * b], t = io.pairs[2 * b + 1];
      if (s > m) m = s;
      if (t > m) m = t;
    }
    need[c] = (c == n_chunks - 1) ? pool.n_scans : (m + 1 < 0 ? 0 : (m + 1 > pool.n_scans ? pool.n_scans : m + 1));
  }
  size_t ws_bytes[kMaxChunks], o_ws[kMaxChunks];
  Carver cv;
  const size_t row_bytes = (size_t)pool.n_max * run_dim * sizeof(double);
  const size_t o_scans = cv.take((size_t)pool.n_scans * row_bytes);
  const size_t o_counts = cv.take((size_t)pool.n_scans * 4);
  const SmallOffsets o = carve_small(cv, io, pool.n_max);
  for (int c = 0; c < n_chunks; ++c) {
    const int rc = b200icp_workspace_bytes(lo[c + 1] - lo[c], pool.n_max, run_dim, &ws_bytes[c]);
    if (rc != B200ICP_OK) return rc;
    o_ws[c] = cv.take(ws_bytes[c]);
  }
  if (!arena_reserve(a, cv.o)) return B200ICP_ERR_CUDA;
  char* d = a.base;
  const double* scans = static_cast<const double*>(pool.data);

  B200_TRY(cudaMemcpyAsync(d + o_counts, pool.counts, (size_t)pool.n_scans * 4, cudaMemcpyHostToDevice, a.copy));
  const int rc_up = upload_small(a, d, o, io);
  if (rc_up != B200ICP_OK) return rc_up;
  if (run_dim != pool.dim) B200_TRY(cudaMemsetAsync(d + o_scans, 0, (size_t)pool.n_scans * row_bytes, a.copy));
  B200_TRY(cudaEventRecord(a.small, a.copy));
  B200_TRY(cudaStreamWaitEvent(a.comp[0], a.small, 0));
  B200_TRY(cudaStreamWaitEvent(a.comp[1], a.small, 0));

  int copied = 0, rc_run = B200ICP_OK;
  for (int c = 0; c < n_chunks; ++c) {
    if (need[c] > copied) {
      if (run_dim == pool.dim) {
        B200_TRY(cudaMemcpyAsync(d + o_scans + (size_t)copied * row_bytes, scans + (size_t)copied * pool.n_max * pool.dim,
                                 (size_t)(need[c] - copied) * row_bytes, cudaMemcpyHostToDevice, a.copy));
      } else {   // (x, y) rows into (x, y, 0) rows
        B200_TRY(cudaMemcpy2DAsync(d + o_scans + (size_t)copied * row_bytes, 24, scans + (size_t)copied * pool.n_max * 2, 16,
                                   16, (size_t)(need[c] - copied) * pool.n_max, cudaMemcpyHostToDevice, a.copy));
      }
      copied = need[c];
    }
    B200_TRY(cudaEventRecord(a.ready[c], a.copy));
    const int nb = lo[c + 1] - lo[c];
    if (nb <= 0) continue;
    cudaStream_t st = a.comp[c & 1];
    B200_TRY(cudaStreamWaitEvent(st, a.ready[c], 0));
    const size_t b0 = (size_t)lo[c];
    b200icp_exchange ex_c{};
    if (io.exchange) {      // the chunk's rows; the arrival flags are raised once, after every chunk (below)
      ex_c = *io.exchange;
      ex_c.peer_row0 += (long long)b0;
      ex_c.flag_offset_bytes = 0;
    }
    rc_run = b200icp_run_ex(reinterpret_cast<const double*>(d + o_scans), reinterpret_cast<const int32_t*>(d + o_counts),
                            pool.n_scans, pool.n_max, run_dim, reinterpret_cast<const int32_t*>(d + o.pairs) + 2 * b0,
                            reinterpret_cast<const double*>(d + o.init) + 16 * b0,
                            reinterpret_cast<const double*>(d + o.eps) + b0,
                            reinterpret_cast<const uint8_t*>(d + o.norm) + b0, nb, io.max_iters, io.stopping_thresh,
                            reinterpret_cast<double*>(d + o.T) + 16 * b0, reinterpret_cast<double*>(d + o.err) + b0,
                            reinterpret_cast<int32_t*>(d + o.iters) + b0, reinterpret_cast<int32_t*>(d + o.status) + b0,
                            io.corr_out ? reinterpret_cast<int32_t*>(d + o.corr) + b0 * pool.n_max : nullptr,
                            d + o_ws[c], ws_bytes[c], st, nullptr, io.exchange ? &ex_c : nullptr, io.flags);
    if (rc_run != B200ICP_OK) break;
  }
  for (int k = 0; k < 2; ++k) {   // results back once both compute streams are done
    B200_TRY(cudaEventRecord(a.done[k], a.comp[k]));
    B200_TRY(cudaStreamWaitEvent(a.copy, a.done[k], 0));
  }
  if (rc_run == B200ICP_OK) {
    if (io.exchange && io.exchange->n_peers > 0 && io.exchange->flag_offset_bytes != 0) {
      const int rc_sig = b200icp_exchange_signal(io.exchange, a.copy);    // all chunks have retired (events above)
      if (rc_sig != B200ICP_OK) return rc_sig;
    }
    const int rc_dn = download_results(a, d, o, io, pool.n_max);
    if (rc_dn != B200ICP_OK) return rc_dn;
  }
  B200_TRY(cudaStreamSynchronize(a.copy));
  return rc_run;
}

int run_host_impl(const HostPool& pool, HostIo io) {
  if (!pool.data || !io.pairs || !io.init || !io.epsilon || !io.normalize || !io.T_out || pool.n_scans <= 0 ||
      pool.n_max <= 0 || io.n_pairs < 0 || io.max_iters < 0 || (pool.dim != 2 && pool.dim != 3) ||
      ((size_t)pool.n_max * pool.dim) % 2 != 0)
    return B200ICP_ERR_BAD_ARG;
  if (io.n_pairs == 0) return B200ICP_OK;
  io.flags |= b200icp::env_flags();
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { cudaGetLastError(); return B200ICP_ERR_NO_DEVICE; }
  std::lock_guard<std::mutex> lock(g_mutex);
  HostArena& a = g_arena[dev];
  if (!arena_init(a)) { cudaGetLastError(); return B200ICP_ERR_CUDA; }

  if (pool.kind != 0) return run_gated(a, pool, io);   // ranges: always the planar kernel, one launch

  // An initial transform that leaves the plane: the reference lifts the scans to z = 0 and carries
  // on in 3-D (icp.py:153-156), so the whole batch goes through the general kernel on a z-padded pool.
  int run_dim = pool.dim;
  if (pool.dim == 2)
    for (int b = 0; b < io.n_pairs; ++b)
      if (!keeps_plane(io.init + (size_t)b * 16)) { run_dim = 3; break; }
  const bool gated = run_dim == 2 && pool.n_max <= kMax2dPoints && io.n_pairs >= kPipelineMinPairs &&
                     !(io.flags & B200ICP_FLAG_HOST_CHUNKED);
  return gated ? run_gated(a, pool, io) : run_chunked(a, pool, io, run_dim);
}

}  // namespace

extern "C" {

int b200icp_run_host_ex(const double* scans, const int32_t* counts, int n_scans, int n_max, int dim,
                        const int32_t* pairs, const double* init, const double* epsilon,
                        const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                        double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                        int32_t* corr_out, const b200icp_exchange* exchange, unsigned flags) {
  if (!counts) return B200ICP_ERR_BAD_ARG;
  HostPool pool{};
  pool.data = scans; pool.kind = 0; pool.n_scans = n_scans; pool.n_max = n_max; pool.dim = dim; pool.counts = counts;
  HostIo io{pairs, init, epsilon, normalize, n_pairs, max_iters, stopping_thresh, T_out, err_out, iters_out,
            status_out, corr_out, exchange, flags};
  return run_host_impl(pool, io);
}

int b200icp_run_host(const double* scans, const int32_t* counts, int n_scans, int n_max, int dim,
                     const int32_t* pairs, const double* init, const double* epsilon,
                     const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                     double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                     int32_t* corr_out) {
  return b200icp_run_host_ex(scans, counts, n_scans, n_max, dim, pairs, init, epsilon, normalize, n_pairs, max_iters,
                             stopping_thresh, T_out, err_out, iters_out, status_out, corr_out, nullptr, 0u);
}

int b200icp_run_host_ranges(const void* ranges, int ranges_dtype, int n_scans, int n_beams, const double* cos_tab,
                            const double* sin_tab, double range_min, double range_max, double off_x, double off_y,
                            const int32_t* pairs, const double* init, const double* epsilon,
                            const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                            double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                            int32_t* corr_out, int32_t* counts_out, const b200icp_exchange* exchange,
                            unsigned flags) {
  if (!cos_tab || !sin_tab || n_beams <= 0 || n_beams > kMax2dPoints || (ranges_dtype != 0 && ranges_dtype != 1))
    return B200ICP_ERR_BAD_ARG;
  HostPool pool{};
  pool.data = ranges; pool.kind = ranges_dtype ? 2 : 1; pool.n_scans = n_scans; pool.n_beams = n_beams;
  pool.n_max = n_beams + (n_beams & 1); pool.dim = 2;
  pool.cos_tab = cos_tab; pool.sin_tab = sin_tab; pool.range_min = range_min; pool.range_max = range_max;
  pool.off_x = off_x; pool.off_y = off_y; pool.counts_out = counts_out;
  HostIo io{pairs, init, epsilon, normalize, n_pairs, max_iters, stopping_thresh, T_out, err_out, iters_out,
            status_out, corr_out, exchange, flags};
  return run_host_impl(pool, io);
}

// ---- two calls in flight: submit / wait ---------------------------------------------------------------
// Lanes 0 and 1 of a device have arenas, streams and pinned words of their own (beside the arena of the
// synchronous entries), so the uploads and the head of call k + 1 run under the queue tail and the copies
// back of call k.
namespace {
struct Lane { HostArena arena; bool busy = false; int rc_run = B200ICP_OK; unsigned gen = 0; };
Lane g_lane[64][2];
}  // namespace

int b200icp_submit_host_ranges(const void* ranges, int ranges_dtype, int n_scans, int n_beams, const double* cos_tab,
                               const double* sin_tab, double range_min, double range_max, double off_x, double off_y,
                               const int32_t* pairs, const double* init, const double* epsilon,
                               const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                               double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                               int32_t* corr_out, int32_t* counts_out, const b200icp_exchange* exchange,
                               unsigned flags, int* ticket_out) {
  if (!ticket_out) return B200ICP_ERR_BAD_ARG;
  *ticket_out = -1;
  if (!ranges || !cos_tab || !sin_tab || n_beams <= 0 || n_beams > kMax2dPoints || (ranges_dtype != 0 && ranges_dtype != 1) ||
      !pairs || !init || !epsilon || !normalize || !T_out || n_scans <= 0 || n_pairs <= 0 || max_iters < 0)
    return B200ICP_ERR_BAD_ARG;
  HostPool pool{};
  pool.data = ranges; pool.kind = ranges_dtype ? 2 : 1; pool.n_scans = n_scans; pool.n_beams = n_beams;
  pool.n_max = n_beams + (n_beams & 1); pool.dim = 2;
  pool.cos_tab = cos_tab; pool.sin_tab = sin_tab; pool.range_min = range_min; pool.range_max = range_max;
  pool.off_x = off_x; pool.off_y = off_y; pool.counts_out = counts_out;
  HostIo io{pairs, init, epsilon, normalize, n_pairs, max_iters, stopping_thresh, T_out, err_out, iters_out,
            status_out, corr_out, exchange, flags | b200icp::env_flags()};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { cudaGetLastError(); return B200ICP_ERR_NO_DEVICE; }
  std::lock_guard<std::mutex> lock(g_mutex);
  int which = -1;
  for (int k = 0; k < 2; ++k)
    if (!g_lane[dev][k].busy) { which = k; break; }
  if (which < 0) return B200ICP_ERR_BUSY;
  Lane& ln = g_lane[dev][which];
  if (!arena_init(ln.arena)) { cudaGetLastError(); return B200ICP_ERR_CUDA; }
  int rc_run = B200ICP_OK;
  const int rc = gated_enqueue(ln.arena, pool, io, &rc_run);
  if (rc != B200ICP_OK) { cudaStreamSynchronize(ln.arena.copy); cudaGetLastError(); return rc; }
  ln.busy = true; ln.rc_run = rc_run; ln.gen = (ln.gen + 1u) & 0x7fffffu;
  *ticket_out = (int)((ln.gen << 8) | ((unsigned)dev << 1) | (unsigned)which);
  return B200ICP_OK;
}

int b200icp_wait_host(int ticket) {
  if (ticket < 0) return B200ICP_ERR_BAD_ARG;
  const int which = ticket & 1, dev = (ticket >> 1) & 63;
  const unsigned gen = (unsigned)ticket >> 8;
  Lane& ln = g_lane[dev][which];
  int rc_run;
  {
    std::lock_guard<std::mutex> lock(g_mutex);
    if (!ln.busy || ln.gen != gen) return B200ICP_ERR_BAD_ARG;      // unknown, or already waited for
    rc_run = ln.rc_run;
  }
  const int rc = gated_finish(ln.arena, rc_run);     // outside the lock: the other lane may be submitted meanwhile
  std::lock_guard<std::mutex> lock(g_mutex);
  ln.busy = false;
  return rc;
}

// run_icp for ONE pair, host arrays in and out: the un-batched drop-in (modules/icp.py:123-189 as
// modules/localization.py:120 calls it).  Everything the launch needs travels in ONE pinned block
// and one host-to-device copy, everything it returns in one device-to-host copy.
int b200icp_run_pair_host(const double* src, int n_src, const double* tgt, int n_tgt, int dim,
                          const double* init, double epsilon, int max_iters, double stopping_thresh,
                          int normalize, double* T_out, double* err_out, int32_t* iters_out,
                          int32_t* status_out, int32_t* corr_out) {
  if (!src || !tgt || !init || !T_out || n_src <= 0 || n_tgt <= 0 || max_iters < 0 || (dim != 2 && dim != 3))
    return B200ICP_ERR_BAD_ARG;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { cudaGetLastError(); return B200ICP_ERR_NO_DEVICE; }
  std::lock_guard<std::mutex> lock(g_mutex);
  HostArena& a = g_arena[dev];
  if (!arena_init(a)) { cudaGetLastError(); return B200ICP_ERR_CUDA; }
  int run_dim = dim;
  if (dim == 2 && !keeps_plane(init)) run_dim = 3;      // the reference's lifting to z = 0 (icp.py:153-156)
  int n_max = n_src > n_tgt ? n_src : n_tgt;
  n_max += n_max & 1;
  const size_t row = (size_t)n_max * run_dim * 8;
  // input block: [init 128][eps 8][counts 8][pair 8][norm 8 (padded)] [source row][target row]
  const size_t in_hdr = 160, in_bytes = in_hdr + 2 * row;
  // output block: [T 128][err 8][iters 4][status 4][corr n_max*4]
  const size_t out_bytes = 144 + (corr_out ? (size_t)n_max * 4 : 0);
  size_t ws_bytes = 0;
  const int rcw = b200icp_workspace_bytes(1, n_max, run_dim, &ws_bytes);
  if (rcw != B200ICP_OK) return rcw;
  Carver cv;
  const size_t o_in = cv.take(in_bytes), o_out = cv.take(out_bytes), o_ws = cv.take(ws_bytes);
  if (!arena_reserve(a, cv.o)) return B200ICP_ERR_CUDA;
  const size_t pin_need = in_bytes + out_bytes + 256;
  if (pin_need > a.pin_cap) {
    if (a.pin) cudaFreeHost(a.pin);
    a.pin = nullptr; a.pin_cap = 0;
    if (cudaHostAlloc(reinterpret_cast<void**>(&a.pin), pin_need * 2, cudaHostAllocDefault) != cudaSuccess) {
      cudaGetLastError();
      return B200ICP_ERR_CUDA;
    }
    a.pin_cap = pin_need * 2;
  }
  char* h = a.pin;
  memcpy(h, init, 128);
  memcpy(h + 128, &epsilon, 8);
  const int32_t counts[2] = {n_src, n_tgt}, pair[2] = {0, 1};
  memcpy(h + 136, counts, 8);
  memcpy(h + 144, pair, 8);
  memset(h + 152, 0, 8);
  h[152] = normalize ? 1 : 0;
  if (run_dim == dim) {
    memcpy(h + in_hdr, src, (size_t)n_src * dim * 8);
    memcpy(h + in_hdr + row, tgt, (size_t)n_tgt * dim * 8);
  } else {      // (x, y) -> (x, y, 0)
    double* s3 = reinterpret_cast<double*>(h + in_hdr);
    double* t3 = reinterpret_cast<double*>(h + in_hdr + row);
    for (int i = 0; i < n_src; ++i) { s3[3 * i] = src[2 * i]; s3[3 * i + 1] = src[2 * i + 1]; s3[3 * i + 2] = 0.0; }
    for (int i = 0; i < n_tgt; ++i) { t3[3 * i] = tgt[2 * i]; t3[3 * i + 1] = tgt[2 * i + 1]; t3[3 * i + 2] = 0.0; }
  }
  char* d = a.base;
  cudaStream_t st = a.comp[0];
  B200_TRY(cudaMemcpyAsync(d + o_in, h, in_bytes, cudaMemcpyHostToDevice, st));
  const int rc_run = b200icp_run_ex(
      reinterpret_cast<const double*>(d + o_in + in_hdr), reinterpret_cast<const int32_t*>(d + o_in + 136), 2, n_max, run_dim,
      reinterpret_cast<const int32_t*>(d + o_in + 144), reinterpret_cast<const double*>(d + o_in),
      reinterpret_cast<const double*>(d + o_in + 128), reinterpret_cast<const uint8_t*>(d + o_in + 152), 1, max_iters,
      stopping_thresh, reinterpret_cast<double*>(d + o_out), reinterpret_cast<double*>(d + o_out + 128),
      reinterpret_cast<int32_t*>(d + o_out + 136), reinterpret_cast<int32_t*>(d + o_out + 140),
      corr_out ? reinterpret_cast<int32_t*>(d + o_out + 144) : nullptr, d + o_ws, ws_bytes, st, nullptr, nullptr,
      B200ICP_FLAG_LATENCY);
  if (rc_run != B200ICP_OK) { cudaStreamSynchronize(st); cudaGetLastError(); return rc_run; }
  char* ho = h + ((in_bytes + 255) / 256) * 256;
  B200_TRY(cudaMemcpyAsync(ho, d + o_out, out_bytes, cudaMemcpyDeviceToHost, st));
  B200_TRY(cudaStreamSynchronize(st));
  memcpy(T_out, ho, 128);
  if (err_out) memcpy(err_out, ho + 128, 8);
  if (iters_out) memcpy(iters_out, ho + 136, 4);
  if (status_out) memcpy(status_out, ho + 140, 4);
  if (corr_out) memcpy(corr_out, ho + 144, (size_t)n_src * 4);
  return B200ICP_OK;
}

void* b200icp_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (bytes == 0) bytes = 1;
  if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return p;
}

int b200icp_host_free(void* p) {
  if (!p) return B200ICP_OK;
  if (cudaFreeHost(p) != cudaSuccess) { cudaGetLastError(); return B200ICP_ERR_CUDA; }
  return B200ICP_OK;
}

int b200icp_host_release(void) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { cudaGetLastError(); return B200ICP_ERR_NO_DEVICE; }
  std::lock_guard<std::mutex> lock(g_mutex);
  HostArena& a = g_arena[dev];
  if (a.base) { cudaFree(a.base); a.base = nullptr; a.cap = 0; }
  for (int k = 0; k < 2; ++k) {                       // the lanes of submit / wait, unless a call is in flight on them
    Lane& ln = g_lane[dev][k];
    if (!ln.busy && ln.arena.base) { cudaFree(ln.arena.base); ln.arena.base = nullptr; ln.arena.cap = 0; }
  }
  return B200ICP_OK;
}

}  // extern "C"
