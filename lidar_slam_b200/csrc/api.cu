// extern "C" entry points of libb200icp.so (see include/b200icp.h).
#include <cuda_runtime.h>
#include <cstdlib>
#include <stdlib.h>

#include "icp2d.cuh"
#include "icp2d_rows.cuh"
#include "icp3d_flow.cuh"

using namespace b200icp;

// Developer overrides from the environment, read ONCE (first use) and OR-ed into every call's
// flags; the launch path itself takes its switches as arguments (include/b200icp.h).
unsigned b200icp::env_flags() {
  static const unsigned f = [] {
    unsigned v = 0;
    if (const char* e = getenv("B200ICP_SWEEP")) if (e[0] == 'e' || e[0] == 'E') v |= B200ICP_FLAG_EXHAUSTIVE;
    if (const char* e = getenv("B200ICP_HOST_GATED")) if (atoi(e) == 0) v |= B200ICP_FLAG_HOST_CHUNKED;
    if (const char* e = getenv("B200ICP_HOST_SLICES")) { const int n = atoi(e); if (n >= 1 && n <= 255) v |= B200ICP_FLAG_SLICES(n); }
    if (const char* e = getenv("B200ICP_PIPELINE_CHUNKS")) { const int n = atoi(e); if (n >= 1 && n <= 255) v |= B200ICP_FLAG_CHUNKS(n); }
    return v;
  }();
  return f;
}

namespace {

constexpr int kMax2d = 2048;  // 6-bit block id x 32 targets
inline bool uses_general_kernel(int n_max, int dim) { return !(dim == 2 && n_max <= kMax2d); }

struct DeviceInfo {
  int sms;
  int ok;
};
DeviceInfo device_info() {
  int dev = 0;
  DeviceInfo d{0, 0};
  if (cudaGetDevice(&dev) != cudaSuccess) return d;
  int major = 0;
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev);
  d.ok = (major == 10);
  return d;
}

// B200ICP_FLAG_EXHAUSTIVE selects the sweep that evaluates every (source, target) candidate
// (same results; kept for A/B timing and as the FP32-issue-rate figure of the inner loop).
// Default: the block-pruned sweep (icp2d.cuh).

template <int R, int W, int MINB, bool PRUNE>
int launch2d_variant(const RunParams& p, cudaStream_t stream, int sms) {
  auto kern = icp2d_kernel<R, W, MINB, PRUNE>;
  const Smem2dLayout L = smem2d_layout(p.n_max, W, PRUNE, R);
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total) != cudaSuccess)
    return B200ICP_ERR_CUDA;
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, W * 32, L.total) != cudaSuccess || per_sm < 1)
    return B200ICP_ERR_CUDA;
  const long long resident = (long long)per_sm * sms;
  const int grid = (int)(p.n_pairs < resident ? p.n_pairs : resident);
  kern<<<grid, W * 32, L.total, stream>>>(p);
  return cudaGetLastError() == cudaSuccess ? B200ICP_OK : B200ICP_ERR_CUDA;
}

// The shipped planar kernel (icp2d_rows.cuh): W warps per pair, R queries per lane.
template <int W, int R, int MINB, bool PRUNE, bool RANGES, bool SRCG = false>
int launch_rows_variant(const RunParams& p, cudaStream_t stream, int sms) {
  auto kern = icp2d_rows_kernel<W, R, MINB, PRUNE, RANGES, SRCG>;
  const Rows2dLayout L = rows2d_layout(p.n_max, W, RANGES, SRCG);
  // attribute + occupancy of this instantiation, remembered per device for the last n_max (a caller
  // that loops over one batch shape pays the two runtime queries once; a racing first call repeats them)
  struct Cached { int n_max = 0, per_sm = 0; };
  static Cached cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return B200ICP_ERR_NO_DEVICE;
  int per_sm = cached[dev].n_max == p.n_max ? cached[dev].per_sm : 0;
  if (per_sm < 1) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total) != cudaSuccess)
      return B200ICP_ERR_CUDA;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, W * 32, L.total) != cudaSuccess || per_sm < 1)
      return B200ICP_ERR_CUDA;
    cached[dev].per_sm = 0;
    cached[dev].n_max = p.n_max;
    cached[dev].per_sm = per_sm;
  }
  const long long resident = (long long)per_sm * sms;
  const int grid = (int)(p.n_pairs < resident ? p.n_pairs : resident);
  kern<<<grid, W * 32, L.total, stream>>>(p);
  return cudaGetLastError() == cudaSuccess ? B200ICP_OK : B200ICP_ERR_CUDA;
}
template <int W, int R, int MINB>
int launch_rows(const RunParams& p, cudaStream_t stream, int sms, unsigned flags) {
  const bool prune = !(flags & B200ICP_FLAG_NO_PRUNE);
  if (p.ranges)
    return prune ? launch_rows_variant<W, R, MINB, true, true>(p, stream, sms)
                 : launch_rows_variant<W, R, MINB, false, true>(p, stream, sms);
  return prune ? launch_rows_variant<W, R, MINB, true, false>(p, stream, sms)
               : launch_rows_variant<W, R, MINB, false, false>(p, stream, sms);
}

// General kernel (icp3d_flow.cuh): three launches per chunk of <= kFlowMaxPairs pairs on the caller's
// stream -- setup, prepare (one CTA per referenced scan), the persistent dataflow kernel.
int launch3d(RunParams p, int dim, void* workspace, size_t workspace_bytes, cudaStream_t stream, int sms, unsigned flags) {
  const int total = p.n_pairs;
  const bool prune = !(flags & (B200ICP_FLAG_NO_PRUNE | B200ICP_FLAG_EXHAUSTIVE));
  const long long row0 = p.peer_row0;
  for (int b0 = 0; b0 < total; b0 += kFlowMaxPairs) {
    const int nb = total - b0 < kFlowMaxPairs ? total - b0 : kFlowMaxPairs;
    RunParams q = p;
    q.n_pairs = nb;
    q.pairs = p.pairs + 2 * (size_t)b0;
    if (p.init) q.init = p.init + 16 * (size_t)b0;
    if (p.epsilon) q.epsilon = p.epsilon + b0;
    if (p.normalize) q.normalize = p.normalize + b0;
    if (p.T_out) q.T_out = p.T_out + 16 * (size_t)b0;
    if (p.err_out) q.err_out = p.err_out + b0;
    if (p.iters_out) q.iters_out = p.iters_out + b0;
    if (p.status_out) q.status_out = p.status_out + b0;
    if (p.corr_out) q.corr_out = p.corr_out + (size_t)b0 * p.n_max;
    if (p.dist2_out) q.dist2_out = p.dist2_out + (size_t)b0 * p.n_max;
#ifdef B200ICP_FLOW_DEBUG
    if (const char* e = getenv("B200ICP_DEBUG_PTR")) q.dist2_out = reinterpret_cast<double*>(strtoull(e, nullptr, 0));
#endif
    q.peer_row0 = row0 + b0;
    if (b0 + nb < total) q.flag_off = 0;      // arrival flags: raised by the last launch only (stream order covers the rest)
    const FlowLayout L = flow_layout(nb, p.n_max);
    if (workspace_bytes < L.total) return B200ICP_ERR_WORKSPACE;
    char* ws = static_cast<char*>(workspace);
    if (b0 > 0 && cudaMemsetAsync(ws, 0, 256, stream) != cudaSuccess) return B200ICP_ERR_CUDA;
    flow_setup_kernel<<<1, 1024, 0, stream>>>(q, dim, L, ws);
    const size_t prep_smem = (size_t)16 * kPrepThreads * sizeof(unsigned);
    flow_prepare_kernel<<<2 * nb, kPrepThreads, prep_smem, stream>>>(q, dim, L, ws);
    int per_sm = 0;
    auto kern = prune ? icp3d_flow_kernel<true> : icp3d_flow_kernel<false>;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kFlowWarps * 32, 0) != cudaSuccess || per_sm < 1)
      return B200ICP_ERR_CUDA;
    long long items = 0;
    items = (long long)nb * L.max_chunks;
    long long grid = (long long)per_sm * sms;
    if (grid > items) grid = items;
    if (grid < 1) grid = 1;
    kern<<<(unsigned)grid, kFlowWarps * 32, 0, stream>>>(q, L, ws);
    if (cudaGetLastError() != cudaSuccess) return B200ICP_ERR_CUDA;
  }
  return B200ICP_OK;
}

int dispatch(RunParams& p, int dim, void* workspace, size_t workspace_bytes, cudaStream_t stream, unsigned flags) {
  flags |= b200icp::env_flags();
  if ((!p.ranges && (!p.scans || !p.counts)) || !p.pairs || p.n_scans <= 0 || p.n_max <= 0 || p.n_pairs < 0 ||
      (dim != 2 && dim != 3))
    return B200ICP_ERR_BAD_ARG;
  if (p.n_pairs == 0) return B200ICP_OK;
  if (!workspace || workspace_bytes < 256) return B200ICP_ERR_WORKSPACE;
  if (!p.ranges && (reinterpret_cast<uintptr_t>(p.scans) & 15u) != 0) return B200ICP_ERR_BAD_ARG;
  const DeviceInfo d = device_info();
  if (!d.ok) return B200ICP_ERR_NO_DEVICE;
  p.queue = reinterpret_cast<unsigned int*>(workspace);
  if (cudaMemsetAsync(workspace, 0, 256, stream) != cudaSuccess) return B200ICP_ERR_CUDA;
  if (dim == 2 && p.n_max <= kMax2d) {
    if (!(flags & B200ICP_FLAG_EXHAUSTIVE) || p.ranges) {
      // row kernel: rows of 32*R source points per warp step; capacity is not a function of W.
      // R = 1 (one query per lane): tighter row boxes, fewer candidate blocks swept than R = 2..4
      // (profiles/r2_row_kernel_cta_shapes.txt: 1.77 ms vs 1.98 ms per 4941-pair step)
      if (p.n_max <= 512) return launch_rows<4, 1, 5>(p, stream, d.sms, flags);
      if (p.n_max <= 1280) {
#ifdef B200ICP_TUNE_VARIANTS
        // developer build: CTA shapes for A/B timing, selected by the environment at every call
        // (tests/devtools/time_variants.py)
        if (const char* e = getenv("B200ICP_VARIANT")) switch (atoi(e)) {
          case 1: return launch_rows_variant<4, 2, 4, true, false>(p, stream, d.sms);
          case 2: return launch_rows_variant<8, 2, 2, true, false>(p, stream, d.sms);
          case 3: return launch_rows_variant<6, 1, 3, true, false>(p, stream, d.sms);
          case 4: return launch_rows_variant<8, 1, 2, true, false>(p, stream, d.sms);
          case 5: return launch_rows_variant<5, 1, 3, true, false>(p, stream, d.sms);
          case 6: return launch_rows_variant<3, 1, 5, true, false>(p, stream, d.sms);
          case 7: return launch_rows_variant<4, 3, 4, true, false>(p, stream, d.sms);
          case 8: return launch_rows_variant<2, 1, 8, true, false>(p, stream, d.sms);
          case 9: return launch_rows_variant<4, 1, 5, true, false, true>(p, stream, d.sms);
          case 10: return launch_rows_variant<4, 1, 4, true, false, true>(p, stream, d.sms);
          case 11: return launch_rows_variant<4, 1, 6, true, false, true>(p, stream, d.sms);
          default: break;
        }
#endif
        // few pairs (the un-batched drop-in): 16 warps on each pair, 1-2 rows per warp and pass
        if ((flags & B200ICP_FLAG_LATENCY) && p.n_pairs <= d.sms) return launch_rows<16, 1, 1>(p, stream, d.sms, flags);
        return launch_rows<4, 1, 4>(p, stream, d.sms, flags);
      }
      return launch_rows<8, 1, 2>(p, stream, d.sms, flags);
    }
    // exhaustive register-tiled kernel (round 1): W warps, up to R rows of 32 source points per
    // warp: capacity 32*R*W >= n_max; 80 registers, 3 CTAs/SM at 1081 points (FMA-bound)
    if (p.n_max <= 256) return launch2d_variant<2, 4, 8, false>(p, stream, d.sms);
    if (p.n_max <= 512) return launch2d_variant<4, 4, 6, false>(p, stream, d.sms);
    if (p.n_max <= 1280) return launch2d_variant<5, 8, 3, false>(p, stream, d.sms);
    return launch2d_variant<8, 8, 2, false>(p, stream, d.sms);
  }
  return launch3d(p, dim, workspace, workspace_bytes, stream, d.sms, flags);
}

__global__ void ffma_probe_kernel(float* sink, int iters) {
  const float a = 1.0f + 1e-7f * threadIdx.x, b = 1e-7f * (blockIdx.x + 1);
  float x[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) x[k] = (float)k;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int rep = 0; rep < 4; ++rep)
#pragma unroll
      for (int k = 0; k < 16; ++k) x[k] = __fmaf_rn(x[k], a, b);
  }
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < 16; ++k) s += x[k];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void pack_poses_kernel(const double* T, const double* err, int n, double* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double* t = T + (size_t)i * 16;
  double4 o;
  o.x = t[3];
  o.y = t[7];
  o.z = atan2(t[4], t[0]);
  o.w = err ? err[i] : 0.0;
  reinterpret_cast<double4*>(out)[i] = o;
}

// Arrival-flag wait of the fused exchange: one thread per flag word, system-scope acquire loads.
__global__ void exchange_wait_kernel(const unsigned int* flags, int n_flags, unsigned int value,
                                     unsigned long long timeout_ns, unsigned int* timed_out) {
  const int k = threadIdx.x;
  if (k >= n_flags) return;
  unsigned long long t0 = 0;
  while (ld_acquire_sys(flags + k) < value) {
    __nanosleep(200);
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    if (t0 == 0) t0 = t;
    else if (t - t0 > timeout_ns) { if (timed_out) *timed_out = 1u; break; }
  }
}

__global__ void exchange_signal_kernel(const RunParams p) {
  if (threadIdx.x == 0) raise_arrival_flags(p);
}

}  // namespace

extern "C" {

int b200icp_version(void) { return B200ICP_VERSION; }

const char* b200icp_error_string(int code) {
  switch (code) {
    case B200ICP_OK: return "ok";
    case B200ICP_ERR_BAD_ARG: return "bad argument (null pointer, non-positive size, dim not 2/3, or misaligned scans)";
    case B200ICP_ERR_TOO_LARGE: return "problem too large for this build";
    case B200ICP_ERR_WORKSPACE: return "workspace missing or smaller than b200icp_workspace_bytes";
    case B200ICP_ERR_CUDA: return "CUDA launch failed";
    case B200ICP_ERR_NO_DEVICE: return "no sm_100 (B200) device is current";
    case B200ICP_ERR_BUSY: return "two submitted host calls are already in flight on this device; wait for one first";
    case B200ICP_ERR_STATIC_SCRATCH:
      return "b200icp_run_static: problem needs more scratch than B200ICP_STATIC_SCRATCH_BYTES (16 MiB); "
             "use b200icp_run with a caller workspace";
    default: return "unknown error code";
  }
}

int b200icp_workspace_bytes(int n_pairs, int n_max, int dim, size_t* bytes_out) {
  if (!bytes_out || n_pairs < 0 || n_max <= 0 || (dim != 2 && dim != 3)) return B200ICP_ERR_BAD_ARG;
  size_t bytes = 256;  // the work-queue counter (padded)
  if (uses_general_kernel(n_max, dim))   // control block, per-pair state, ring, partial sums, one prepared copy per referenced scan
    bytes = flow_layout(n_pairs < kFlowMaxPairs ? (n_pairs > 0 ? n_pairs : 1) : kFlowMaxPairs, n_max).total;
  *bytes_out = bytes;
  return B200ICP_OK;
}

static int apply_exchange(RunParams& p, const b200icp_exchange* ex) {
  if (!ex) return B200ICP_OK;
  if (ex->n_peers < 0 || ex->n_peers > kMaxPeers || (ex->n_peers > 0 && !ex->peer_pose_out) || ex->peer_row0 < 0 ||
      ex->flag_offset_bytes < 0 || (ex->flag_offset_bytes & 3) || ex->flag_index < 0)
    return B200ICP_ERR_BAD_ARG;
  p.n_peers = ex->n_peers; p.peer_row0 = ex->peer_row0;
  p.flag_off = ex->n_peers > 0 ? ex->flag_offset_bytes : 0; p.flag_idx = ex->flag_index; p.flag_val = ex->flag_value;
  for (int k = 0; k < ex->n_peers; ++k) {
    if (!ex->peer_pose_out[k] || (reinterpret_cast<uintptr_t>(ex->peer_pose_out[k]) & 31u)) return B200ICP_ERR_BAD_ARG;
    p.peer_out[k] = ex->peer_pose_out[k];
  }
  return B200ICP_OK;
}

static int run_common(const double* scans, const int32_t* counts, int n_scans, int n_max, int dim,
                      const int32_t* pairs, const double* init, const double* epsilon,
                      const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                      double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                      int32_t* corr_out, void* workspace, size_t workspace_bytes, void* stream,
                      const b200icp_exchange* exchange,
                      const unsigned int* ready_scans = nullptr, unsigned flags = 0) {
  if (!init || !epsilon || !normalize || !T_out || max_iters < 0) return B200ICP_ERR_BAD_ARG;
  if (ready_scans && uses_general_kernel(n_max, dim)) return B200ICP_ERR_BAD_ARG;   // planar kernel only
  RunParams p{};
  p.scans = scans; p.counts = counts; p.n_scans = n_scans; p.n_max = n_max;
  p.pairs = pairs; p.init = init; p.epsilon = epsilon; p.normalize = normalize;
  p.n_pairs = n_pairs; p.max_iters = max_iters; p.stopping_thresh = stopping_thresh;
  p.T_out = T_out; p.err_out = err_out; p.iters_out = iters_out; p.status_out = status_out;
  p.corr_out = corr_out; p.dist2_out = nullptr; p.nearest_only = 0;
  p.ready_scans = ready_scans;
  p.gate_timeout_ns = 5000000000ull;
  const int rc_ex = apply_exchange(p, exchange);
  if (rc_ex != B200ICP_OK) return rc_ex;
  return dispatch(p, dim, workspace, workspace_bytes, static_cast<cudaStream_t>(stream), flags);
}

int b200icp_run_ex(const double* scans, const int32_t* counts, int n_scans, int n_max, int dim,
                   const int32_t* pairs, const double* init, const double* epsilon,
                   const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                   double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                   int32_t* corr_out, void* workspace, size_t workspace_bytes, void* stream,
                   const unsigned int* ready_scans, const b200icp_exchange* exchange, unsigned flags) {
  return run_common(scans, counts, n_scans, n_max, dim, pairs, init, epsilon, normalize, n_pairs, max_iters,
                    stopping_thresh, T_out, err_out, iters_out, status_out, corr_out, workspace,
                    workspace_bytes, stream, exchange, ready_scans, flags);
}

int b200icp_run_ranges(const void* ranges, int ranges_dtype, int n_scans, int n_beams, const double* cos_tab,
                       const double* sin_tab, double range_min, double range_max, double off_x, double off_y,
                       const int32_t* pairs, const double* init, const double* epsilon,
                       const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                       double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                       int32_t* corr_out, int32_t* counts_out, void* workspace, size_t workspace_bytes,
                       void* stream, const unsigned int* ready_scans, const b200icp_exchange* exchange,
                       unsigned flags) {
  if (!ranges || (ranges_dtype != 0 && ranges_dtype != 1) || !cos_tab || !sin_tab || n_beams <= 0 ||
      n_beams > kMax2d || !init || !epsilon || !normalize || !T_out || max_iters < 0)
    return B200ICP_ERR_BAD_ARG;
  if ((reinterpret_cast<uintptr_t>(ranges) & (ranges_dtype ? 3u : 7u)) != 0) return B200ICP_ERR_BAD_ARG;
  RunParams p{};
  p.ranges = ranges; p.ranges_f32 = ranges_dtype; p.n_beams = n_beams; p.cos_tab = cos_tab; p.sin_tab = sin_tab;
  p.range_min = range_min; p.range_max = range_max; p.off_x = off_x; p.off_y = off_y; p.counts_out = counts_out;
  p.n_scans = n_scans; p.n_max = n_beams + (n_beams & 1);
  p.pairs = pairs; p.init = init; p.epsilon = epsilon; p.normalize = normalize;
  p.n_pairs = n_pairs; p.max_iters = max_iters; p.stopping_thresh = stopping_thresh;
  p.T_out = T_out; p.err_out = err_out; p.iters_out = iters_out; p.status_out = status_out;
  p.corr_out = corr_out;
  p.ready_scans = ready_scans; p.gate_timeout_ns = 5000000000ull;
  const int rc_ex = apply_exchange(p, exchange);
  if (rc_ex != B200ICP_OK) return rc_ex;
  return dispatch(p, 2, workspace, workspace_bytes, static_cast<cudaStream_t>(stream), flags);
}

int b200icp_run(const double* scans, const int32_t* counts, int n_scans, int n_max, int dim,
                const int32_t* pairs, const double* init, const double* epsilon,
                const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                int32_t* corr_out, void* workspace, size_t workspace_bytes, void* stream) {
  return run_common(scans, counts, n_scans, n_max, dim, pairs, init, epsilon, normalize, n_pairs, max_iters,
                    stopping_thresh, T_out, err_out, iters_out, status_out, corr_out, workspace,
                    workspace_bytes, stream, nullptr);
}

int b200icp_run_exchange(const double* scans, const int32_t* counts, int n_scans, int n_max, int dim,
                         const int32_t* pairs, const double* init, const double* epsilon,
                         const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                         double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                         int32_t* corr_out, void* workspace, size_t workspace_bytes, void* stream,
                         double* const* peer_pose_out, int n_peers, long long peer_row0) {
  const b200icp_exchange ex{peer_pose_out, n_peers, peer_row0, 0, 0, 0u};
  return run_common(scans, counts, n_scans, n_max, dim, pairs, init, epsilon, normalize, n_pairs, max_iters,
                    stopping_thresh, T_out, err_out, iters_out, status_out, corr_out, workspace,
                    workspace_bytes, stream, &ex);
}

int b200icp_run_gated(const double* scans, const int32_t* counts, int n_scans, int n_max, int dim,
                      const int32_t* pairs, const double* init, const double* epsilon,
                      const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                      double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                      int32_t* corr_out, void* workspace, size_t workspace_bytes, void* stream,
                      const unsigned int* ready_scans, double* const* peer_pose_out, int n_peers,
                      long long peer_row0) {
  if (!ready_scans) return B200ICP_ERR_BAD_ARG;
  const b200icp_exchange ex{peer_pose_out, n_peers, peer_row0, 0, 0, 0u};
  return run_common(scans, counts, n_scans, n_max, dim, pairs, init, epsilon, normalize, n_pairs, max_iters,
                    stopping_thresh, T_out, err_out, iters_out, status_out, corr_out, workspace,
                    workspace_bytes, stream, &ex, ready_scans);
}

int b200icp_nearest(const double* scans, const int32_t* counts, int n_scans, int n_max, int dim,
                    const int32_t* pairs, int n_pairs, int32_t* corr_out, double* dist2_out,
                    void* workspace, size_t workspace_bytes, void* stream) {
  if (!corr_out) return B200ICP_ERR_BAD_ARG;
  RunParams p{};
  p.scans = scans; p.counts = counts; p.n_scans = n_scans; p.n_max = n_max;
  p.pairs = pairs; p.n_pairs = n_pairs; p.max_iters = 0; p.stopping_thresh = 0.0;
  p.corr_out = corr_out; p.dist2_out = dist2_out; p.nearest_only = 1;
  return dispatch(p, dim, workspace, workspace_bytes, static_cast<cudaStream_t>(stream), 0);
}

int b200icp_pack_poses(const double* T, const double* err, int n, double* out, void* stream) {
  if (!T || !out || n < 0) return B200ICP_ERR_BAD_ARG;
  if (n == 0) return B200ICP_OK;
  pack_poses_kernel<<<(n + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(T, err, n, out);
  return cudaGetLastError() == cudaSuccess ? B200ICP_OK : B200ICP_ERR_CUDA;
}

int b200icp_exchange_wait(const unsigned int* flags, int n_flags, unsigned int value, int timeout_ms,
                          unsigned int* timed_out, void* stream) {
  if (!flags || n_flags <= 0 || n_flags > kMaxPeers || timeout_ms <= 0) return B200ICP_ERR_BAD_ARG;
  exchange_wait_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(flags, n_flags, value,
                                                                        (unsigned long long)timeout_ms * 1000000ull, timed_out);
  return cudaGetLastError() == cudaSuccess ? B200ICP_OK : B200ICP_ERR_CUDA;
}

int b200icp_exchange_signal(const b200icp_exchange* exchange, void* stream) {
  if (!exchange || exchange->n_peers <= 0 || exchange->flag_offset_bytes <= 0) return B200ICP_ERR_BAD_ARG;
  RunParams p{};
  const int rc = apply_exchange(p, exchange);
  if (rc != B200ICP_OK) return rc;
  exchange_signal_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(p);
  return cudaGetLastError() == cudaSuccess ? B200ICP_OK : B200ICP_ERR_CUDA;
}

int b200icp_ffma_probe(float* sink, int n_blocks, int iters, void* stream) {
  if (!sink || n_blocks <= 0 || iters <= 0) return B200ICP_ERR_BAD_ARG;
  ffma_probe_kernel<<<n_blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(sink, iters);
  return cudaGetLastError() == cudaSuccess ? B200ICP_OK : B200ICP_ERR_CUDA;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// b200icp_run_static: b200icp_run with its scratch in the library's device image (no allocation
// anywhere on the call path).  No counterpart in the reference; added
// for callers that cannot call cudaMalloc (graph capture, serving loops).
//
// In host code `g_static_scratch` is only a host-side stand-in for the device array: its device
// address is looked up once per device with cudaGetSymbolAddress (with that device current) and
// that pointer is what the kernels get.  Every call on a device shares the array; calls that
// overlap on the GPU race (include/b200icp.h).
// ---------------------------------------------------------------------------------------------
#include <mutex>

static __device__ __align__(256) unsigned char g_static_scratch[B200ICP_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mutex;
void* g_static_ptr[kStaticMaxDevices];        // device address of g_static_scratch, per device ordinal

int static_scratch_for_current_device(void** out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kStaticMaxDevices) return B200ICP_ERR_NO_DEVICE;
  std::lock_guard<std::mutex> lock(g_static_mutex);
  if (!g_static_ptr[dev]) {
    void* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_scratch);
    if (e != cudaSuccess || !ptr) {
      (void)cudaGetLastError();
      return e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? B200ICP_ERR_NO_DEVICE : B200ICP_ERR_CUDA;
    }
    g_static_ptr[dev] = ptr;
  }
  *out = g_static_ptr[dev];
  return B200ICP_OK;
}
}  // namespace

extern "C" int b200icp_run_static(const double* scans, const int32_t* counts, int n_scans, int n_max, int dim,
                                  const int32_t* pairs, const double* init, const double* epsilon,
                                  const uint8_t* normalize, int n_pairs, int max_iters, double stopping_thresh,
                                  double* T_out, double* err_out, int32_t* iters_out, int32_t* status_out,
                                  int32_t* corr_out, void* stream) {
  size_t need = 0;
  const int rc = b200icp_workspace_bytes(n_pairs, n_max, dim, &need);
  if (rc != B200ICP_OK) return rc;
  if (need > (size_t)B200ICP_STATIC_SCRATCH_BYTES) return B200ICP_ERR_STATIC_SCRATCH;
  void* scratch = nullptr;
  const int rs = static_scratch_for_current_device(&scratch);
  if (rs != B200ICP_OK) return rs;
  // dispatch() clears the control block (first 256 bytes) on `stream`; the setup kernel of the general
  // path rewrites ring, per-pair state and slot headers, so nothing else has to start at zero.
  return run_common(scans, counts, n_scans, n_max, dim, pairs, init, epsilon, normalize, n_pairs, max_iters,
                    stopping_thresh, T_out, err_out, iters_out, status_out, corr_out, scratch,
                    (size_t)B200ICP_STATIC_SCRATCH_BYTES, stream, nullptr);
}
