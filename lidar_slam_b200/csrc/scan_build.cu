// Scan construction on the device: the producer of the scan
// pool the ICP kernels read.
//
//   ranges_to_scans_kernel   get_lidar_data            modules/localization.py:252-307
//   range_stats / filter     statistical_outliers_removal modules/localization.py:223-250
//
// Arithmetic is float64 with explicit round-to-nearest intrinsics (no FMA contraction), in the
// reference's operation order, so the pool is bit-identical to the reference's ragged list:
//   x = fl(fl(r * cos a) + 0.13323),  y = fl(fl(r * sin a) + 0.0)       localization.py:290-301
// cos a / sin a come from a caller-supplied table (numpy's libm values on the host), because the
// device's cos()/sin() are not bit-identical to glibc's.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200icp.h"

namespace {

constexpr int kStatBlocks = 592;   // 4 x 148 SMs
constexpr int kStatThreads = 256;

// One warp per scan: gate the beams, compact the survivors (order kept), write the padded row.
__global__ void ranges_to_scans_kernel(const double* __restrict__ ranges, int n_scans, int n_beams,
                                       const double* __restrict__ cos_tab,
                                       const double* __restrict__ sin_tab, double range_min,
                                       double range_max, double off_x, double off_y,
                                       double* __restrict__ scans, int32_t* __restrict__ counts,
                                       int n_max) {
  const int lane = threadIdx.x & 31;
  const int scan = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (scan >= n_scans) return;
  const double* r_row = ranges + (size_t)scan * n_beams;
  double2* out = reinterpret_cast<double2*>(scans + (size_t)scan * n_max * 2);
  int base = 0;
  for (int b0 = 0; b0 < n_beams; b0 += 32) {
    const int b = b0 + lane;
    double r = 0.0;
    bool valid = false;
    if (b < n_beams) {
      r = r_row[b];
      valid = (r >= range_min) && (r <= range_max);   // NaN fails both, as in numpy
    }
    const unsigned m = __ballot_sync(0xffffffffu, valid);
    if (valid) {
      const int pos = base + __popc(m & ((1u << lane) - 1u));
      if (pos < n_max) {
        double2 p;
        p.x = __dadd_rn(__dmul_rn(r, cos_tab[b]), off_x);
        p.y = __dadd_rn(__dmul_rn(r, sin_tab[b]), off_y);
        out[pos] = p;
      }
    }
    base += __popc(m);
  }
  const int n = base < n_max ? base : n_max;
  for (int i = n + lane; i < n_max; i += 32) out[i] = make_double2(0.0, 0.0);
  if (lane == 0) counts[scan] = n;
}

__device__ __forceinline__ double point_range(double2 p) {
  // np.sqrt(np.sum(points**2, axis=1)): x*x + y*y without contraction, IEEE sqrt
  return __dsqrt_rn(__dadd_rn(__dmul_rn(p.x, p.x), __dmul_rn(p.y, p.y)));
}

__device__ __forceinline__ double block_sum(double v, double* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  double t = 0.0;
  if (warp == 0) {
    t = lane < (blockDim.x >> 5) ? sh[lane] : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  }
  __syncthreads();
  return t;   // valid in warp 0
}

// pass 0: partial sums of d and the point count; pass 1: partial sums of (d - mean)^2.
// Deterministic: fixed grid, per-block partials, combined in block order by the finish kernel.
__global__ void range_stats_kernel(const double* __restrict__ scans,
                                   const int32_t* __restrict__ counts, int n_scans, int n_max,
                                   int pass, const double* __restrict__ stats,
                                   double* __restrict__ partial) {
  __shared__ double sh[32];
  const double mean = pass ? stats[0] : 0.0;
  double acc = 0.0, cnt = 0.0;
  const int warps_per_block = blockDim.x >> 5;
  const int lane = threadIdx.x & 31;
  for (int s = blockIdx.x * warps_per_block + (threadIdx.x >> 5); s < n_scans;
       s += gridDim.x * warps_per_block) {
    const int n = min(counts[s], n_max);
    const double2* row = reinterpret_cast<const double2*>(scans + (size_t)s * n_max * 2);
    for (int i = lane; i < n; i += 32) {
      const double d = point_range(row[i]);
      if (pass) { const double e = d - mean; acc += e * e; } else { acc += d; }
    }
    if (lane == 0) cnt += (double)n;
  }
  const double a = block_sum(acc, sh);
  const double c = block_sum(cnt, sh);
  if (threadIdx.x == 0) { partial[2 * blockIdx.x] = a; partial[2 * blockIdx.x + 1] = c; }
}

// stats = {mean, std, threshold, n_points}
__global__ void range_stats_finish_kernel(const double* __restrict__ partial, int n_blocks, int pass,
                                          double k_std, double* __restrict__ stats) {
  __shared__ double sh[32];
  double a = 0.0, c = 0.0;
  for (int i = threadIdx.x; i < n_blocks; i += blockDim.x) { a += partial[2 * i]; c += partial[2 * i + 1]; }
  a = block_sum(a, sh);
  c = block_sum(c, sh);
  if (threadIdx.x == 0) {
    if (pass == 0) {
      stats[3] = c;
      stats[0] = a / c;
    } else {
      stats[1] = sqrt(a / stats[3]);
      stats[2] = stats[0] + k_std * stats[1];
    }
  }
}

// One warp per scan: keep the points with range < threshold, in place, order kept.
__global__ void range_filter_kernel(double* __restrict__ scans, int32_t* __restrict__ counts,
                                    int n_scans, int n_max, const double* __restrict__ stats) {
  const int lane = threadIdx.x & 31;
  const int scan = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (scan >= n_scans) return;
  const double thr = stats[2];
  double2* row = reinterpret_cast<double2*>(scans + (size_t)scan * n_max * 2);
  const int n = min(counts[scan], n_max);
  int base = 0;
  for (int i0 = 0; i0 < n; i0 += 32) {
    const int i = i0 + lane;
    double2 p = make_double2(0.0, 0.0);
    bool keep = false;
    if (i < n) {
      p = row[i];
      keep = point_range(p) < thr;
    }
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    __syncwarp();                       // every lane has read its point before any lane writes
    if (keep) row[base + __popc(m & ((1u << lane) - 1u))] = p;
    base += __popc(m);
    __syncwarp();
  }
  for (int i = base + lane; i < n; i += 32) row[i] = make_double2(0.0, 0.0);
  if (lane == 0) counts[scan] = base;
}

}  // namespace

extern "C" {

int b200icp_scans_from_ranges(const double* ranges, int n_scans, int n_beams, const double* cos_tab,
                              const double* sin_tab, double range_min, double range_max,
                              double off_x, double off_y, double* scans_out, int32_t* counts_out,
                              int n_max, void* stream) {
  if (!ranges || !cos_tab || !sin_tab || !scans_out || !counts_out || n_scans < 0 || n_beams <= 0 ||
      n_max <= 0 || (n_max & 1) || (reinterpret_cast<uintptr_t>(scans_out) & 15u))
    return B200ICP_ERR_BAD_ARG;
  if (n_scans == 0) return B200ICP_OK;
  const int warps = 8;
  ranges_to_scans_kernel<<<(n_scans + warps - 1) / warps, warps * 32, 0,
                           static_cast<cudaStream_t>(stream)>>>(
      ranges, n_scans, n_beams, cos_tab, sin_tab, range_min, range_max, off_x, off_y, scans_out,
      counts_out, n_max);
  return cudaGetLastError() == cudaSuccess ? B200ICP_OK : B200ICP_ERR_CUDA;
}

int b200icp_range_filter_workspace_bytes(size_t* bytes_out) {
  if (!bytes_out) return B200ICP_ERR_BAD_ARG;
  *bytes_out = (size_t)kStatBlocks * 2 * sizeof(double);
  return B200ICP_OK;
}

int b200icp_range_filter(double* scans, int32_t* counts, int n_scans, int n_max, double k_std,
                         double* stats_out, void* workspace, size_t workspace_bytes, void* stream) {
  if (!scans || !counts || !stats_out || n_scans < 0 || n_max <= 0 ||
      (reinterpret_cast<uintptr_t>(scans) & 15u))
    return B200ICP_ERR_BAD_ARG;
  if (!workspace || workspace_bytes < (size_t)kStatBlocks * 2 * sizeof(double)) return B200ICP_ERR_WORKSPACE;
  if (n_scans == 0) return B200ICP_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  double* partial = static_cast<double*>(workspace);
  for (int pass = 0; pass < 2; ++pass) {
    range_stats_kernel<<<kStatBlocks, kStatThreads, 0, st>>>(scans, counts, n_scans, n_max, pass,
                                                             stats_out, partial);
    range_stats_finish_kernel<<<1, 256, 0, st>>>(partial, kStatBlocks, pass, k_std, stats_out);
  }
  const int warps = 8;
  range_filter_kernel<<<(n_scans + warps - 1) / warps, warps * 32, 0, st>>>(scans, counts, n_scans,
                                                                            n_max, stats_out);
  return cudaGetLastError() == cudaSuccess ? B200ICP_OK : B200ICP_ERR_CUDA;
}

}  // extern "C"
