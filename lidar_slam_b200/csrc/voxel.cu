// Voxel-grid down-sampling on the device; replaces voxel_downsample
// (modules/icp.py:4-27, duplicate code/icp_warm_up/icp.py:4-34):
//   idx = floor((p - min(p)) / voxel_size)                       icp.py:19
//   group by idx, lexicographic (x,y,z) order of the voxel ids    icp.py:22 (np.unique axis=0)
//   mean of the points of each voxel                              icp.py:24-26
// A stable radix sort on the packed 63-bit voxel key keeps the points of a voxel in
// their original order, and the per-voxel sum runs sequentially in that order, which is
// the order numpy's axis-0 reduction uses -- so the means are bit-identical.
#include <cuda_runtime.h>
#include <cub/cub.cuh>

#include "../../include/b200icp.h"

namespace {

constexpr int kAxisBits = 21;  // voxel index range per axis: [0, 2^21)

struct VoxelWs {
  double* mins;            // [3] (+1 pad)
  unsigned long long* keys_in;
  unsigned long long* keys_out;
  int* idx_in;
  int* idx_out;
  int* head;               // [n] 1 at the first point of each voxel run
  int* slot;               // [n] inclusive scan of head
  int* overflow;           // [1]
  void* cub_temp;
  size_t cub_bytes;
};

size_t align_up(size_t v) { return (v + 255) & ~(size_t)255; }

size_t cub_temp_bytes(int n) {
  size_t sort_bytes = 0, scan_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                  (int*)nullptr, (int*)nullptr, n, 0, 3 * kAxisBits);
  cub::DeviceScan::InclusiveSum(nullptr, scan_bytes, (int*)nullptr, (int*)nullptr, n);
  return sort_bytes > scan_bytes ? sort_bytes : scan_bytes;
}

size_t carve(VoxelWs& w, void* base, int n) {
  size_t o = 0;
  auto take = [&](size_t bytes) { void* p = base ? (char*)base + o : nullptr; o += align_up(bytes); return p; };
  w.mins = (double*)take(4 * sizeof(double));
  w.keys_in = (unsigned long long*)take((size_t)n * 8);
  w.keys_out = (unsigned long long*)take((size_t)n * 8);
  w.idx_in = (int*)take((size_t)n * 4);
  w.idx_out = (int*)take((size_t)n * 4);
  w.head = (int*)take((size_t)n * 4);
  w.slot = (int*)take((size_t)n * 4);
  w.overflow = (int*)take(4);
  w.cub_bytes = cub_temp_bytes(n);
  w.cub_temp = take(w.cub_bytes);
  return o;
}

__global__ void min_kernel(const double* pts, int n, double* mins) {
  __shared__ double sh[3][32];
  double m[3] = {1.0 / 0.0, 1.0 / 0.0, 1.0 / 0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    for (int a = 0; a < 3; ++a) m[a] = fmin(m[a], pts[3 * (size_t)i + a]);
  for (int a = 0; a < 3; ++a)
    for (int o = 16; o > 0; o >>= 1) m[a] = fmin(m[a], __shfl_xor_sync(0xffffffffu, m[a], o));
  if ((threadIdx.x & 31) == 0)
    for (int a = 0; a < 3; ++a) sh[a][threadIdx.x >> 5] = m[a];
  __syncthreads();
  if (threadIdx.x < 3) {
    double v = sh[threadIdx.x][0];
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) v = fmin(v, sh[threadIdx.x][w]);
    // single block launch: plain store
    mins[threadIdx.x] = v;
  }
}

__global__ void key_kernel(const double* pts, int n, double voxel, const double* mins,
                           unsigned long long* keys, int* idx, int* overflow) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  unsigned long long key = 0;
  for (int a = 0; a < 3; ++a) {
    const double v = floor((pts[3 * (size_t)i + a] - mins[a]) / voxel);
    long long c = (long long)v;
    if (!(v >= 0.0) || v >= (double)(1 << kAxisBits)) { *overflow = 1; c = 0; }
    key = (key << kAxisBits) | (unsigned long long)c;
  }
  keys[i] = key;
  idx[i] = i;
}

__global__ void head_kernel(const unsigned long long* keys, int n, int* head) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) head[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
}

__global__ void mean_kernel(const double* pts, const unsigned long long* keys, const int* order,
                            const int* head, const int* slot, int n, double* out, int* n_out,
                            const int* overflow) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) *n_out = *overflow ? -1 : slot[n - 1];
  if (i >= n || !head[i]) return;
  double sx = 0.0, sy = 0.0, sz = 0.0;
  int cnt = 0;
  const unsigned long long k = keys[i];
  for (int j = i; j < n && keys[j] == k; ++j) {
    const double* q = pts + 3 * (size_t)order[j];
    sx += q[0]; sy += q[1]; sz += q[2];
    ++cnt;
  }
  double* o = out + 3 * (size_t)(slot[i] - 1);
  o[0] = sx / (double)cnt; o[1] = sy / (double)cnt; o[2] = sz / (double)cnt;
}

}  // namespace

extern "C" {

int b200icp_voxel_workspace_bytes(int n_points, size_t* bytes_out) {
  if (!bytes_out || n_points <= 0) return B200ICP_ERR_BAD_ARG;
  VoxelWs w;
  *bytes_out = carve(w, nullptr, n_points);
  return B200ICP_OK;
}

int b200icp_voxel_downsample(const double* points, int n_points, double voxel_size, double* out,
                             int32_t* n_out, void* workspace, size_t workspace_bytes, void* stream_) {
  if (!points || !out || !n_out || n_points <= 0 || !(voxel_size > 0.0)) return B200ICP_ERR_BAD_ARG;
  if (!workspace) return B200ICP_ERR_WORKSPACE;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  VoxelWs w;
  if (carve(w, workspace, n_points) > workspace_bytes) return B200ICP_ERR_WORKSPACE;
  const int n = n_points, tb = 256, gb = (n + tb - 1) / tb;
  cudaMemsetAsync(w.overflow, 0, 4, stream);
  min_kernel<<<1, 1024, 0, stream>>>(points, n, w.mins);
  key_kernel<<<gb, tb, 0, stream>>>(points, n, voxel_size, w.mins, w.keys_in, w.idx_in, w.overflow);
  size_t tmp = w.cub_bytes;
  cub::DeviceRadixSort::SortPairs(w.cub_temp, tmp, w.keys_in, w.keys_out, w.idx_in, w.idx_out, n, 0,
                                  3 * kAxisBits, stream);
  head_kernel<<<gb, tb, 0, stream>>>(w.keys_out, n, w.head);
  tmp = w.cub_bytes;
  cub::DeviceScan::InclusiveSum(w.cub_temp, tmp, w.head, w.slot, n, stream);
  mean_kernel<<<gb, tb, 0, stream>>>(points, w.keys_out, w.idx_out, w.head, w.slot, n, out, n_out, w.overflow);
  return cudaGetLastError() == cudaSuccess ? B200ICP_OK : B200ICP_ERR_CUDA;
}

}  // extern "C"
