// Texture mapping on the device (second half): RGB-D frames are
// back-projected, carried to the world frame with the robot pose and splatted onto the grid.
//
//   texture_paint_kernel    the frame loop of generate_texture_map     modules/texture_mapping.py:49-96
//                           get_depth_image                            modules/texture_mapping.py:132-143
//                           get_rgbi_rgbj                              modules/texture_mapping.py:145-160
//                           vectorized_generate_point_cloud            modules/texture_mapping.py:162-224
//                           transform_point_cloud / transform_points   texture_mapping.py:226-240, utils.py:140-159
//   texture_resolve_kernel  initial map + astype(float32)/255          modules/texture_mapping.py:46-47, :98
//
// The reference paints frame after frame with a fancy-index assignment, so a cell keeps the colour
// of the LAST frame that reaches it and, inside that frame, of the LAST pixel in row-major order.
// That is a maximum over (frame, pixel), so every pixel of every frame is processed in parallel
// and folded into the cell with one 64-bit atomicMax on the key
//     (frame + 1) << 44 | pixel << 24 | r << 16 | g << 8 | b .
// Arithmetic follows the reference's dtypes step by step: the depth chain is float32
// (disparity is float32 and numpy keeps python scalars weak), everything from the RGB
// look-up coordinates on is float64.  HBM-bound byte work: 4 B disparity + 3 B colour per pixel.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200icp.h"

namespace {

struct TexParams {
  const float* disparity;    // [n_frames][h][w] float32 (read_image(..., is_disparity=True))
  const uint8_t* rgb;        // [n_frames][h][w][3] uint8, RGB order
  const double* T_wr;        // [n_frames][16] world <- robot, built by the caller from the pose
  double Kinv[9];            // np.linalg.inv(K)
  double T_co[16];           // np.linalg.inv(T_oc): camera <- optical
  double T_rc[16];           // robot <- camera
  int n_frames, h, w, frame_base;
  int width, height;         // grid
  double res, min_x, min_y;
  unsigned long long* keys;  // [width*height]
};

__device__ __forceinline__ void xform(const double* T, double& x, double& y, double& z) {
  // (T @ [x y z 1]^T)[:3], products and sums rounded one by one, left to right
  const double nx = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T[0], x), __dmul_rn(T[1], y)), __dmul_rn(T[2], z)), T[3]);
  const double ny = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T[4], x), __dmul_rn(T[5], y)), __dmul_rn(T[6], z)), T[7]);
  const double nz = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T[8], x), __dmul_rn(T[9], y)), __dmul_rn(T[10], z)), T[11]);
  x = nx; y = ny; z = nz;
}

__global__ void __launch_bounds__(256) texture_paint_kernel(const TexParams p) {
  __shared__ double sT[16];
  const int frame = blockIdx.y;
  if (threadIdx.x < 16) sT[threadIdx.x] = p.T_wr[(size_t)frame * 16 + threadIdx.x];
  __syncthreads();
  const int npix = p.h * p.w;
  const float* disp = p.disparity + (size_t)frame * npix;
  const uint8_t* rgb = p.rgb + (size_t)frame * npix * 3;
  const unsigned long long fkey = (unsigned long long)(p.frame_base + frame + 1) << 44;
  for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < npix; pix += gridDim.x * blockDim.x) {
    const int i = pix / p.w, j = pix - i * p.w;
    // texture_mapping.py:142-143 in float32: dd = (-0.00304 * d) + 3.31 ; depth = 1.03 / dd
    const float dd = __fadd_rn(__fmul_rn(-0.00304f, disp[pix]), 3.31f);
    const float depth = __fdiv_rn(1.03f, dd);
    // texture_mapping.py:158-159 (called with the DEPTH as `dd`, :200): float64 except the
    // float32 product 7877.07 * depth
    const double t = (double)__fmul_rn(7877.07f, depth);
    const double rgbi = __ddiv_rn(__dsub_rn(__dadd_rn(__dmul_rn(526.37, (double)i), 19276.0), t), 585.051);
    const double rgbj = __ddiv_rn(__dadd_rn(__dmul_rn(526.37, (double)j), 16662.0), 585.051);
    if (!(0.0 <= rgbi && rgbi < (double)p.h && 0.0 <= rgbj && rgbj < (double)p.w)) continue;   // :203
    int ri = (int)rgbi, rj = (int)rgbj;                                                          // :206-207
    ri = min(max(ri, 0), p.h - 1); rj = min(max(rj, 0), p.w - 1);
    // :195-197  xyz = (K_inv @ [j, i, 1]) * depth
    const double fj = (double)j, fi = (double)i, dz = (double)depth;
    double x = __dmul_rn(__dadd_rn(__dadd_rn(__dmul_rn(p.Kinv[0], fj), __dmul_rn(p.Kinv[1], fi)), p.Kinv[2]), dz);
    double y = __dmul_rn(__dadd_rn(__dadd_rn(__dmul_rn(p.Kinv[3], fj), __dmul_rn(p.Kinv[4], fi)), p.Kinv[5]), dz);
    double z = __dmul_rn(__dadd_rn(__dadd_rn(__dmul_rn(p.Kinv[6], fj), __dmul_rn(p.Kinv[7], fi)), p.Kinv[8]), dz);
    xform(p.T_co, x, y, z);      // optical -> camera     :222
    xform(p.T_rc, x, y, z);      // camera  -> robot      :68
    xform(sT, x, y, z);          // robot   -> world      :82
    if (!(isfinite(x) && isfinite(y))) continue;           // depth = inf/nan when dd == 0: never in range
    // world2grid (ogm.py:118-119) and the in-grid test (:91-92)
    const double gxf = ceil(__ddiv_rn(__dsub_rn(x, p.min_x), p.res));
    const double gyf = ceil(__ddiv_rn(__dsub_rn(y, p.min_y), p.res));
    if (!(gxf >= 1.0 && gxf <= (double)p.width && gyf >= 1.0 && gyf <= (double)p.height)) continue;
    const int gx = (int)gxf - 1, gy = (int)gyf - 1;
    const uint8_t* c = rgb + ((size_t)ri * p.w + rj) * 3;
    const unsigned long long key = fkey | ((unsigned long long)pix << 24) |
                                   ((unsigned long long)c[0] << 16) | ((unsigned long long)c[1] << 8) | c[2];
    atomicMax(&p.keys[(size_t)gx * p.height + gy], key);
  }
}

// texture = painted ? colour : grid_map value on all three channels; then float32 / 255
__global__ void texture_resolve_kernel(const unsigned long long* __restrict__ keys,
                                       const uint8_t* __restrict__ grid, float* __restrict__ out,
                                       long long n_cells) {
  const long long c = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_cells) return;
  const unsigned long long k = keys[c];
  unsigned r, g, b;
  if (k) { r = (unsigned)(k >> 16) & 255u; g = (unsigned)(k >> 8) & 255u; b = (unsigned)k & 255u; }
  else { r = g = b = grid ? grid[c] : 0u; }
  out[3 * c + 0] = __fdiv_rn((float)r, 255.0f);
  out[3 * c + 1] = __fdiv_rn((float)g, 255.0f);
  out[3 * c + 2] = __fdiv_rn((float)b, 255.0f);
}

}  // namespace

extern "C" {

int b200icp_texture_paint(unsigned long long* keys, int width, int height, double res, double min_x,
                          double min_y, const float* disparity, const uint8_t* rgb, int n_frames,
                          int img_h, int img_w, int frame_base, const double* T_wr,
                          const double* K_inv, const double* T_cam_from_optical,
                          const double* T_robot_from_cam, void* stream) {
  if (!keys || !disparity || !rgb || !T_wr || !K_inv || !T_cam_from_optical || !T_robot_from_cam ||
      width <= 0 || height <= 0 || !(res > 0.0) || n_frames < 0 || img_h <= 0 || img_w <= 0 ||
      frame_base < 0)
    return B200ICP_ERR_BAD_ARG;
  if ((long long)img_h * img_w >= (1 << 20) || (long long)frame_base + n_frames >= (1 << 20) - 1)
    return B200ICP_ERR_TOO_LARGE;
  if (n_frames == 0) return B200ICP_OK;
  TexParams p{};
  p.disparity = disparity; p.rgb = rgb; p.T_wr = T_wr;
  for (int k = 0; k < 9; ++k) p.Kinv[k] = K_inv[k];
  for (int k = 0; k < 16; ++k) { p.T_co[k] = T_cam_from_optical[k]; p.T_rc[k] = T_robot_from_cam[k]; }
  p.n_frames = n_frames; p.h = img_h; p.w = img_w; p.frame_base = frame_base;
  p.width = width; p.height = height; p.res = res; p.min_x = min_x; p.min_y = min_y; p.keys = keys;
  const int npix = img_h * img_w;
  int bx = (npix + 255) / 256;
  if (bx > 148 * 8) bx = 148 * 8;
  for (int f0 = 0; f0 < n_frames; f0 += 65535) {          // gridDim.y limit
    const int nf = n_frames - f0 < 65535 ? n_frames - f0 : 65535;
    TexParams q = p;
    q.disparity = disparity + (size_t)f0 * npix; q.rgb = rgb + (size_t)f0 * npix * 3;
    q.T_wr = T_wr + (size_t)f0 * 16; q.frame_base = frame_base + f0; q.n_frames = nf;
    texture_paint_kernel<<<dim3(bx, nf), 256, 0, static_cast<cudaStream_t>(stream)>>>(q);
  }
  return cudaGetLastError() == cudaSuccess ? B200ICP_OK : B200ICP_ERR_CUDA;
}

int b200icp_texture_resolve(const unsigned long long* keys, const uint8_t* grid, float* texture_out,
                            int width, int height, void* stream) {
  if (!keys || !texture_out || width <= 0 || height <= 0) return B200ICP_ERR_BAD_ARG;
  const long long n = (long long)width * height;
  texture_resolve_kernel<<<(unsigned)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      keys, grid, texture_out, n);
  return cudaGetLastError() == cudaSuccess ? B200ICP_OK : B200ICP_ERR_CUDA;
}

}  // extern "C"
