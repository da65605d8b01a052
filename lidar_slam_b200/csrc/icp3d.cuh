// Fused general (3-D, or planar scans too large for the 2-D kernel) ICP: one CTA per
// pair, clouds stay in global memory (L2 resident), the target is streamed through a
// shared-memory fp32 tile.  Same exactness scheme as icp2d.cuh; the solve is the 3x3
// Kabsch of kabsch3.cuh.
//
// Replaces modules/icp.py:123-189 for the call shape of
// code/icp_warm_up/test_icp.py:78-88 (3-D clouds, 24 yaw initialisations).
#pragma once
#include "icp_common.cuh"
#include "kabsch3.cuh"

namespace b200icp {

constexpr int kTile3d = 2048;   // targets per shared-memory tile (SoA fp32: 24 KB)
constexpr int kSums3d = 16;     // sum u(3) v(3) u v^T(9) sse
constexpr int kRed3d = 22;      // + min/max per axis over moved source and matched target

__device__ __forceinline__ void load_point(const double* base, int dim, int i, double& x, double& y, double& z) {
  if (dim == 3) {
    x = base[3 * (size_t)i]; y = base[3 * (size_t)i + 1]; z = base[3 * (size_t)i + 2];
  } else {
    const double2 v = reinterpret_cast<const double2*>(base)[i];
    x = v.x; y = v.y; z = 0.0;
  }
}

struct Affine3 {  // rows 0..2 of the homogeneous transform
  double m[12];
  __device__ __forceinline__ void apply(double sx, double sy, double sz, double& x, double& y, double& z) const {
    x = fma(m[0], sx, fma(m[1], sy, fma(m[2], sz, m[3])));
    y = fma(m[4], sx, fma(m[5], sy, fma(m[6], sz, m[7])));
    z = fma(m[8], sx, fma(m[9], sy, fma(m[10], sz, m[11])));
  }
};

template <int R, int TH>
__global__ void __launch_bounds__(TH, 1) icp3d_kernel(const RunParams p, const int dim) {
  __shared__ __align__(16) float tile_x[kTile3d];
  __shared__ __align__(16) float tile_y[kTile3d];
  __shared__ __align__(16) float tile_z[kTile3d];
  __shared__ double red[2][TH / 32][kRed3d];
  __shared__ int s_pair;
  constexpr int W = TH / 32;
  const double kInf = __longlong_as_double(0x7ff0000000000000LL);
  constexpr float kTruncRel = 2.44140625e-4f;  // 2^-12 >= the 10 truncated mantissa bits (2^-13)
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  for (;;) {
    __syncthreads();
    if (tid == 0) s_pair = (int)atomicAdd(p.queue, 1u);
    __syncthreads();
    const int pair = s_pair;
    if (pair >= p.n_pairs) break;

    const int si = p.pairs[2 * pair], ti = p.pairs[2 * pair + 1];
    int early = -1, ns = 0, nt = 0;
    if (si < 0 || ti < 0 || si >= p.n_scans || ti >= p.n_scans) {
      early = B200ICP_STATUS_SKIPPED;
    } else {
      ns = min(p.counts[si], p.n_max);
      nt = min(p.counts[ti], p.n_max);
      if (ns <= 0 || nt <= 0) early = B200ICP_STATUS_EMPTY;
    }
    if (early >= 0) {
      if (tid == 0) {
        if (p.status_out) p.status_out[pair] = early;
        if (p.iters_out) p.iters_out[pair] = 0;
        if (p.err_out) p.err_out[pair] = __longlong_as_double(0x7ff8000000000000LL);
        if (p.T_out)
          for (int k = 0; k < 16; ++k) p.T_out[(size_t)pair * 16 + k] = p.init ? p.init[(size_t)pair * 16 + k] : 0.0;
      }
      if (p.corr_out)
        for (int j = tid; j < p.n_max; j += TH) p.corr_out[(size_t)pair * p.n_max + j] = -1;
      continue;
    }
    const double* src = p.scans + (size_t)si * p.n_max * dim;
    const double* tgt = p.scans + (size_t)ti * p.n_max * dim;

    Affine3 T;
    double bottom[4] = {0.0, 0.0, 0.0, 1.0};
#pragma unroll
    for (int k = 0; k < 12; ++k) T.m[k] = (k % 5 == 0) ? 1.0 : 0.0;
    if (!p.nearest_only) {
#pragma unroll
      for (int k = 0; k < 12; ++k) T.m[k] = p.init[(size_t)pair * 16 + k];
#pragma unroll
      for (int k = 0; k < 4; ++k) bottom[k] = p.init[(size_t)pair * 16 + 12 + k];
    }

    // ---- target statistics: max |coord| (fp32) and centroid (shift origin) -----
    float amax = 0.f;
    double cx = 0.0, cy = 0.0, cz = 0.0;
    for (int i = tid; i < nt; i += TH) {
      double x, y, z;
      load_point(tgt, dim, i, x, y, z);
      amax = fmaxf(amax, fmaxf(fabsf((float)x), fmaxf(fabsf((float)y), fabsf((float)z))));
      cx += x; cy += y; cz += z;
    }
    amax = warp_maxf(amax);
    cx = warp_sum(cx); cy = warp_sum(cy); cz = warp_sum(cz);
    if (lane == 0) { red[0][warp][0] = (double)amax; red[0][warp][1] = cx; red[0][warp][2] = cy; red[0][warp][3] = cz; }
    __syncthreads();
    float tmax = 0.f;
    double ox = 0.0, oy = 0.0, oz = 0.0;
#pragma unroll
    for (int w = 0; w < W; ++w) {
      tmax = fmaxf(tmax, (float)red[0][w][0]);
      ox += red[0][w][1]; oy += red[0][w][2]; oz += red[0][w][3];
    }
    tmax *= (1.0f + 4.0f * kU);
    ox /= (double)nt; oy /= (double)nt; oz /= (double)nt;
    __syncthreads();

    const double epsilon = p.nearest_only ? 0.0 : p.epsilon[pair];
    const bool normalize = p.nearest_only ? false : (p.normalize[pair] != 0);
    const int nblk32 = (nt + 31) >> 5;
    const int group = (nblk32 + 1023) >> 10;      // 32-target blocks per bookkeeping block
    const int bk_targets = 32 * group;
    StopState st{0.0, 0, 0};
    int passes = 0;

    for (;;) {  // ------------------------------ one ICP pass ------------------------
      double acc[kRed3d];
#pragma unroll
      for (int k = 0; k < kSums3d; ++k) acc[k] = 0.0;
#pragma unroll
      for (int k = 0; k < 3; ++k) { acc[kSums3d + 2 * k] = kInf; acc[kSums3d + 2 * k + 1] = -kInf; }

      for (int cbase = 0; cbase < ns; cbase += TH * R) {
        float qx[R], qy[R], qz[R];
        unsigned best[R], second[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int j = cbase + tid + TH * r;
          qx[r] = qy[r] = qz[r] = 0.f;
          best[r] = second[r] = 0xffffffffu;
          if (j < ns) {
            double sx, sy, sz, x, y, z;
            load_point(src, dim, j, sx, sy, sz);
            T.apply(sx, sy, sz, x, y, z);
            qx[r] = (float)x; qy[r] = (float)y; qz[r] = (float)z;
          }
        }
        // ---- fp32 sweep, target streamed through the shared tile ----------------
        float bm[R];
#pragma unroll
        for (int r = 0; r < R; ++r) bm[r] = __int_as_float(0x7f800000);
        for (int tbase = 0; tbase < nt; tbase += kTile3d) {
          __syncthreads();
          for (int i = tid; i < kTile3d; i += TH) {
            float x = kBigCoord, y = kBigCoord, z = kBigCoord;
            if (tbase + i < nt) {
              double dx, dy, dz;
              load_point(tgt, dim, tbase + i, dx, dy, dz);
              x = (float)dx; y = (float)dy; z = (float)dz;
            }
            tile_x[i] = x; tile_y[i] = y; tile_z[i] = z;
          }
          __syncthreads();
          const int nb = (min(kTile3d, nt - tbase) + 31) >> 5;
          for (int b = 0; b < nb; ++b) {
            const int gb = (tbase >> 5) + b;
            const float4* vx = reinterpret_cast<const float4*>(tile_x + b * 32);
            const float4* vy = reinterpret_cast<const float4*>(tile_y + b * 32);
            const float4* vz = reinterpret_cast<const float4*>(tile_z + b * 32);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float4 X = vx[i], Y = vy[i], Z = vz[i];
#pragma unroll
              for (int r = 0; r < R; ++r) {
                const float d0 = dist2_3d(qx[r], qy[r], qz[r], X.x, Y.x, Z.x);
                const float d1 = dist2_3d(qx[r], qy[r], qz[r], X.y, Y.y, Z.y);
                const float d2 = dist2_3d(qx[r], qy[r], qz[r], X.z, Y.z, Z.z);
                const float d3 = dist2_3d(qx[r], qy[r], qz[r], X.w, Y.w, Z.w);
                bm[r] = fmin3(fmin3(bm[r], d0, d1), d2, d3);
              }
            }
            if ((gb + 1) % group == 0 || gb == nblk32 - 1) {
              const unsigned id = (unsigned)(gb / group);
#pragma unroll
              for (int r = 0; r < R; ++r) {
                const unsigned pk = (__float_as_uint(bm[r]) & 0xfffffc00u) | id;
                second[r] = min(second[r], max(best[r], pk));
                best[r] = min(best[r], pk);
                bm[r] = __int_as_float(0x7f800000);
              }
            }
          }
        }
        // ---- exact resolution ------------------------------------------------------
        int corr[R];
        unsigned ambiguous = 0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int j = cbase + tid + TH * r;
          corr[r] = -1;
          if (j < ns) {
            const int blk = (int)(best[r] & 1023u);
            const float bval = __uint_as_float(best[r] & 0xfffffc00u);
            const float sval = __uint_as_float(second[r] & 0xfffffc00u);
            const float a_bound = fmaxf(fabsf(qx[r]), fmaxf(fabsf(qy[r]), fabsf(qz[r]))) * (1.0f + 4.0f * kU) + tmax;
            const float thr = candidate_threshold(bval, a_bound, kTruncRel);
            if (!(sval > thr)) {
              ambiguous |= 1u << r;
            } else {
              double sx, sy, sz, x, y, z;
              load_point(src, dim, j, sx, sy, sz);
              T.apply(sx, sy, sz, x, y, z);
              double bd = kInf;
              int bi = -1;
              const int lo = blk * bk_targets, hi = min(nt, lo + bk_targets);
              for (int i = lo; i < hi; ++i) {
                double ax, ay, az;
                load_point(tgt, dim, i, ax, ay, az);
                if (dist2_3d(qx[r], qy[r], qz[r], (float)ax, (float)ay, (float)az) <= thr) {
                  const double ex = x - ax, ey = y - ay, ez = z - az;
                  const double d = ex * ex + ey * ey + ez * ez;
                  if (d < bd) { bd = d; bi = i; }
                }
              }
              if (bi < 0) ambiguous |= 1u << r;
              corr[r] = bi;
            }
          }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
          unsigned vote = __ballot_sync(0xffffffffu, (ambiguous >> r) & 1u);
          while (vote) {
            const int src_lane = __ffs(vote) - 1;
            vote &= vote - 1;
            const float fx = __shfl_sync(0xffffffffu, qx[r], src_lane);
            const float fy = __shfl_sync(0xffffffffu, qy[r], src_lane);
            const float fz = __shfl_sync(0xffffffffu, qz[r], src_lane);
            const unsigned bpk = __shfl_sync(0xffffffffu, best[r], src_lane);
            const float a_bound = fmaxf(fabsf(fx), fmaxf(fabsf(fy), fabsf(fz))) * (1.0f + 4.0f * kU) + tmax;
            const float thr = candidate_threshold(__uint_as_float(bpk & 0xfffffc00u), a_bound, kTruncRel);
            const int j = cbase + (tid - lane) + src_lane + TH * r;
            double sx, sy, sz, x, y, z;
            load_point(src, dim, j, sx, sy, sz);
            T.apply(sx, sy, sz, x, y, z);
            double bd = kInf;
            int bi = 0x7fffffff;
            for (int i = lane; i < nt; i += 32) {
              double ax, ay, az;
              load_point(tgt, dim, i, ax, ay, az);
              if (!(dist2_3d(fx, fy, fz, (float)ax, (float)ay, (float)az) > thr)) {
                const double ex = x - ax, ey = y - ay, ez = z - az;
                const double d = ex * ex + ey * ey + ez * ez;
                if (d < bd || bi == 0x7fffffff) { bd = d; bi = i; }
              }
            }
            warp_argmin(bd, bi);
            if (lane == src_lane) corr[r] = (bi == 0x7fffffff) ? 0 : bi;
          }
        }
        // ---- float64 sums ----------------------------------------------------------
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int j = cbase + tid + TH * r;
          if (j < ns) {
            double sx, sy, sz, x, y, z, ax, ay, az;
            load_point(src, dim, j, sx, sy, sz);
            T.apply(sx, sy, sz, x, y, z);
            load_point(tgt, dim, corr[r], ax, ay, az);
            const double ux = x - ox, uy = y - oy, uz = z - oz;
            const double vx = ax - ox, vy = ay - oy, vz = az - oz;
            acc[0] += ux; acc[1] += uy; acc[2] += uz;
            acc[3] += vx; acc[4] += vy; acc[5] += vz;
            acc[6] = fma(ux, vx, acc[6]);   acc[7] = fma(ux, vy, acc[7]);   acc[8] = fma(ux, vz, acc[8]);
            acc[9] = fma(uy, vx, acc[9]);   acc[10] = fma(uy, vy, acc[10]); acc[11] = fma(uy, vz, acc[11]);
            acc[12] = fma(uz, vx, acc[12]); acc[13] = fma(uz, vy, acc[13]); acc[14] = fma(uz, vz, acc[14]);
            const double ex = x - ax, ey = y - ay, ez = z - az;
            const double d = ex * ex + ey * ey + ez * ez;
            acc[15] += d;
            if (normalize) {
              acc[16] = fmin(acc[16], fmin(x, ax)); acc[17] = fmax(acc[17], fmax(x, ax));
              acc[18] = fmin(acc[18], fmin(y, ay)); acc[19] = fmax(acc[19], fmax(y, ay));
              acc[20] = fmin(acc[20], fmin(z, az)); acc[21] = fmax(acc[21], fmax(z, az));
            }
            if (p.corr_out) p.corr_out[(size_t)pair * p.n_max + j] = corr[r];
            if (p.nearest_only && p.dist2_out) p.dist2_out[(size_t)pair * p.n_max + j] = d;
          }
        }
      }  // source chunks

#pragma unroll
      for (int k = 0; k < kSums3d; ++k) acc[k] = warp_sum(acc[k]);
      if (normalize) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          acc[kSums3d + 2 * k] = warp_min(acc[kSums3d + 2 * k]);
          acc[kSums3d + 2 * k + 1] = warp_max(acc[kSums3d + 2 * k + 1]);
        }
      }
      double(*slot)[kRed3d] = red[passes & 1];
      if (lane == 0) {
#pragma unroll
        for (int k = 0; k < kRed3d; ++k) slot[warp][k] = acc[k];
      }
      __syncthreads();
      double tot[kRed3d];
#pragma unroll
      for (int k = 0; k < kRed3d; ++k) tot[k] = slot[0][k];
      for (int w = 1; w < W; ++w) {
#pragma unroll
        for (int k = 0; k < kSums3d; ++k) tot[k] += slot[w][k];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          tot[kSums3d + 2 * k] = fmin(tot[kSums3d + 2 * k], slot[w][kSums3d + 2 * k]);
          tot[kSums3d + 2 * k + 1] = fmax(tot[kSums3d + 2 * k + 1], slot[w][kSums3d + 2 * k + 1]);
        }
      }
      passes += 1;

      double err = tot[15];
      if (normalize) {
        const double wx = tot[17] - tot[16], wy = tot[19] - tot[18], wz = tot[21] - tot[20];
        const double diag = sqrt(wx * wx + wy * wy + wz * wz);
        err = err / (diag * diag * (double)ns);
      }
      int status;
      if (p.nearest_only) {
        status = B200ICP_STATUS_EPSILON;
      } else {
        const double nn = (double)ns, inv_n = 1.0 / nn;
        const double mq[3] = {tot[0] * inv_n, tot[1] * inv_n, tot[2] * inv_n};
        const double mp[3] = {tot[3] * inv_n, tot[4] * inv_n, tot[5] * inv_n};
        double C[9], Rm[9];
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
          for (int b = 0; b < 3; ++b) C[a * 3 + b] = tot[6 + a * 3 + b] - nn * mq[a] * mp[b];
        kabsch_rotation_3x3(C, Rm);
        const double qb[3] = {ox + mq[0], oy + mq[1], oz + mq[2]};
        const double pb[3] = {ox + mp[0], oy + mp[1], oz + mp[2]};
        double dt[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) dt[a] = pb[a] - (Rm[a * 3] * qb[0] + Rm[a * 3 + 1] * qb[1] + Rm[a * 3 + 2] * qb[2]);
        // T <- [Rm dt; 0 1] * T   (bottom row of T may be anything the caller passed)
        double n[12];
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
          for (int c = 0; c < 4; ++c)
            n[a * 4 + c] = Rm[a * 3] * T.m[c] + Rm[a * 3 + 1] * T.m[4 + c] + Rm[a * 3 + 2] * T.m[8 + c] + dt[a] * bottom[c];
#pragma unroll
        for (int k = 0; k < 12; ++k) T.m[k] = n[k];
        status = stop_test(st, err, epsilon, p.max_iters, p.stopping_thresh);
      }
      if (status >= 0) {
        if (tid == 0) {
          if (p.T_out) {
            double* To = p.T_out + (size_t)pair * 16;
#pragma unroll
            for (int k = 0; k < 12; ++k) To[k] = T.m[k];
#pragma unroll
            for (int k = 0; k < 4; ++k) To[12 + k] = bottom[k];
          }
          if (p.err_out) p.err_out[pair] = err;
          if (p.iters_out) p.iters_out[pair] = passes;
          if (p.status_out) p.status_out[pair] = status;
        }
        if (p.corr_out)
          for (int j = ns + tid; j < p.n_max; j += TH) p.corr_out[(size_t)pair * p.n_max + j] = -1;
        break;
      }
    }
  }
}

}  // namespace b200icp
