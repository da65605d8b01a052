// Fused general (3-D, or planar scans too large for the 2-D kernel) ICP.  One thread-block
// CLUSTER aligns one pair: the source cloud is cut into one slice per CTA of the cluster, every
// CTA sweeps its slice against the whole target (streamed through a shared-memory fp32 tile,
// packed fp32x2 math), the per-pass sums are reduced CTA -> cluster through distributed shared
// memory behind one hardware cluster barrier, and every CTA solves the same 3x3 Kabsch problem
// redundantly so the pose never leaves registers.  The cluster size is chosen at launch from
// the batch shape (few big clouds -> 16 CTAs per pair, many small ones -> 1).  Same exactness
// scheme as icp2d.cuh; the solve is kabsch3.cuh.
//
// Replaces modules/icp.py:123-189 for the call shape of
// code/icp_warm_up/test_icp.py:78-88 (3-D clouds, 24 yaw initialisations).
#pragma once
#include <cooperative_groups.h>

#include "icp_common.cuh"
#include "kabsch3.cuh"

namespace b200icp {

constexpr int kTile3d = 1536;   // targets per shared-memory tile (fp32 [x4 y4 z4] groups: 18 KB), two tiles in flight
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

// One 32-target block of the sweep for R source points per thread.  The tile holds groups of
// four targets as [x0..x3 y0..y3 z0..z3]; every LDS.128 is a warp-wide broadcast.  Per two
// candidates: 3 FADD2 + FMUL2 + 2 FFMA2, bit-identical to dist2_3d() (same operation order).
template <int R>
__device__ __forceinline__ void sweep3d_block(const float* tile_blk, const f32x2 (&qxx)[R],
                                              const f32x2 (&qyy)[R], const f32x2 (&qzz)[R],
                                              float (&bm)[R]) {
  const ulonglong2* tp = reinterpret_cast<const ulonglong2*>(tile_blk);
#pragma unroll
  for (int g = 0; g < 8; ++g) {
    const ulonglong2 X = tp[3 * g], Y = tp[3 * g + 1], Z = tp[3 * g + 2];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      f32x2 dx = sub2(qxx[r], X.x), dy = sub2(qyy[r], Y.x), dz = sub2(qzz[r], Z.x);
      const f32x2 d01 = fma2(dx, dx, fma2(dy, dy, mul2(dz, dz)));
      dx = sub2(qxx[r], X.y); dy = sub2(qyy[r], Y.y); dz = sub2(qzz[r], Z.y);
      const f32x2 d23 = fma2(dx, dx, fma2(dy, dy, mul2(dz, dz)));
      float d0, d1, d2, d3;
      unpack2(d01, d0, d1);
      unpack2(d23, d2, d3);
      bm[r] = fmin3(fmin3(bm[r], d0, d1), d2, d3);
    }
  }
}

// The sweep of one chunk (NR rows of TH source points) against the whole target.  The target's
// fp32 copy (built once per pair in the workspace, already in tile layout) is streamed through
// two shared-memory tiles with TMA bulk copies: thread 0 issues tile k+1 while the CTA sweeps
// tile k; one mbarrier per tile signals arrival, one __syncthreads per tile frees it.
struct TilePipe {
  float* tile;          // [2][kTile3d * 3]
  uint64_t* full;       // [2] mbarriers
  uint32_t phase[2];
};

template <int NR, int R, int TH>
__device__ __forceinline__ void sweep3d_chunk(TilePipe& tp, const float* tgt32, int nt, int group,
                                              int nblk32, const float (&qx)[R], const float (&qy)[R],
                                              const float (&qz)[R], unsigned (&best)[R],
                                              unsigned (&second)[R]) {
  const int tid = threadIdx.x;
  f32x2 qxx[NR], qyy[NR], qzz[NR];
  float bm[NR];
#pragma unroll
  for (int r = 0; r < NR; ++r) {
    qxx[r] = pack2(qx[r], qx[r]); qyy[r] = pack2(qy[r], qy[r]); qzz[r] = pack2(qz[r], qz[r]);
    bm[r] = __int_as_float(0x7f800000);
  }
  const int n_tiles = (nt + kTile3d - 1) / kTile3d;
  auto issue = [&](int k) {
    const int first = k * kTile3d;
    const int blocks = (min(kTile3d, nt - first) + 31) >> 5;          // whole 32-target blocks (the copy is padded)
    const uint32_t bytes = (uint32_t)blocks * 32u * 12u;
    uint64_t* bar = tp.full + (k & 1);
    mbar_expect_tx(bar, bytes);
    tma_load_1d(tp.tile + (size_t)(k & 1) * kTile3d * 3, tgt32 + (size_t)first * 3, bytes, bar);
  };
  if (tid == 0) issue(0);
  for (int k = 0; k < n_tiles; ++k) {
    if (tid == 0 && k + 1 < n_tiles) issue(k + 1);                     // its buffer was freed by the barrier below
    mbar_wait(tp.full + (k & 1), tp.phase[k & 1]);
    tp.phase[k & 1] ^= 1u;
    const float* tile = tp.tile + (size_t)(k & 1) * kTile3d * 3;
    const int tbase = k * kTile3d;
    const int nb = (min(kTile3d, nt - tbase) + 31) >> 5;
    for (int b = 0; b < nb; ++b) {
      const int gb = (tbase >> 5) + b;
      sweep3d_block<NR>(tile + b * 96, qxx, qyy, qzz, bm);
      if ((gb + 1) % group == 0 || gb == nblk32 - 1) {
        const unsigned id = (unsigned)(gb / group);
#pragma unroll
        for (int r = 0; r < NR; ++r) {
          const unsigned pk = (__float_as_uint(bm[r]) & 0xfffffc00u) | id;
          second[r] = min(second[r], max(best[r], pk));
          best[r] = min(best[r], pk);
          bm[r] = __int_as_float(0x7f800000);
        }
      }
    }
    __syncthreads();
  }
}

template <int R, int TH>
__global__ void __launch_bounds__(TH, 2) icp3d_kernel(const RunParams p, const int dim) {
  __shared__ __align__(128) float tiles[2 * kTile3d * 3];  // two tiles of four-target groups: x4 y4 z4
  __shared__ __align__(8) uint64_t full_bar[2];
  __shared__ double wpart[TH / 32][kRed3d];             // per-warp partial sums of this CTA
  __shared__ double part[2][kRed3d];                    // this CTA's sums, by pass parity (read by the cluster)
  __shared__ double s_tot[kRed3d];                      // cluster totals
  __shared__ int s_pair;
  constexpr int W = TH / 32;
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  const int G = (int)cluster.num_blocks();              // CTAs per pair
  const int rank = (int)cluster.block_rank();
  TilePipe tp;
  tp.tile = tiles; tp.full = full_bar; tp.phase[0] = tp.phase[1] = 0u;
  if (threadIdx.x == 0) {
    mbar_init(&full_bar[0], 1);
    mbar_init(&full_bar[1], 1);
    fence_mbar_init();
  }
  __syncthreads();
  // this cluster's fp32 copy of the current target, in tile layout, padded to whole 32-target blocks
  float* tgt32 = p.tgt32 + (size_t)(blockIdx.x / G) * p.tgt32_stride;
  const double kInf = __longlong_as_double(0x7ff0000000000000LL);
  constexpr float kTruncRel = 2.44140625e-4f;  // 2^-12 >= the 10 truncated mantissa bits (2^-13)
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  for (;;) {
    if (rank == 0 && tid == 0) s_pair = (int)atomicAdd(p.queue, 1u);
    cluster.sync();
    const int pair = *cluster.map_shared_rank(&s_pair, 0);
    cluster.sync();   // every CTA has read rank 0's s_pair: rank 0 may overwrite it, or exit
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
      if (tid == 0 && rank == 0) {
        if (p.status_out) p.status_out[pair] = early;
        if (p.iters_out) p.iters_out[pair] = 0;
        if (p.err_out) p.err_out[pair] = __longlong_as_double(0x7ff8000000000000LL);
        if (p.T_out)
          for (int k = 0; k < 16; ++k) p.T_out[(size_t)pair * 16 + k] = p.init ? p.init[(size_t)pair * 16 + k] : 0.0;
        if (p.init) {
          const double* T0 = p.init + (size_t)pair * 16;
          store_pose_to_peers(p, pair, T0[3], T0[7], T0[0], T0[4], __longlong_as_double(0x7ff8000000000000LL));
        }
      }
      if (p.corr_out)
        for (int j = rank * TH + tid; j < p.n_max; j += G * TH) p.corr_out[(size_t)pair * p.n_max + j] = -1;
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
    if (lane == 0) { wpart[warp][0] = (double)amax; wpart[warp][1] = cx; wpart[warp][2] = cy; wpart[warp][3] = cz; }
    __syncthreads();
    float tmax = 0.f;
    double ox = 0.0, oy = 0.0, oz = 0.0;
#pragma unroll
    for (int w = 0; w < W; ++w) {
      tmax = fmaxf(tmax, (float)wpart[w][0]);
      ox += wpart[w][1]; oy += wpart[w][2]; oz += wpart[w][3];
    }
    tmax *= (1.0f + 4.0f * kU);
    ox /= (double)nt; oy /= (double)nt; oz /= (double)nt;
    __syncthreads();
    // ---- fp32 copy of the target (once per pair; each CTA converts a share, TMA reads it) ----
    {
      const int np32 = ((nt + 31) >> 5) << 5;
      for (int i = rank * TH + tid; i < np32; i += G * TH) {
        float x = kBigCoord, y = kBigCoord, z = kBigCoord;
        if (i < nt) {
          double dx, dy, dz;
          load_point(tgt, dim, i, dx, dy, dz);
          x = (float)dx; y = (float)dy; z = (float)dz;
        }
        float* g = tgt32 + (size_t)(i >> 2) * 12 + (i & 3);
        g[0] = x; g[4] = y; g[8] = z;
      }
      __threadfence();
      fence_proxy_async_global();     // generic-proxy stores -> visible to the TMA (async proxy) reads
      cluster.sync();
    }

    const double epsilon = p.nearest_only ? 0.0 : p.epsilon[pair];
    const bool normalize = p.nearest_only ? false : (p.normalize[pair] != 0);
    const int nblk32 = (nt + 31) >> 5;
    const int group = (nblk32 + 1023) >> 10;      // 32-target blocks per bookkeeping block
    const int bk_targets = 32 * group;
    // source slice of this CTA: rows of TH points dealt out evenly over the cluster, processed in
    // chunks of nr <= R rows (rows spread evenly over the chunks so that no chunk is nearly empty)
    const int rows_total = (ns + TH - 1) / TH;
    const int row_lo = (int)((long long)rows_total * rank / G), row_hi = (int)((long long)rows_total * (rank + 1) / G);
    const int n_chunks = (row_hi - row_lo + R - 1) / R;
    StopState st{0.0, 0, 0};
    int passes = 0;

    for (;;) {  // ------------------------------ one ICP pass ------------------------
      double acc[kRed3d];
#pragma unroll
      for (int k = 0; k < kSums3d; ++k) acc[k] = 0.0;
#pragma unroll
      for (int k = 0; k < 3; ++k) { acc[kSums3d + 2 * k] = kInf; acc[kSums3d + 2 * k + 1] = -kInf; }

      for (int ch = 0; ch < n_chunks; ++ch) {
        const int rows_here = row_hi - row_lo;
        const int row0 = row_lo + (int)((long long)rows_here * ch / n_chunks);
        const int nr = row_lo + (int)((long long)rows_here * (ch + 1) / n_chunks) - row0;
        const int cbase = row0 * TH;
        float qx[R], qy[R], qz[R];
        unsigned best[R], second[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int j = cbase + tid + TH * r;
          qx[r] = qy[r] = qz[r] = 0.f;
          best[r] = second[r] = 0xffffffffu;
          if (r < nr && j < ns) {
            double sx, sy, sz, x, y, z;
            load_point(src, dim, j, sx, sy, sz);
            T.apply(sx, sy, sz, x, y, z);
            qx[r] = (float)x; qy[r] = (float)y; qz[r] = (float)z;
          }
        }
        // ---- fp32 sweep, target streamed through the shared tile ----------------
        switch (nr) {
          case 1: sweep3d_chunk<1, R, TH>(tp, tgt32, nt, group, nblk32, qx, qy, qz, best, second); break;
          case 2: if constexpr (R >= 2) sweep3d_chunk<2, R, TH>(tp, tgt32, nt, group, nblk32, qx, qy, qz, best, second); break;
          case 3: if constexpr (R >= 3) sweep3d_chunk<3, R, TH>(tp, tgt32, nt, group, nblk32, qx, qy, qz, best, second); break;
          case 4: if constexpr (R >= 4) sweep3d_chunk<4, R, TH>(tp, tgt32, nt, group, nblk32, qx, qy, qz, best, second); break;
          default: break;
        }
        // ---- exact resolution ------------------------------------------------------
        // Rescan of the best bookkeeping block: fp32 filter on the fp32 copy (L2), float64 on
        // the few candidates that pass.  The R rows of a thread are walked together, group by
        // group, so that their 3*R independent 16-byte loads are in flight at once.
        int corr[R];
        unsigned ambiguous = 0, live = 0;
        float thr[R];
        int blo[R];
        double px[R], py[R], pz[R], bd[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int j = cbase + tid + TH * r;
          corr[r] = -1; thr[r] = 0.f; blo[r] = 0; bd[r] = kInf;
          px[r] = py[r] = pz[r] = 0.0;
          if (r < nr && j < ns) {
            const int blk = (int)(best[r] & 1023u);
            const float bval = __uint_as_float(best[r] & 0xfffffc00u);
            const float sval = __uint_as_float(second[r] & 0xfffffc00u);
            const float a_bound = fmaxf(fabsf(qx[r]), fmaxf(fabsf(qy[r]), fabsf(qz[r]))) * (1.0f + 4.0f * kU) + tmax;
            thr[r] = candidate_threshold(bval, a_bound, kTruncRel);
            if (!(sval > thr[r])) {
              ambiguous |= 1u << r;
            } else {
              live |= 1u << r;
              blo[r] = blk * bk_targets;
              double sx, sy, sz;
              load_point(src, dim, j, sx, sy, sz);
              T.apply(sx, sy, sz, px[r], py[r], pz[r]);
            }
          }
        }
        for (int off = 0; off < bk_targets; off += 4) {      // bk_targets is a multiple of 32
          float4 X[R], Y[R], Z[R];
#pragma unroll
          for (int r = 0; r < R; ++r) {
            // rows that are not live read group 0 of the copy (always present) and ignore it
            const int i0 = ((live >> r) & 1u) ? min(blo[r] + off, ((nt - 1) >> 2) << 2) : 0;
            const float4* g4 = reinterpret_cast<const float4*>(tgt32 + (size_t)(i0 >> 2) * 12);
            X[r] = __ldcg(g4); Y[r] = __ldcg(g4 + 1); Z[r] = __ldcg(g4 + 2);
          }
#pragma unroll
          for (int r = 0; r < R; ++r) {
            if (!((live >> r) & 1u)) continue;
            const int i0 = blo[r] + off;
            if (i0 >= nt) continue;
            const float fxs[4] = {X[r].x, X[r].y, X[r].z, X[r].w};
            const float fys[4] = {Y[r].x, Y[r].y, Y[r].z, Y[r].w};
            const float fzs[4] = {Z[r].x, Z[r].y, Z[r].z, Z[r].w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int i = i0 + u;
              if (i < nt && dist2_3d(qx[r], qy[r], qz[r], fxs[u], fys[u], fzs[u]) <= thr[r]) {
                double ax, ay, az;
                load_point(tgt, dim, i, ax, ay, az);
                const double ex = px[r] - ax, ey = py[r] - ay, ez = pz[r] - az;
                const double d = ex * ex + ey * ey + ez * ez;
                if (d < bd[r]) { bd[r] = d; corr[r] = i; }
              }
            }
          }
        }
#pragma unroll
        for (int r = 0; r < R; ++r)
          if (((live >> r) & 1u) && corr[r] < 0) ambiguous |= 1u << r;
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
          if (r < nr && j < ns) {
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
      // CTA: warp partials -> part[parity]; cluster: every CTA adds the G parts in rank order
      // (distributed shared memory), so all CTAs hold bit-identical totals and take the same
      // decisions.  Entries 16/18/20 are minima, 17/19/21 maxima, the rest sums.
      if (lane == 0) {
#pragma unroll
        for (int k = 0; k < kRed3d; ++k) wpart[warp][k] = acc[k];
      }
      __syncthreads();
      const bool is_min = lane >= kSums3d && ((lane - kSums3d) & 1) == 0;
      const bool is_max = lane >= kSums3d && ((lane - kSums3d) & 1) == 1;
      if (warp == 0 && lane < kRed3d) {
        double v = wpart[0][lane];
        for (int w = 1; w < W; ++w) {
          const double o = wpart[w][lane];
          v = is_min ? fmin(v, o) : (is_max ? fmax(v, o) : v + o);
        }
        part[passes & 1][lane] = v;
      }
      cluster.sync();
      if (warp == 0 && lane < kRed3d) {
        double v = *cluster.map_shared_rank(&part[passes & 1][lane], 0);
        for (int g = 1; g < G; ++g) {
          const double o = *cluster.map_shared_rank(&part[passes & 1][lane], g);
          v = is_min ? fmin(v, o) : (is_max ? fmax(v, o) : v + o);
        }
        s_tot[lane] = v;
      }
      __syncthreads();
      double tot[kRed3d];
#pragma unroll
      for (int k = 0; k < kRed3d; ++k) tot[k] = s_tot[k];
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
        if (tid == 0 && rank == 0) {
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
          store_pose_to_peers(p, pair, T.m[3], T.m[7], T.m[0], T.m[4], err);
        }
        if (p.corr_out && rank == 0)
          for (int j = ns + tid; j < p.n_max; j += TH) p.corr_out[(size_t)pair * p.n_max + j] = -1;
        break;
      }
    }
  }
}

}  // namespace b200icp
