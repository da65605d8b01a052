// Fused planar ICP, row engine (the shipped 2-D kernel).  One CTA aligns one (source, target) scan
// pair from staging to the final pose; the iterate-until-converged loop stays on the device.
//
// Replaces modules/icp.py:123-189 (run_icp) for 2-D scans, i.e. the call shapes of
// modules/localization.py:120-123 and main.py:162; in RANGES mode it also replaces
// get_lidar_data (modules/localization.py:252-307) for the two scans of the pair, so that the
// caller ships 8 (or 4) bytes per beam instead of 16 per point.
//
// What changed against icp2d.cuh (round 1, kept as the exhaustive register-tiled variant): the
// profile of that kernel showed 23 % of its issue slots in the candidate arithmetic and 77 % in
// per-row fixed work (address arithmetic, 64-bit mask loops, spills, predicates for the ragged
// tail, reductions).  Here a warp walks ROWS of 32*R consecutive source points (R queries per lane,
// 32 apart; R = 1 ships: R = 2..4 pay every warp-level step -- row box, block bounds, seeds,
// ballots, the shared-memory loads of a target block -- once per 64..128 queries but their wider
// row boxes admit more blocks, measured slower; api.cu); ragged tails are clamped
// duplicates instead of predicates; target blocks beyond the scan are +inf boxes instead of
// tests; matches live in a 2-byte shared array instead of a register queue; the bounding box of
// the normalised error is a separate light pass that only normalised pairs execute.
//
// Per pair:
//   stage   TMA bulk copies (cp.async.bulk + mbarrier) of the float64 scans into shared memory
//           (RANGES: coalesced L2 loads of the two range rows, gate, polar -> xy with the
//           caller's cos/sin tables in the reference's operation order, ballot compaction);
//           fp32 copy of the target in [x4 y4] groups, bounding boxes of its 32-point blocks.
//   pass k  per row: q = T_k s (float64, rounded once to fp32 for the search); row box
//           (CREDUX.MIN/MAX); lower bound of every target block against the row box with the
//           candidate's own monotone arithmetic (exact, see icp2d.cuh); blocks that overlap the
//           row are swept first, then every block whose bound is within the largest candidate
//           threshold of the row; float64 resolution of the candidates of the best block (or,
//           when a second block is within the slack, of all targets); float64 sums.
//           Then warp shuffles, ONE shared-memory exchange and ONE __syncthreads per pass,
//           closed-form planar Kabsch, pose composition, the reference's three stop tests.
#pragma once
#include "icp2d.cuh"

namespace b200icp {

struct Rows2dLayout {
  int np;   // padded point capacity (multiple of 64)
  size_t off_t64, off_s64, off_t32, off_bb, off_red, off_corr, off_rows, off_ctl, total;
};
__host__ __device__ inline Rows2dLayout rows2d_layout(int n_max, int warps, bool ranges, bool src_global = false) {
  Rows2dLayout L;
  L.np = ((n_max + 63) / 64) * 64;
  size_t o = 0;
  L.off_t64 = o; o += (size_t)L.np * 16;
  L.off_s64 = o; o += src_global ? 0 : (size_t)L.np * 16;    // SRCG: the source stays in the pool (L1/L2)
  L.off_t32 = o; o += (size_t)L.np * 8;
  L.off_bb = o; o += (size_t)64 * 16;                        // target-block boxes (xmin, xmax, ymin, ymax)
  L.off_red = o; o += (size_t)2 * warps * kRed2d * 8;        // per-pass exchange, double-buffered
  L.off_corr = o; o += (size_t)L.np * 2;                     // matches of the current pass (uint16)
  L.off_rows = o; o += ranges ? (size_t)2 * 128 * 4 : 0;     // RANGES: valid beams per 32-beam row, row offsets
  o = (o + 15) / 16 * 16;
  L.off_ctl = o; o += 64;
  L.total = o;
  return L;
}

// One 32-target block against the R queries of every lane.  The eight group-of-four minima of a
// query's new best block are kept (gbest) so that the resolution can tell which groups hold
// candidates without evaluating the block a second time.
// FIRST: the first block a row sweeps -- it becomes the best block unconditionally (best = second = "none"
// before it), so the words and the group minima are written instead of compared and selected.
template <int R, bool FIRST = false>
__device__ __forceinline__ void sweep_block_rows(const ulonglong2* tp, int b, const f32x2 (&qxx)[R],
                                                 const f32x2 (&qyy)[R], unsigned (&best)[R],
                                                 unsigned (&second)[R], float (&gbest)[R][8]) {
  float gm[R][8];
  const ulonglong2* tb = tp + b * (kBlockTargets / 2);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const ulonglong2 X = tb[2 * i];
    const ulonglong2 Y = tb[2 * i + 1];
#pragma unroll
    for (int q = 0; q < R; ++q) {
      f32x2 dx = sub2(qxx[q], X.x), dy = sub2(qyy[q], Y.x);
      const f32x2 d01 = fma2(dx, dx, mul2(dy, dy));
      dx = sub2(qxx[q], X.y); dy = sub2(qyy[q], Y.y);
      const f32x2 d23 = fma2(dx, dx, mul2(dy, dy));
      float d0, d1, d2, d3;
      unpack2(d01, d0, d1);
      unpack2(d23, d2, d3);
      gm[q][i] = fminf(fmin3(d0, d1, d2), d3);
    }
  }
#pragma unroll
  for (int q = 0; q < R; ++q) {
    const float bm = fminf(fmin3(fmin3(fmin3(gm[q][0], gm[q][1], gm[q][2]), gm[q][3], gm[q][4]), gm[q][5], gm[q][6]),
                           gm[q][7]);
    const unsigned pk = (__float_as_uint(bm) & 0xffffffc0u) | (unsigned)b;
    if constexpr (FIRST) {
#pragma unroll
      for (int i = 0; i < 8; ++i) gbest[q][i] = gm[q][i];
      second[q] = 0xffffffffu;
      best[q] = pk;
    } else {
      if (pk < best[q]) {
#pragma unroll
        for (int i = 0; i < 8; ++i) gbest[q][i] = gm[q][i];
      }
      second[q] = min(second[q], max(best[q], pk));
      best[q] = min(best[q], pk);
    }
  }
}

__device__ __forceinline__ double load_range(const RunParams& p, int scan, int beam) {
  const size_t at = (size_t)scan * p.n_beams + beam;
  // L2 loads (ld.global.cg): a row that arrives while the kernel runs (gated launch) must never be
  // served from an L1 line fetched before it was resident
  return p.ranges_f32 ? (double)__ldcg(reinterpret_cast<const float*>(p.ranges) + at)
                      : __ldcg(reinterpret_cast<const double*>(p.ranges) + at);
}

// W warps per CTA, R queries per lane; PRUNE = false sweeps every block (same bits; A/B and tests);
// RANGES: the pool is [n_scans, n_beams] ranges, the scans are built here.
// SRCG (points in only): the float64 source is read from the pool in global memory (each pass reads
// every source point once, coalesced, L1-resident after the first pass) instead of a shared copy:
// 30 KB instead of 47 KB per CTA at 1081 points, so that a fifth CTA fits per SM.
template <int W, int R, int MINB, bool PRUNE, bool RANGES, bool SRCG = false>
__global__ void __launch_bounds__(W * 32, MINB) icp2d_rows_kernel(const RunParams p) {
  static_assert(!(RANGES && SRCG), "ranges-in builds the source in shared memory");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const Rows2dLayout L = rows2d_layout(p.n_max, W, RANGES, SRCG);
  double2* t64 = reinterpret_cast<double2*>(smem_raw + L.off_t64);
  double2* s64 = reinterpret_cast<double2*>(smem_raw + L.off_s64);       // unused with SRCG
  const double2* src_pts = s64;
  float* t32 = reinterpret_cast<float*>(smem_raw + L.off_t32);
  float4* bb = reinterpret_cast<float4*>(smem_raw + L.off_bb);
  double* red = reinterpret_cast<double*>(smem_raw + L.off_red);
  unsigned short* corr_s = reinterpret_cast<unsigned short*>(smem_raw + L.off_corr);
  int* s_rows = reinterpret_cast<int*>(smem_raw + L.off_rows);
  uint64_t* mbar = reinterpret_cast<uint64_t*>(smem_raw + L.off_ctl);
  int* s_ctl = reinterpret_cast<int*>(smem_raw + L.off_ctl + 16);

  constexpr int kRowQ = 32 * R;
  const int tid = threadIdx.x;
  const int lane = tid & 31;
  const int warp = tid >> 5;
  constexpr float kTruncRel = 1.52587890625e-5f;  // 2^-16 >= the 6 truncated mantissa bits (2^-17)
  const float kInfF = __int_as_float(0x7f800000);
  const double kInf = __longlong_as_double(0x7ff0000000000000LL);
  const double kNaN = __longlong_as_double(0x7ff8000000000000LL);
  const unsigned full = 0xffffffffu;

  if (tid == 0) {
    mbar_init(mbar, 1);
    fence_mbar_init();
  }
  uint32_t parity = 0;
  unsigned retired = 0;   // pairs this thread (tid 0) has retired, announced to the peers when the CTA leaves the queue

  for (;;) {
    __syncthreads();  // previous pair fully consumed; mbarrier initialised
    if (tid == 0) {
      const int pr = (int)atomicAdd(p.queue, 1u);
      int resident = 1;
      if (p.ready_scans && pr < p.n_pairs) {
        // Gated launch: wait until the copy stream has delivered both scans (range rows) of this
        // pair.  Bounded: after gate_timeout_ns the pair is given up (status NOT_RESIDENT, never
        // staged), queue[5] tells the host side, and from then on nobody waits.
        const int a = p.pairs[2 * pr], b = p.pairs[2 * pr + 1];
        if (a >= 0 && b >= 0 && a < p.n_scans && b < p.n_scans) {
          const unsigned need = (unsigned)max(a, b) + 1u;
          if (ld_acquire_sys(p.ready_scans) < need) {
            volatile unsigned* gave_up = p.queue + 5;
            unsigned long long t0, t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
            while (ld_acquire_sys(p.ready_scans) < need && *gave_up == 0u) {
              __nanosleep(500);
              asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
              if (t1 - t0 > p.gate_timeout_ns) *gave_up = 1u;
            }
            resident = ld_acquire_sys(p.ready_scans) >= need ? 1 : 0;
          }
          fence_proxy_async_global();
        }
      }
      s_ctl[0] = pr;
      s_ctl[1] = resident;
    }
    __syncthreads();
    const int pair = s_ctl[0];
    if (pair >= p.n_pairs) {
      if (tid == 0) pairs_retired(p, p.queue + 6, retired);    // this CTA's pairs, announced once
      break;
    }

    const int si = p.pairs[2 * pair], ti = p.pairs[2 * pair + 1];
    int early = -1;
    int ns = 0, nt = 0;
    if (si < 0 || ti < 0 || si >= p.n_scans || ti >= p.n_scans) {
      early = B200ICP_STATUS_SKIPPED;
    } else if (!s_ctl[1]) {
      early = B200ICP_STATUS_NOT_RESIDENT;
    } else if (!RANGES) {
      ns = min(p.counts[si], p.n_max);
      nt = min(p.counts[ti], p.n_max);
      if (ns <= 0 || nt <= 0) early = B200ICP_STATUS_EMPTY;
    }
    // ---- initial transform (planar part) ------------------------------------
    double r00 = 1, r01 = 0, r10 = 0, r11 = 1, tx = 0, ty = 0, zs = 1;
    if (early < 0 && !p.nearest_only) {
      const double* T0 = p.init + (size_t)pair * 16;
      r00 = T0[0]; r01 = T0[1]; tx = T0[3];
      r10 = T0[4]; r11 = T0[5]; ty = T0[7];
      zs = T0[10];
      const bool planar = T0[2] == 0.0 && T0[6] == 0.0 && T0[8] == 0.0 && T0[9] == 0.0 &&
                          T0[11] == 0.0 && T0[12] == 0.0 && T0[13] == 0.0 && T0[14] == 0.0 &&
                          T0[15] == 1.0;
      if (!planar) early = B200ICP_STATUS_NONPLANAR;
    }

    if (early < 0) {
      if constexpr (RANGES) {
        // ---- build both scans from their range rows (get_lidar_data, localization.py:288-301) ---
        // x = fl(fl(r cos a) + off_x), y = fl(fl(r sin a) + off_y), beams outside [rmin, rmax] (and
        // NaN) dropped, beam order kept.  Rows of 32 beams: [0, nrows) source, [nrows, 2 nrows) target.
        const int nrows = (p.n_beams + 31) >> 5;
        // kStageRows rows per step: their loads are issued together (one L2 latency per step instead of
        // one per row; a ballot between dependent loads kept them serial: 17 rows x ~0.4 us per pair)
        constexpr int kStageRows = 6;
        const double kNoRange = 0.0;               // placeholder of a lane without a beam (have[k] decides)
        for (int row0 = warp; row0 < 2 * nrows; row0 += W * kStageRows) {
          double rr[kStageRows];
          bool have[kStageRows];
#pragma unroll
          for (int k = 0; k < kStageRows; ++k) {
            const int row = row0 + k * W;
            const int b = (row < nrows ? row : row - nrows) * 32 + lane;
            have[k] = row < 2 * nrows && b < p.n_beams;
            rr[k] = have[k] ? load_range(p, row < nrows ? si : ti, b) : kNoRange;
          }
#pragma unroll
          for (int k = 0; k < kStageRows; ++k) {
            const int row = row0 + k * W;
            if (row < 2 * nrows) {               // warp-uniform
              const bool valid = have[k] && (rr[k] >= p.range_min) && (rr[k] <= p.range_max);
              const unsigned m = __ballot_sync(full, valid);
              if (lane == 0) s_rows[row] = __popc(m);
            }
          }
        }
        __syncthreads();
        if (warp < 2) {   // warp 0: source offsets, warp 1 (or warp 0 again when W == 1): target offsets
          for (int half = warp; half < 2; half += (W >= 2 ? 2 : 1)) {
            const int* cnt = s_rows + half * nrows;
            int* off = s_rows + 128 + half * nrows;
            const int c0 = 2 * lane < nrows ? cnt[2 * lane] : 0;
            const int c1 = 2 * lane + 1 < nrows ? cnt[2 * lane + 1] : 0;
            int incl = c0 + c1;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
              const int up = __shfl_up_sync(full, incl, o);
              if (lane >= o) incl += up;
            }
            const int excl = incl - (c0 + c1);
            if (2 * lane < nrows) off[2 * lane] = excl;
            if (2 * lane + 1 < nrows) off[2 * lane + 1] = excl + c0;
            if (lane == 31) s_ctl[2 + half] = incl;
          }
        }
        __syncthreads();
        ns = min(s_ctl[2], p.n_max);
        nt = min(s_ctl[3], p.n_max);
        for (int row0 = warp; row0 < 2 * nrows; row0 += W * kStageRows) {
          double rr[kStageRows];
          bool have[kStageRows];
#pragma unroll
          for (int k = 0; k < kStageRows; ++k) {
            const int row = row0 + k * W;
            const int b = (row < nrows ? row : row - nrows) * 32 + lane;
            have[k] = row < 2 * nrows && b < p.n_beams;
            rr[k] = have[k] ? load_range(p, row < nrows ? si : ti, b) : kNoRange;
          }
#pragma unroll
          for (int k = 0; k < kStageRows; ++k) {
            const int row = row0 + k * W;
            if (row < 2 * nrows) {               // warp-uniform
              const bool is_src = row < nrows;
              const int b = (is_src ? row : row - nrows) * 32 + lane;
              const double r = rr[k];
              const bool valid = have[k] && (r >= p.range_min) && (r <= p.range_max);
              const unsigned m = __ballot_sync(full, valid);
              if (valid) {
                const int pos = s_rows[128 + row] + __popc(m & ((1u << lane) - 1u));
                if (pos < p.n_max) {
                  double2 pt;
                  pt.x = __dadd_rn(__dmul_rn(r, __ldg(p.cos_tab + b)), p.off_x);
                  pt.y = __dadd_rn(__dmul_rn(r, __ldg(p.sin_tab + b)), p.off_y);
                  (is_src ? s64 : t64)[pos] = pt;
                }
              }
            }
          }
        }
        if (tid == 0 && p.counts_out) { p.counts_out[si] = ns; p.counts_out[ti] = nt; }
        if (ns <= 0 || nt <= 0) early = B200ICP_STATUS_EMPTY;   // uniform: read from shared memory
        __syncthreads();
      } else {
        // ---- stage both scans with TMA bulk copies --------------------------------
        if constexpr (SRCG) src_pts = reinterpret_cast<const double2*>(p.scans) + (size_t)si * p.n_max;
        if (tid == 0) {
          fence_proxy_async();
          mbar_expect_tx(mbar, (uint32_t)((SRCG ? 0 : ns) + nt) * 16u);
          if constexpr (!SRCG) tma_load_1d(s64, p.scans + (size_t)si * p.n_max * 2, (uint32_t)ns * 16u, mbar);
          tma_load_1d(t64, p.scans + (size_t)ti * p.n_max * 2, (uint32_t)nt * 16u, mbar);
        }
        mbar_wait(mbar, parity);
        parity ^= 1u;
      }
    }
    if (early >= 0) {
      if (tid == 0) {
        if (p.status_out) p.status_out[pair] = early;
        if (p.iters_out) p.iters_out[pair] = 0;
        if (p.err_out) p.err_out[pair] = kNaN;
        if (p.T_out)
          for (int k = 0; k < 16; ++k) p.T_out[(size_t)pair * 16 + k] = p.init ? p.init[(size_t)pair * 16 + k] : 0.0;
        if (p.init) {
          const double* T0 = p.init + (size_t)pair * 16;
          store_pose_to_peers(p, pair, T0[3], T0[7], T0[0], T0[4], kNaN);
        }
        ++retired;
      }
      if (p.corr_out)
        for (int j = tid; j < p.n_max; j += W * 32) p.corr_out[(size_t)pair * p.n_max + j] = -1;
      continue;
    }

    // ---- fp32 copy of the target (+ sentinels), float64 sentinels, max |coord|, centroid ------
    const int nblk = (nt + kBlockTargets - 1) / kBlockTargets;
    float amax = 0.f;
    double cx = 0.0, cy = 0.0;
    for (int i = tid; i < L.np; i += W * 32) {
      float2 v = make_float2(kBigCoord, kBigCoord);
      if (i < nt) {
        const double2 t = t64[i];
        v = make_float2((float)t.x, (float)t.y);
        amax = fmaxf(amax, fmaxf(fabsf(v.x), fabsf(v.y)));
        cx += t.x; cy += t.y;
      } else {
        t64[i] = make_double2(1.0e300, 1.0e300);   // a distance of +inf: never a match, no index test needed
      }
      float* g = t32 + (i >> 2) * 8 + (i & 3);
      g[0] = v.x; g[4] = v.y;
    }
    amax = warp_maxf(amax);
    cx = warp_sum(cx);
    cy = warp_sum(cy);
    if (lane == 0) { red[warp * 3 + 0] = (double)amax; red[warp * 3 + 1] = cx; red[warp * 3 + 2] = cy; }
    __syncthreads();
    // bounding boxes of the real fp32 targets of every 32-target block; blocks past the scan get a
    // box at +inf (their bound is +inf: never seeds, never within a threshold)
    for (int b = warp; b < 64; b += W) {
      float4 box = make_float4(kInfF, kInfF, kInfF, kInfF);
      if (b < nblk) {
        const int i = b * kBlockTargets + lane;
        const float2 v = target32(t32, i);
        const bool real = i < nt;
        box = make_float4(redux_minf(real ? v.x : kInfF), redux_maxf(real ? v.x : -kInfF),
                          redux_minf(real ? v.y : kInfF), redux_maxf(real ? v.y : -kInfF));
      }
      if (lane == 0) bb[b] = box;
    }
    float tmax = 0.f;
    double ox = 0.0, oy = 0.0;
#pragma unroll
    for (int w = 0; w < W; ++w) {
      tmax = fmaxf(tmax, (float)red[w * 3 + 0]);
      ox += red[w * 3 + 1]; oy += red[w * 3 + 2];
    }
    tmax = tmax * (1.0f + 4.0f * kU);   // fp32 rounding of the inputs
    ox /= (double)nt; oy /= (double)nt;  // shift origin for the one-pass covariance
    __syncthreads();                     // red[] is reused by the passes below; bb[] complete

    const double epsilon = p.nearest_only ? 0.0 : p.epsilon[pair];
    const bool normalize = p.nearest_only ? false : (p.normalize[pair] != 0);
    const unsigned vmask_lo = nblk >= 32 ? full : ((1u << nblk) - 1u);
    const unsigned vmask_hi = nblk <= 32 ? 0u : (nblk >= 64 ? full : ((1u << (nblk - 32)) - 1u));
    const int rows = (ns + kRowQ - 1) / kRowQ;
    const ulonglong2* tp = reinterpret_cast<const ulonglong2*>(t32);
    StopState st{0.0, 0, 0};
    int passes = 0;
    unsigned tiles = 0;     // (32-source row) x (32-target block) tiles this warp swept for this pair

    for (;;) {  // ------------------------------ one ICP pass ------------------------
      double acc[kSums2d];
#pragma unroll
      for (int k = 0; k < kSums2d; ++k) acc[k] = 0.0;

#pragma unroll 1
      for (int row = warp; row < rows; row += W) {
        const int j0 = row * kRowQ + lane;
        double qdx[R], qdy[R];
        float fqx[R], fqy[R];
#pragma unroll
        for (int q = 0; q < R; ++q) {
          // ragged tail: slots past the scan repeat its last point (no predicates below; they are
          // left out of the sums)
          const double2 s = src_pts[min(j0 + 32 * q, ns - 1)];
          qdx[q] = fma(r00, s.x, fma(r01, s.y, tx));
          qdy[q] = fma(r10, s.x, fma(r11, s.y, ty));
          fqx[q] = (float)qdx[q]; fqy[q] = (float)qdy[q];
        }
        // row box, this lane's two block bounds
        float xlo = fqx[0], xhi = fqx[0], ylo = fqy[0], yhi = fqy[0];
#pragma unroll
        for (int q = 1; q < R; ++q) {
          xlo = fminf(xlo, fqx[q]); xhi = fmaxf(xhi, fqx[q]);
          ylo = fminf(ylo, fqy[q]); yhi = fmaxf(yhi, fqy[q]);
        }
        const float4 qb = make_float4(redux_minf(xlo), redux_maxf(xhi), redux_minf(ylo), redux_maxf(yhi));
        const float lb0 = box_lb(qb, bb[lane]);
        const float lb1 = box_lb(qb, bb[lane + 32]);
        unsigned m_lo, m_hi;
        if constexpr (PRUNE) {
          // seeds: every block whose box overlaps the row box (bound 0: it passes any threshold,
          // so sweeping it first is never wasted); none -> the block of smallest bound
          m_lo = __ballot_sync(full, lb0 == 0.f) & vmask_lo;
          m_hi = __ballot_sync(full, lb1 == 0.f) & vmask_hi;
          if ((m_lo | m_hi) == 0u) {
            const unsigned k0 = (__float_as_uint(lb0) & 0xffffffc0u) | (unsigned)lane;
            const unsigned k1 = (__float_as_uint(lb1) & 0xffffffc0u) | (unsigned)(lane + 32);
            const unsigned f = __reduce_min_sync(full, min(k0, k1)) & 63u;
            m_lo = (f < 32u ? (1u << f) : 0u) & vmask_lo;
            m_hi = (f >= 32u ? (1u << (f - 32u)) : 0u) & vmask_hi;
          }
        } else {
          m_lo = vmask_lo; m_hi = vmask_hi;
        }

        unsigned best[R], second[R];
        float gbest[R][8], thr[R], a_bound[R];
        f32x2 qxx[R], qyy[R];
#pragma unroll
        for (int q = 0; q < R; ++q) {
          qxx[q] = pack2(fqx[q], fqx[q]); qyy[q] = pack2(fqy[q], fqy[q]);
          a_bound[q] = fmaxf(fabsf(fqx[q]), fabsf(fqy[q])) * (1.0f + 4.0f * kU) + tmax;
        }
        unsigned done_lo = m_lo, done_hi = m_hi;
        {   // the first block of the row (the seeds are never empty: a scan has at least one block)
          const int b0 = m_lo ? __ffs((int)m_lo) - 1 : 32 + __ffs((int)m_hi) - 1;
          if (m_lo) m_lo &= m_lo - 1u; else m_hi &= m_hi - 1u;
          tiles += (unsigned)R;
          sweep_block_rows<R, true>(tp, b0, qxx, qyy, best, second, gbest);
        }
        // the other seeds: blocks that overlap the row box (or, with PRUNE off, every block)
        while (m_lo) {
          const int b = __ffs((int)m_lo) - 1;
          m_lo &= m_lo - 1u;
          tiles += (unsigned)R;
          sweep_block_rows<R>(tp, b, qxx, qyy, best, second, gbest);
        }
        while (m_hi) {
          const int b = 32 + __ffs((int)m_hi) - 1;
          m_hi &= m_hi - 1u;
          tiles += (unsigned)R;
          sweep_block_rows<R>(tp, b, qxx, qyy, best, second, gbest);
        }
        // the threshold the resolution uses (final once no further block is swept)
        auto thresholds = [&]() -> float {
          float thr_max = 0.f;
#pragma unroll
          for (int q = 0; q < R; ++q) {
            thr[q] = candidate_threshold(__uint_as_float(best[q] & 0xffffffc0u), a_bound[q], kTruncRel);
            thr_max = fmaxf(thr_max, thr[q]);
          }
          return thr_max;
        };
        const float thr_seed = thresholds();
        if constexpr (PRUNE) {
          // the other blocks that can still hold a candidate of any query of this row: within the row's
          // bound, and then only if at least one lane's OWN point-to-box bound is within that lane's
          // threshold (the candidate arithmetic on a degenerate box: a block refused here holds no target
          // with d32 <= thr(q) for any query of the row) -- typically the neighbours of the overlapping
          // blocks, which only the queries at the ends of the row can reach
          const float bound = redux_maxf(thr_seed);
          m_lo = __ballot_sync(full, !(lb0 > bound)) & vmask_lo & ~done_lo;
          m_hi = __ballot_sync(full, !(lb1 > bound)) & vmask_hi & ~done_hi;
          bool swept_more = false;
          auto admitted = [&](int b) -> bool {
            const float4 box = bb[b];
            bool need = false;
#pragma unroll
            for (int q = 0; q < R; ++q)
              need = need || !(box_lb(make_float4(fqx[q], fqx[q], fqy[q], fqy[q]), box) > thr[q]);
            return __any_sync(full, need);
          };
          while (m_lo) {
            const int b = __ffs((int)m_lo) - 1;
            m_lo &= m_lo - 1u;
            if (!admitted(b)) continue;
            swept_more = true;
            tiles += (unsigned)R;
            sweep_block_rows<R>(tp, b, qxx, qyy, best, second, gbest);
          }
          while (m_hi) {
            const int b = 32 + __ffs((int)m_hi) - 1;
            m_hi &= m_hi - 1u;
            if (!admitted(b)) continue;
            swept_more = true;
            tiles += (unsigned)R;
            sweep_block_rows<R>(tp, b, qxx, qyy, best, second, gbest);
          }
          if (swept_more) thresholds();     // best may have improved: the final thresholds
        }

        // exact resolution in float64
        int c[R];
        bool amb[R];
#pragma unroll
        for (int q = 0; q < R; ++q) {
          c[q] = -1;
          const float sval = __uint_as_float(second[q] & 0xffffffc0u);
          // a second swept block within the slack makes the query ambiguous; no second block at all
          // (every other block was pruned, or the target has one block) does not
          amb[q] = second[q] != 0xffffffffu && !(sval > thr[q]);
          if (!amb[q]) {
            const int blk = (int)(best[q] & 63u);
            // groups of four targets whose minimum (kept from the sweep) is within the slack: all
            // four members are re-evaluated in float64 -- a superset of the candidates the slack
            // admits, so the result is the float64 argmin (lowest index on exact ties).  Slots past
            // the scan hold a sentinel whose distance is +inf.
            unsigned gmask = 0;
#pragma unroll
            for (int g = 0; g < 8; ++g)
              if (gbest[q][g] <= thr[q]) gmask |= 1u << g;
            double bd = kInf;
            const double2* tb = t64 + blk * kBlockTargets;
            while (gmask) {
              const int g = __ffs((int)gmask) - 1;
              gmask &= gmask - 1u;
#pragma unroll
              for (int u = 0; u < 4; ++u) {
                const double2 t = tb[4 * g + u];
                const double ex = qdx[q] - t.x, ey = qdy[q] - t.y;
                const double d = ex * ex + ey * ey;
                if (d < bd) { bd = d; c[q] = blk * kBlockTargets + 4 * g + u; }
              }
            }
            if (c[q] < 0) amb[q] = true;  // non-finite data only
          }
        }
        // rare: a second block is within the slack -> the warp scans all targets
#pragma unroll
        for (int q = 0; q < R; ++q) {
          unsigned vote = __ballot_sync(full, amb[q]);
          while (vote) {
            const int src_lane = __ffs((int)vote) - 1;
            vote &= vote - 1u;
            const float fx = __shfl_sync(full, fqx[q], src_lane);
            const float fy = __shfl_sync(full, fqy[q], src_lane);
            const float thr_s = __shfl_sync(full, thr[q], src_lane);
            const double sdx = __shfl_sync(full, qdx[q], src_lane);
            const double sdy = __shfl_sync(full, qdy[q], src_lane);
            double bd = kInf;
            int bi = 0x7fffffff;
            for (int i = lane; i < nt; i += 32) {
              const float2 t = target32(t32, i);
              // !(d > thr) also admits NaN so that non-finite data still yields an index
              if (!(dist2_2d(fx, fy, t.x, t.y) > thr_s)) {
                const double2 td = t64[i];
                const double ex = sdx - td.x, ey = sdy - td.y;
                const double d = ex * ex + ey * ey;
                if (d < bd || bi == 0x7fffffff) { bd = d; bi = i; }
              }
            }
            warp_argmin(bd, bi);
            if (lane == src_lane) c[q] = (bi == 0x7fffffff) ? 0 : bi;
          }
        }
        // float64 sums over the matched pairs
#pragma unroll
        for (int q = 0; q < R; ++q) {
          const int j = j0 + 32 * q;
          if (j < ns) {
            const double2 t = t64[c[q]];
            const double ux = qdx[q] - ox, uy = qdy[q] - oy, vx = t.x - ox, vy = t.y - oy;
            acc[0] += ux; acc[1] += uy; acc[2] += vx; acc[3] += vy;
            acc[4] = fma(ux, vx, acc[4]); acc[5] = fma(ux, vy, acc[5]);
            acc[6] = fma(uy, vx, acc[6]); acc[7] = fma(uy, vy, acc[7]);
            const double ex = qdx[q] - t.x, ey = qdy[q] - t.y;
            const double d = ex * ex + ey * ey;
            acc[8] += d;
            corr_s[j] = (unsigned short)c[q];
            if (p.nearest_only && p.dist2_out) p.dist2_out[(size_t)pair * p.n_max + j] = d;
          }
        }
      }  // rows

#pragma unroll
      for (int k = 0; k < kSums2d; ++k) acc[k] = warp_sum(acc[k]);
      double* slot = red + (size_t)(passes & 1) * W * kRed2d;
      if (normalize) {
        // bounding box of the moved source and the matched targets (icp.py:91-95): a separate light
        // pass over this warp's rows (its own matches, written above by the same lanes), so that the
        // row loop carries no extrema; only pairs with a normalised error come here
        double mn[4] = {kInf, kInf, kInf, kInf}, mx[4] = {-kInf, -kInf, -kInf, -kInf};
        for (int row = warp; row < rows; row += W) {
#pragma unroll
          for (int q = 0; q < R; ++q) {
            const int j = row * kRowQ + 32 * q + lane;
            if (j < ns) {
              const double2 s = src_pts[j];
              const double x = fma(r00, s.x, fma(r01, s.y, tx)), y = fma(r10, s.x, fma(r11, s.y, ty));
              const double2 t = t64[corr_s[j]];
              mn[0] = fmin(mn[0], x); mx[0] = fmax(mx[0], x);
              mn[1] = fmin(mn[1], y); mx[1] = fmax(mx[1], y);
              mn[2] = fmin(mn[2], t.x); mx[2] = fmax(mx[2], t.x);
              mn[3] = fmin(mn[3], t.y); mx[3] = fmax(mx[3], t.y);
            }
          }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) { mn[k] = warp_min(mn[k]); mx[k] = warp_max(mx[k]); }
        if (lane == 0) {
#pragma unroll
          for (int k = 0; k < 4; ++k) { slot[warp * kRed2d + kSums2d + 2 * k] = mn[k]; slot[warp * kRed2d + kSums2d + 2 * k + 1] = mx[k]; }
        }
      }
      // exchange: one shared-memory round per pass, double-buffered by pass parity
      if (lane == 0) {
#pragma unroll
        for (int k = 0; k < kSums2d; ++k) slot[warp * kRed2d + k] = acc[k];
      }
      __syncthreads();
      // Column k of the W partial rows is reduced by lane k (same order as a sequential sum over
      // the warps), then every lane fetches the totals it needs by shuffle.
      double tot[kRed2d];
      {
        const int kcol = lane < kRed2d ? lane : kRed2d - 1;
        double col = (kcol < kSums2d || normalize) ? slot[kcol] : 0.0;
#pragma unroll
        for (int w = 1; w < W; ++w) {
          if (kcol < kSums2d || normalize) {
            const double o = slot[w * kRed2d + kcol];
            col = kcol < kSums2d ? col + o : ((kcol & 1) ? fmin(col, o) : fmax(col, o));   // 9, 11, ..: minima
          }
        }
#pragma unroll
        for (int k = 0; k < kSums2d; ++k) tot[k] = __shfl_sync(full, col, k);
#pragma unroll
        for (int k = kSums2d; k < kRed2d; ++k) tot[k] = normalize ? __shfl_sync(full, col, k) : 0.0;
      }
      passes += 1;

      // error of this pass (before the update), solve, compose, stop tests
      double err = tot[8];
      if (normalize) {
        const double wx = fmax(tot[10], tot[14]) - fmin(tot[9], tot[13]);
        const double wy = fmax(tot[12], tot[16]) - fmin(tot[11], tot[15]);
        const double diag = sqrt(wx * wx + wy * wy);
        err = err / (diag * diag * (double)ns);
      }
      int status;
      if (p.nearest_only) {
        status = B200ICP_STATUS_EPSILON;
      } else {
        const double inv_n = 1.0 / (double)ns;
        const double mqx = tot[0] * inv_n, mqy = tot[1] * inv_n;
        const double mpx = tot[2] * inv_n, mpy = tot[3] * inv_n;
        const double nn = (double)ns;
        const double c00 = tot[4] - nn * mqx * mpx, c01 = tot[5] - nn * mqx * mpy;
        const double c10 = tot[6] - nn * mqy * mpx, c11 = tot[7] - nn * mqy * mpy;
        double d00, d01, d10, d11, dz;
        if (c00 * c11 - c01 * c10 >= 0.0) {
          const double num = c01 - c10, den = c00 + c11;
          const double h = hypot(num, den);
          const double co = (h == 0.0) ? 1.0 : den / h, sn = (h == 0.0) ? 0.0 : num / h;
          d00 = co; d01 = -sn; d10 = sn; d11 = co; dz = 1.0;
        } else {
          const double num = c01 + c10, den = c00 - c11;
          const double h = hypot(num, den);
          const double co = den / h, sn = num / h;
          d00 = co; d01 = sn; d10 = sn; d11 = -co; dz = -1.0;
        }
        const double qbx = ox + mqx, qby = oy + mqy, pbx = ox + mpx, pby = oy + mpy;
        const double dtx = pbx - (d00 * qbx + d01 * qby);
        const double dty = pby - (d10 * qbx + d11 * qby);
        const double n00 = d00 * r00 + d01 * r10, n01 = d00 * r01 + d01 * r11;
        const double n10 = d10 * r00 + d11 * r10, n11 = d10 * r01 + d11 * r11;
        const double ntx = d00 * tx + d01 * ty + dtx, nty = d10 * tx + d11 * ty + dty;
        r00 = n00; r01 = n01; r10 = n10; r11 = n11; tx = ntx; ty = nty; zs *= dz;
        status = stop_test(st, err, epsilon, p.max_iters, p.stopping_thresh);
      }
      if (status >= 0) {
        if (lane == 0) atomicAdd(reinterpret_cast<unsigned long long*>(p.queue) + 1, (unsigned long long)tiles);
        if (tid == 0) {
          if (p.T_out) {
            double* T = p.T_out + (size_t)pair * 16;
            T[0] = r00; T[1] = r01; T[2] = 0.0; T[3] = tx;
            T[4] = r10; T[5] = r11; T[6] = 0.0; T[7] = ty;
            T[8] = 0.0; T[9] = 0.0; T[10] = zs; T[11] = 0.0;
            T[12] = 0.0; T[13] = 0.0; T[14] = 0.0; T[15] = 1.0;
          }
          if (p.err_out) p.err_out[pair] = err;
          if (p.iters_out) p.iters_out[pair] = passes;
          if (p.status_out) p.status_out[pair] = status;
          store_pose_to_peers(p, pair, tx, ty, r00, r10, err);
          ++retired;
        }
        if (p.corr_out) {
          // every lane wrote the matches of its own rows; the barrier of the exchange above made
          // them visible to the whole CTA
          int* out = p.corr_out + (size_t)pair * p.n_max;
          for (int j = tid; j < p.n_max; j += W * 32) out[j] = j < ns ? (int)corr_s[j] : -1;
        }
        break;
      }
    }
  }
}

}  // namespace b200icp
