// Fused planar ICP: one CTA aligns one (source, target) scan pair from staging to
// the final pose; the whole iterate-until-converged loop stays on the device.
//
// Replaces modules/icp.py:123-189 (run_icp) for 2-D scans, i.e. the call shapes
// of modules/localization.py:120-123 and main.py:162.
//
// Per pair:
//   stage   TMA bulk copies (cp.async.bulk + mbarrier) bring the float64 source
//           and target scans into shared memory; the target is also kept as
//           fp32 (x,y) pairs for the search.                      [icp.py:153-161]
//   pass k  1. q = T_k s in float64, rounded to fp32 for the search [icp.py:116]
//           2. fp32 nearest-neighbour sweep over 32-target blocks, targets broadcast
//              from shared memory as LDS.128 (two targets), packed fp32x2 math,
//              FMNMX3 minima; per row of 32 queries the best and second-best
//              (value | block id) are kept.  Default (PRUNE): one row at a time,
//              blocks whose bounding-box lower bound exceeds the candidate
//              threshold are skipped -- exact, see sweep_row_pruned.  PRUNE=false
//              (B200ICP_SWEEP=exhaustive): every block, R rows per thread in
//              registers.                                           [icp.py:117]
//           3. exactness: the best block is rescanned, every candidate within a
//              proven slack of the fp32 minimum is re-evaluated in float64 from
//              the original coordinates; if a second block is within the slack
//              the warp rescans all targets cooperatively.  Result == float64
//              brute force (lowest index on exact ties).
//           4. float64 sums (centroids, cross-covariance, SSE, bbox), warp
//              shuffle + one shared-memory exchange.              [icp.py:55-61,88-95]
//           5. closed-form planar Kabsch (rotation or reflection branch), pose
//              composition, the reference's three stop tests.  [icp.py:62-72,119,172-181]
#pragma once
#include "icp_common.cuh"

namespace b200icp {

#ifdef B200ICP_PHASE_PROFILE
// developer instrumentation: cycles per phase, summed over warp 0 of every CTA into queue[16 + 2*i]
#define PHASE_DECL long long ph_t = clock64(); long long ph_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#define PHASE(i) do { const long long now_ = clock64(); ph_acc[i] += now_ - ph_t; ph_t = now_; } while (0)
#define PHASE_FLUSH do { if (tid == 0) for (int i_ = 0; i_ < 8; ++i_) \
    atomicAdd(reinterpret_cast<unsigned long long*>(p.queue) + 8 + i_, (unsigned long long)ph_acc[i_]); } while (0)
#else
#define PHASE_DECL
#define PHASE(i) do { } while (0)
#define PHASE_FLUSH do { } while (0)
#endif

constexpr int kSums2d = 9;    // sum qx qy px py qxpx qxpy qypx qypy sse   (shifted by origin)
constexpr int kRed2d = 17;    // + min/max of qx qy px py

struct Smem2dLayout {
  int np;          // padded point capacity (multiple of 32)
  size_t off_t64, off_s64, off_t32, off_red, off_part, off_bb, off_rb, off_ctl, total;
};
// prune = true: the block-pruned sweep (bounding boxes of the target blocks and of the source rows
// instead of the shared-row exchange area).
__host__ __device__ inline Smem2dLayout smem2d_layout(int n_max, int warps, bool prune, int rows_per_thread) {
  Smem2dLayout L;
  L.np = ((n_max + 31) / 32) * 32;
  size_t o = 0;
  L.off_t64 = o; o += (size_t)L.np * 16;
  L.off_s64 = o; o += (size_t)L.np * 16;
  L.off_t32 = o; o += (size_t)L.np * 8;
  L.off_red = o; o += (size_t)2 * warps * kRed2d * 8;
  L.off_part = o; o += prune ? 0 : (size_t)(warps - 1) * warps * 32 * 8;   // shared rows: (best, second) per warp and lane
  L.off_bb = o; o += prune ? (size_t)64 * 16 : 0;                           // target-block boxes (xmin, xmax, ymin, ymax)
  L.off_rb = o; o += prune ? (size_t)warps * 128 : 0;                       // per-warp scratch: bbox of the matched pairs
  L.off_ctl = o; o += 64;
  L.total = o;
  return L;
}

// Shared-memory layout of the fp32 target copy: groups of four targets stored as
// [x0 x1 x2 x3 y0 y1 y2 y3], so one LDS.128 yields two (x,x) register pairs ready for the
// packed fp32x2 instructions.
__device__ __forceinline__ float2 target32(const float* t32, int i) {
  const float* g = t32 + (i >> 2) * 8 + (i & 3);
  return make_float2(g[0], g[4]);
}

// One 32-target block of the NN sweep for NR source points per thread (rows of 32 sources per warp).
// Per 4 candidates: 2x(FADD2, FADD2, FMUL2, FFMA2) + 2 FMNMX3, i.e. 8 FP32-pipe issue slots
// carrying the 16 algorithmic lane-instructions, plus 0.5 ALU slot per candidate.
template <int NR, int RMAX>
__device__ __forceinline__ void sweep_block(const ulonglong2* tp, int b, const f32x2 (&qxx)[NR], const f32x2 (&qyy)[NR],
                                            unsigned (&best)[RMAX], unsigned (&second)[RMAX]) {
  float bm[NR];
#pragma unroll
  for (int r = 0; r < NR; ++r) bm[r] = __int_as_float(0x7f800000);
#pragma unroll
  for (int i = 0; i < kBlockTargets / 4; ++i) {
    const ulonglong2 X = tp[b * (kBlockTargets / 2) + 2 * i];
    const ulonglong2 Y = tp[b * (kBlockTargets / 2) + 2 * i + 1];
#pragma unroll
    for (int r = 0; r < NR; ++r) {
      f32x2 dx = sub2(qxx[r], X.x), dy = sub2(qyy[r], Y.x);
      const f32x2 d01 = fma2(dx, dx, mul2(dy, dy));
      dx = sub2(qxx[r], X.y); dy = sub2(qyy[r], Y.y);
      const f32x2 d23 = fma2(dx, dx, mul2(dy, dy));
      float d0, d1, d2, d3;
      unpack2(d01, d0, d1);
      unpack2(d23, d2, d3);
      bm[r] = fmin3(fmin3(bm[r], d0, d1), d2, d3);
    }
  }
#pragma unroll
  for (int r = 0; r < NR; ++r) {
    const unsigned pk = (__float_as_uint(bm[r]) & 0xffffffc0u) | (unsigned)b;
    second[r] = min(second[r], max(best[r], pk));
    best[r] = min(best[r], pk);
  }
}

// One block for ONE row, keeping the minima of its eight groups of four targets: when the block
// becomes the row's best block they are saved in gbest[], so that the resolution can tell which
// groups hold candidates without evaluating the block's 32 distances a second time.
__device__ __forceinline__ void sweep_block_row(const ulonglong2* tp, int b, f32x2 qxx, f32x2 qyy, unsigned& best,
                                                unsigned& second, float (&gbest)[kBlockTargets / 4]) {
  float gm[kBlockTargets / 4];
#pragma unroll
  for (int i = 0; i < kBlockTargets / 4; ++i) {
    const ulonglong2 X = tp[b * (kBlockTargets / 2) + 2 * i];
    const ulonglong2 Y = tp[b * (kBlockTargets / 2) + 2 * i + 1];
    f32x2 dx = sub2(qxx, X.x), dy = sub2(qyy, Y.x);
    const f32x2 d01 = fma2(dx, dx, mul2(dy, dy));
    dx = sub2(qxx, X.y); dy = sub2(qyy, Y.y);
    const f32x2 d23 = fma2(dx, dx, mul2(dy, dy));
    float d0, d1, d2, d3;
    unpack2(d01, d0, d1);
    unpack2(d23, d2, d3);
    gm[i] = fminf(fmin3(d0, d1, d2), d3);
  }
  static_assert(kBlockTargets == 32, "eight groups per block");
  const float bm = fminf(fmin3(fmin3(fmin3(gm[0], gm[1], gm[2]), gm[3], gm[4]), gm[5], gm[6]), gm[7]);
  const unsigned pk = (__float_as_uint(bm) & 0xffffffc0u) | (unsigned)b;
  if (pk < best) {
#pragma unroll
    for (int i = 0; i < kBlockTargets / 4; ++i) gbest[i] = gm[i];
  }
  second = min(second, max(best, pk));
  best = min(best, pk);
}

// Exhaustive sweep over the blocks [blk_lo, blk_hi).
template <int NR, int RMAX>
__device__ __forceinline__ void sweep2d(const float* t32, int blk_lo, int blk_hi, const float (&qx)[RMAX],
                                        const float (&qy)[RMAX], unsigned (&best)[RMAX],
                                        unsigned (&second)[RMAX]) {
  f32x2 qxx[NR], qyy[NR];
#pragma unroll
  for (int r = 0; r < NR; ++r) { qxx[r] = pack2(qx[r], qx[r]); qyy[r] = pack2(qy[r], qy[r]); }
  const ulonglong2* tp = reinterpret_cast<const ulonglong2*>(t32);
  for (int b = blk_lo; b < blk_hi; ++b) sweep_block<NR, RMAX>(tp, b, qxx, qyy, best, second);
}

// ---------------------------------------------------------------------------------------------
// Block-pruned sweep.  Exactly the same matches as the exhaustive sweep, with the blocks that
// cannot matter left out.
//
// Every target block b has a bounding box B_b of its fp32 points, every source row r (32 queries
// of one warp) a bounding box Q_r of its fp32 queries.  lb(Q_r, B_b) is computed with the SAME
// operations as a candidate distance (FADD, FADD, FMUL, FFMA on the per-axis gaps); rounding to
// nearest is monotone, so lb(Q_r, B_b) <= d32(q, t) for every q in the row and every t in the
// block -- no error term is needed.
//   1. seeds: per row the two blocks of smallest lb are swept first (REDUX.MIN over the lanes,
//      lane l evaluates blocks l and l + 32).
//   2. bound: thr(q) = candidate_threshold(best(q)) is the value the resolution phase will use to
//      decide which targets are candidates and whether a second block is ambiguous; it is
//      monotone in best(q), and best(q) only decreases, so thr_now(q) >= thr_final(q).  A block
//      with lb(Q_r, B_b) > max_q thr_now(q) for every row of the warp therefore holds no target
//      with d32 <= thr_final(q) for any query: it is neither the best block, nor a reason for
//      the ambiguity fallback, nor does it hold a candidate.  Such blocks are skipped.
//   3. the other blocks are swept as usual.
// The (best, second) lists then cover a superset of all blocks with a minimum <= thr_final, which
// is all the resolution reads; the float64 result is the exhaustive one (same tie rule).
// Returns the number of blocks swept (for the evaluated-candidate counter).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float box_lb(const float4 q, const float4 b) {   // boxes: (xmin, xmax, ymin, ymax)
  const float ex = fmaxf(fmaxf(__fsub_rn(b.x, q.y), __fsub_rn(q.x, b.y)), 0.f);
  const float ey = fmaxf(fmaxf(__fsub_rn(b.z, q.w), __fsub_rn(q.z, b.w)), 0.f);
  return __fmaf_rn(ex, ex, __fmul_rn(ey, ey));
}

// fp32 warp reductions in one instruction (CREDUX.MIN/MAX.F32, sm_100a); NaN operands are ignored.
__device__ __forceinline__ float redux_minf(float v) {
  float r;
  asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
  return r;
}
__device__ __forceinline__ float redux_maxf(float v) {
  float r;
  asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
  return r;
}

// One row (32 queries, one per lane; `valid` = this lane holds a real source point).
__device__ __forceinline__ int sweep_row_pruned(const float* t32, const float4* bb, int nblk, bool valid, float qx,
                                                float qy, float tmax, float trunc_rel, unsigned& best_out,
                                                unsigned& second_out, float& thr_out,
                                                float (&gbest)[kBlockTargets / 4]) {
  const int lane = threadIdx.x & 31;
  const float kInfF = __int_as_float(0x7f800000);
  const bool has0 = lane < nblk, has1 = lane + 32 < nblk;
  // row box and this lane's two block bounds
  const float4 q = make_float4(redux_minf(valid ? qx : kInfF), redux_maxf(valid ? qx : -kInfF),
                               redux_minf(valid ? qy : kInfF), redux_maxf(valid ? qy : -kInfF));
  unsigned lb0 = 0xffffffffu, lb1 = 0xffffffffu;
  if (has0) lb0 = __float_as_uint(box_lb(q, bb[lane]));
  if (has1) lb1 = __float_as_uint(box_lb(q, bb[lane + 32]));
  // 1. seeds: the two blocks of smallest bound
  unsigned k0 = has0 ? ((lb0 & 0xffffffc0u) | (unsigned)lane) : 0xffffffffu;
  unsigned k1 = has1 ? ((lb1 & 0xffffffc0u) | (unsigned)(lane + 32)) : 0xffffffffu;
  const unsigned f = __reduce_min_sync(0xffffffffu, min(k0, k1));
  if (k0 == f) k0 = 0xffffffffu;
  if (k1 == f) k1 = 0xffffffffu;
  const unsigned s = __reduce_min_sync(0xffffffffu, min(k0, k1));
  unsigned long long m = 0;
  if (f != 0xffffffffu) m |= 1ull << (f & 63u);
  if (s != 0xffffffffu) m |= 1ull << (s & 63u);

  unsigned best = 0xffffffffu, second = 0xffffffffu;
  const f32x2 qxx = pack2(qx, qx), qyy = pack2(qy, qy);
  const ulonglong2* tp = reinterpret_cast<const ulonglong2*>(t32);
  unsigned long long done = 0;
  int swept = 0;
  const float a_bound = fmaxf(fabsf(qx), fabsf(qy)) * (1.0f + 4.0f * kU) + tmax;
  float thr;
#pragma unroll 1
  for (int phase = 0;; ++phase) {
    swept += __popcll(m);
    done |= m;
    while (m) {
      const int b = __ffsll((long long)m) - 1;
      m &= m - 1;
      sweep_block_row(tp, b, qxx, qyy, best, second, gbest);
    }
    // the threshold the resolution uses (final once no further block is swept)
    thr = candidate_threshold(__uint_as_float(best & 0xffffffc0u), a_bound, trunc_rel);
    if (phase == 1) break;
    // 2. the other blocks that can still hold a candidate of any query of this row
    const unsigned bound = __reduce_max_sync(0xffffffffu, valid ? __float_as_uint(thr) : 0u);
    m = ((unsigned long long)__ballot_sync(0xffffffffu, has0 && !(lb0 > bound)) |
         ((unsigned long long)__ballot_sync(0xffffffffu, has1 && !(lb1 > bound)) << 32)) & ~done;
    if (m == 0) break;
  }
  thr_out = thr;
  best_out = best;
  second_out = second;
  return swept;
}

template <int R, int W, int MINB, bool PRUNE>
__global__ void __launch_bounds__(W * 32, MINB) icp2d_kernel(const RunParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const Smem2dLayout L = smem2d_layout(p.n_max, W, PRUNE, R);
  double2* t64 = reinterpret_cast<double2*>(smem_raw + L.off_t64);
  double2* s64 = reinterpret_cast<double2*>(smem_raw + L.off_s64);
  float* t32 = reinterpret_cast<float*>(smem_raw + L.off_t32);
  double* red = reinterpret_cast<double*>(smem_raw + L.off_red);
  uint2* part = reinterpret_cast<uint2*>(smem_raw + L.off_part);
  float4* bb = reinterpret_cast<float4*>(smem_raw + L.off_bb);
  double* wscr = reinterpret_cast<double*>(smem_raw + L.off_rb) + (threadIdx.x >> 5) * 16;
  uint64_t* mbar = reinterpret_cast<uint64_t*>(smem_raw + L.off_ctl);
  int* s_pair = reinterpret_cast<int*>(smem_raw + L.off_ctl + 16);

  const int tid = threadIdx.x;
  const int lane = tid & 31;
  const int warp = tid >> 5;
  constexpr float kTruncRel = 1.52587890625e-5f;  // 2^-16 >= the 6 truncated mantissa bits (2^-17)
  const float kInfF32 = __int_as_float(0x7f800000);
  const double kInf = __longlong_as_double(0x7ff0000000000000LL);

  if (tid == 0) {
    mbar_init(mbar, 1);
    fence_mbar_init();
  }
  uint32_t parity = 0;
  unsigned retired = 0;   // pairs this thread (tid 0) has retired, announced to the peers when the CTA leaves the queue
  PHASE_DECL

  for (;;) {
    __syncthreads();  // previous pair fully consumed; mbarrier initialised
    PHASE(7);
    if (tid == 0) {
      const int pr = (int)atomicAdd(p.queue, 1u);
      s_pair[0] = pr;
      int resident = 1;
      if (p.ready_scans && pr < p.n_pairs) {
        // Gated launch: wait until the copy stream has delivered both scans of this pair.  Bounded:
        // a caller whose copies never arrive must not hang the device.  After 5 s the pair is given
        // up (status NOT_RESIDENT, it is not staged), queue[5] tells the host side, and from then on
        // nobody waits: pairs whose scans are still missing are given up at once.
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
      s_pair[1] = resident;
    }
    __syncthreads();
    const int pair = s_pair[0];
    if (pair >= p.n_pairs) {
      if (tid == 0) pairs_retired(p, p.queue + 6, retired);    // this CTA's pairs, announced once
      break;
    }

    const int si = p.pairs[2 * pair], ti = p.pairs[2 * pair + 1];
    int early = -1;
    int ns = 0, nt = 0;
    if (si < 0 || ti < 0 || si >= p.n_scans || ti >= p.n_scans) {
      early = B200ICP_STATUS_SKIPPED;
    } else if (!s_pair[1]) {
      early = B200ICP_STATUS_NOT_RESIDENT;
    } else {
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
    if (early >= 0) {
      if (tid == 0) {
        if (p.status_out) p.status_out[pair] = early;
        if (p.iters_out) p.iters_out[pair] = 0;
        if (p.err_out) p.err_out[pair] = __longlong_as_double(0x7ff8000000000000LL);
        if (p.T_out)
          for (int k = 0; k < 16; ++k) p.T_out[(size_t)pair * 16 + k] = p.init ? p.init[(size_t)pair * 16 + k] : 0.0;
        if (p.init) {
          const double* T0 = p.init + (size_t)pair * 16;
          store_pose_to_peers(p, pair, T0[3], T0[7], T0[0], T0[4], __longlong_as_double(0x7ff8000000000000LL));
        }
        ++retired;
      }
      if (p.corr_out)
        for (int j = tid; j < p.n_max; j += W * 32) p.corr_out[(size_t)pair * p.n_max + j] = -1;
      continue;
    }

    // ---- stage both scans with TMA bulk copies --------------------------------
    if (tid == 0) {
      fence_proxy_async();
      mbar_expect_tx(mbar, (uint32_t)(ns + nt) * 16u);
      tma_load_1d(s64, p.scans + (size_t)si * p.n_max * 2, (uint32_t)ns * 16u, mbar);
      tma_load_1d(t64, p.scans + (size_t)ti * p.n_max * 2, (uint32_t)nt * 16u, mbar);
    }
    mbar_wait(mbar, parity);
    parity ^= 1u;

    // fp32 copy of the target (+ sentinels), max |coord| and centroid of the target
    float amax = 0.f;
    double cx = 0.0, cy = 0.0;
    for (int i = tid; i < L.np; i += W * 32) {
      float2 v = make_float2(kBigCoord, kBigCoord);
      if (i < nt) {
        double2 t = t64[i];
        v = make_float2((float)t.x, (float)t.y);
        amax = fmaxf(amax, fmaxf(fabsf(v.x), fabsf(v.y)));
        cx += t.x; cy += t.y;
      }
      float* g = t32 + (i >> 2) * 8 + (i & 3);
      g[0] = v.x; g[4] = v.y;
    }
    amax = warp_maxf(amax);
    cx = warp_sum(cx);
    cy = warp_sum(cy);
    if (lane == 0) { red[warp * 3 + 0] = (double)amax; red[warp * 3 + 1] = cx; red[warp * 3 + 2] = cy; }
    __syncthreads();
    if constexpr (PRUNE) {   // bounding boxes of the real fp32 targets of every 32-target block
      const float kInfF = __int_as_float(0x7f800000);
      for (int b = warp; b < (nt + kBlockTargets - 1) / kBlockTargets; b += W) {
        const int i = b * kBlockTargets + lane;
        float2 v = target32(t32, i);
        const bool real = i < nt;
        const float xlo = warp_minf(real ? v.x : kInfF), xhi = warp_maxf(real ? v.x : -kInfF);
        const float ylo = warp_minf(real ? v.y : kInfF), yhi = warp_maxf(real ? v.y : -kInfF);
        if (lane == 0) bb[b] = make_float4(xlo, xhi, ylo, yhi);
      }
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
    __syncthreads();                     // red[] is reused by the passes below
    if constexpr (PRUNE) {
      // The pose and the origin are warp-uniform: the pruned kernel keeps them in its per-warp
      // scratch (read back where needed) instead of 16 registers per thread across the row loop.
      if (lane == 0) {
        wscr[8] = r00; wscr[9] = r01; wscr[10] = r10; wscr[11] = r11; wscr[12] = tx; wscr[13] = ty;
        wscr[14] = ox; wscr[15] = oy;
      }
      __syncwarp();
    }

    PHASE(0);
    const double epsilon = p.nearest_only ? 0.0 : p.epsilon[pair];
    const bool normalize = p.nearest_only ? false : (p.normalize[pair] != 0);
    const int nblk = (nt + kBlockTargets - 1) / kBlockTargets;
    // Rows of 32 source points.  Every warp owns `base` rows and sweeps them against all
    // targets.  The rows % W rows left over are SHARED: every warp sweeps all of them against its
    // 1/W of the target blocks, the per-warp (best, second) pairs meet in shared memory, and
    // shared row e is then merged, resolved and summed by warp e (register slot `base`).  All
    // warps therefore carry the same sweep load (34 rows on 8 warps: 4.25 each instead of 5/4).
    const int rows = (ns + 31) >> 5;
    const int base = rows / W, extra = rows % W;
    const int nr = base + (warp < extra ? 1 : 0);
    const int n_shared_chunks = (extra + R - 1) / R;
    const int sblk_lo = (warp * nblk) / W, sblk_hi = ((warp + 1) * nblk) / W;
    auto row_of = [&](int r) { return r < base ? warp * base + r : base * W + warp; };
    StopState st{0.0, 0, 0};
    int passes = 0;
    unsigned tiles = 0;     // (32-source row) x (32-target block) tiles this warp swept for this pair

    for (;;) {  // ------------------------------ one ICP pass ------------------------
      double acc[kRed2d];
#pragma unroll
      for (int k = 0; k < kSums2d; ++k) acc[k] = 0.0;
#pragma unroll
      for (int k = 0; k < 4; ++k) { acc[kSums2d + 2 * k] = kInf; acc[kSums2d + 2 * k + 1] = -kInf; }
      int corr[R];
#pragma unroll
      for (int r = 0; r < R; ++r) corr[r] = -1;

      float qx[R], qy[R];
      unsigned best[R], second[R];
      if constexpr (PRUNE) {
        // bounding box of the matched pairs (normalised error only): kept per warp in shared
        // memory instead of 16 registers per thread (minima and maxima are order-independent)
        if (normalize && lane < 8) wscr[lane] = (lane & 1) ? -kInf : kInf;
        __syncwarp();
        // One rolled loop over this warp's rows: transform, pruned fp32 sweep, exact resolution,
        // sums.  (A rolled loop keeps the kernel small: the exhaustive variant's five-row unrolled
        // phases no longer pay once the sweep is pruned -- instruction fetch was the top stall.)
#pragma unroll 1
        for (int r = 0; r < nr; ++r) {
          const int j = row_of(r) * 32 + lane;
          const bool v = j < ns;
          double qdx = 0.0, qdy = 0.0;
          if (v) {
            const volatile double* ws = wscr;
            const double2 s = s64[j];
            qdx = fma(ws[8], s.x, fma(ws[9], s.y, ws[12]));
            qdy = fma(ws[10], s.x, fma(ws[11], s.y, ws[13]));
          }
          const float fqx = (float)qdx, fqy = (float)qdy;
          unsigned bst, snd;
          float thr;
          float gbest[kBlockTargets / 4];
#pragma unroll
          for (int g = 0; g < kBlockTargets / 4; ++g) gbest[g] = kInfF32;
          tiles += (unsigned)sweep_row_pruned(t32, bb, nblk, v, fqx, fqy, tmax, kTruncRel, bst, snd, thr, gbest);
          // 3. exact resolution in float64
          int c = -1;
          bool amb = false;
          if (v) {
            const float sval = __uint_as_float(snd & 0xffffffc0u);
            if (!(sval > thr)) {
              amb = true;
            } else {
              const int blk = (int)(bst & 63u);
              // groups of four targets whose minimum (kept from the sweep) is within the slack: all
              // four members are re-evaluated in float64 -- a superset of the candidates the slack
              // admits, so the result is the float64 argmin (lowest index on exact ties).
              unsigned gmask = 0;
#pragma unroll
              for (int g = 0; g < kBlockTargets / 4; ++g)
                if (gbest[g] <= thr) gmask |= 1u << g;
              double bd = kInf;
              while (gmask) {
                const int g = __ffs(gmask) - 1;
                gmask &= gmask - 1;
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                  const int idx = blk * kBlockTargets + 4 * g + u;
                  if (idx < nt) {
                    const double2 t = t64[idx];
                    const double ex = qdx - t.x, ey = qdy - t.y;
                    const double d = ex * ex + ey * ey;
                    if (d < bd) { bd = d; c = idx; }
                  }
                }
              }
              if (c < 0) amb = true;  // cannot happen for finite data; be safe
            }
          }
          // 3b. rare: a second block is within the slack -> the warp scans all targets
          unsigned vote = __ballot_sync(0xffffffffu, amb);
          while (vote) {
            const int src_lane = __ffs(vote) - 1;
            vote &= vote - 1;
            const float fx = __shfl_sync(0xffffffffu, fqx, src_lane);
            const float fy = __shfl_sync(0xffffffffu, fqy, src_lane);
            const float thr_s = __shfl_sync(0xffffffffu, thr, src_lane);
            const double sdx = __shfl_sync(0xffffffffu, qdx, src_lane);
            const double sdy = __shfl_sync(0xffffffffu, qdy, src_lane);
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
            if (lane == src_lane) c = (bi == 0x7fffffff) ? 0 : bi;
          }
          // 4. float64 sums over the matched pairs
          if (v) {
            const double2 t = t64[c];
            const double wox = static_cast<const volatile double*>(wscr)[14];
            const double woy = static_cast<const volatile double*>(wscr)[15];
            const double ux = qdx - wox, uy = qdy - woy, vx = t.x - wox, vy = t.y - woy;
            acc[0] += ux; acc[1] += uy; acc[2] += vx; acc[3] += vy;
            acc[4] = fma(ux, vx, acc[4]); acc[5] = fma(ux, vy, acc[5]);
            acc[6] = fma(uy, vx, acc[6]); acc[7] = fma(uy, vy, acc[7]);
            const double ex = qdx - t.x, ey = qdy - t.y;
            acc[8] += ex * ex + ey * ey;
            if (p.nearest_only && p.dist2_out)
              p.dist2_out[(size_t)pair * p.n_max + j] = ex * ex + ey * ey;
          }
          if (normalize) {   // warp-uniform
            double2 t = make_double2(0.0, 0.0);
            if (v) t = t64[c];
            const double m0 = warp_min(v ? qdx : kInf), m1 = warp_max(v ? qdx : -kInf);
            const double m2 = warp_min(v ? qdy : kInf), m3 = warp_max(v ? qdy : -kInf);
            const double m4 = warp_min(v ? t.x : kInf), m5 = warp_max(v ? t.x : -kInf);
            const double m6 = warp_min(v ? t.y : kInf), m7 = warp_max(v ? t.y : -kInf);
            if (lane == 0) {
              wscr[0] = fmin(wscr[0], m0); wscr[1] = fmax(wscr[1], m1);
              wscr[2] = fmin(wscr[2], m2); wscr[3] = fmax(wscr[3], m3);
              wscr[4] = fmin(wscr[4], m4); wscr[5] = fmax(wscr[5], m5);
              wscr[6] = fmin(wscr[6], m6); wscr[7] = fmax(wscr[7], m7);
            }
          }
          // the row's match enters the register queue at the top: after the loop row r sits in slot R - nr + r
#pragma unroll
          for (int i = 0; i + 1 < R; ++i) corr[i] = corr[i + 1];
          corr[R - 1] = c;
        }
        if (normalize) {
          __syncwarp();
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[kSums2d + k] = wscr[k];
        }
        PHASE(1);
      } else {
      // 1+2. transform and fp32 sweep: first the shared rows (chunks of <= R rows, this warp's
      // share of the target blocks, results to shared memory), last this warp's own rows.
      for (int ph = 0; ph <= n_shared_chunks; ++ph) {
        const bool own = ph == n_shared_chunks;
        const int cnt = own ? base : min(R, extra - ph * R);
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int row = own ? row_of(r) : base * W + ph * R + r;
          const int j = row * 32 + lane;
          qx[r] = 0.f; qy[r] = 0.f;
          best[r] = 0xffffffffu; second[r] = 0xffffffffu;
          if (r < (own ? nr : cnt) && j < ns) {
            const double2 s = s64[j];
            qx[r] = (float)fma(r00, s.x, fma(r01, s.y, tx));
            qy[r] = (float)fma(r10, s.x, fma(r11, s.y, ty));
          }
        }
        const int blk_lo = own ? 0 : sblk_lo, blk_hi = own ? nblk : sblk_hi;
        switch (cnt) {
#define B200ICP_SWEEP_CASE(N) \
  case N:                     \
    if constexpr (N <= R) sweep2d<N, R>(t32, blk_lo, blk_hi, qx, qy, best, second); \
    break;
          B200ICP_SWEEP_CASE(1) B200ICP_SWEEP_CASE(2) B200ICP_SWEEP_CASE(3) B200ICP_SWEEP_CASE(4)
          B200ICP_SWEEP_CASE(5) B200ICP_SWEEP_CASE(6) B200ICP_SWEEP_CASE(7) B200ICP_SWEEP_CASE(8)
          B200ICP_SWEEP_CASE(9) B200ICP_SWEEP_CASE(10) B200ICP_SWEEP_CASE(11) B200ICP_SWEEP_CASE(12)
#undef B200ICP_SWEEP_CASE
          default: break;
        }
        if (!own) {
#pragma unroll
          for (int r = 0; r < R; ++r)
            if (r < cnt) part[((ph * R + r) * W + warp) * 32 + lane] = make_uint2(best[r], second[r]);
        }
        tiles += (unsigned)(cnt * (blk_hi - blk_lo));
      }
      }
      PHASE(1);
      if (!PRUNE && extra > 0) {
        __syncthreads();        // the shared rows' partial results are complete
        if (warp < extra) {     // merge the W partial top-2 lists of shared row `warp` into slot `base`
          unsigned b = 0xffffffffu, s2 = 0xffffffffu;
#pragma unroll
          for (int w = 0; w < W; ++w) {
            const uint2 o = part[(warp * W + w) * 32 + lane];
            s2 = min(min(s2, o.y), max(b, o.x));
            b = min(b, o.x);
          }
#pragma unroll
          for (int r = 0; r < R; ++r)
            if (r == base) { best[r] = b; second[r] = s2; }
        }
      }
      PHASE(2);
      if (nr > 0) {
        if constexpr (!PRUNE) {
        // 3. exact resolution in float64
        unsigned ambiguous = 0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int j = row_of(r) * 32 + lane;
          if (r < nr && j < ns) {
            const int blk = (int)(best[r] & 63u);
            const float bval = __uint_as_float(best[r] & 0xffffffc0u);
            const float sval = __uint_as_float(second[r] & 0xffffffc0u);
            const float a_bound = fmaxf(fabsf(qx[r]), fabsf(qy[r])) * (1.0f + 4.0f * kU) + tmax;
            const float thr = candidate_threshold(bval, a_bound, kTruncRel);
            if (!(sval > thr)) {
              ambiguous |= 1u << r;
            } else {
              // fp32 filter on groups of four targets (packed math, bit-identical to the sweep):
              // a group whose minimum is within the slack has all four members re-evaluated in
              // float64 -- a superset of the candidates the slack admits, so the result is the
              // same float64 argmin (lowest index on exact ties: groups and members ascend).
              unsigned gmask = 0;
              const ulonglong2* tb = reinterpret_cast<const ulonglong2*>(t32 + blk * (2 * kBlockTargets));
              const f32x2 qxx = pack2(qx[r], qx[r]), qyy = pack2(qy[r], qy[r]);
#pragma unroll
              for (int g = 0; g < kBlockTargets / 4; ++g) {
                const ulonglong2 X = tb[2 * g], Y = tb[2 * g + 1];
                f32x2 dx = sub2(qxx, X.x), dy = sub2(qyy, Y.x);
                const f32x2 d01 = fma2(dx, dx, mul2(dy, dy));
                dx = sub2(qxx, X.y); dy = sub2(qyy, Y.y);
                const f32x2 d23 = fma2(dx, dx, mul2(dy, dy));
                float d0, d1, d2, d3;
                unpack2(d01, d0, d1);
                unpack2(d23, d2, d3);
                if (fminf(fmin3(d0, d1, d2), d3) <= thr) gmask |= 1u << g;
              }
              const double2 s = s64[j];
              const double qdx = fma(r00, s.x, fma(r01, s.y, tx));
              const double qdy = fma(r10, s.x, fma(r11, s.y, ty));
              double bd = kInf;
              int bi = -1;
              while (gmask) {
                const int g = __ffs(gmask) - 1;
                gmask &= gmask - 1;
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                  const int idx = blk * kBlockTargets + 4 * g + u;
                  if (idx < nt) {
                    const double2 t = t64[idx];
                    const double ex = qdx - t.x, ey = qdy - t.y;
                    const double d = ex * ex + ey * ey;
                    if (d < bd) { bd = d; bi = idx; }
                  }
                }
              }
              if (bi < 0) ambiguous |= 1u << r;  // cannot happen for finite data; be safe
              corr[r] = bi;
            }
          }
        }
        // 3b. rare: a second block is within the slack -> the warp scans all targets
#pragma unroll
        for (int r = 0; r < R; ++r) {
          unsigned vote = __ballot_sync(0xffffffffu, (ambiguous >> r) & 1u);
          while (vote) {
            const int src_lane = __ffs(vote) - 1;
            vote &= vote - 1;
            const float fx = __shfl_sync(0xffffffffu, qx[r], src_lane);
            const float fy = __shfl_sync(0xffffffffu, qy[r], src_lane);
            const unsigned bpk = __shfl_sync(0xffffffffu, best[r], src_lane);
            const float a_bound = fmaxf(fabsf(fx), fabsf(fy)) * (1.0f + 4.0f * kU) + tmax;
            const float thr = candidate_threshold(__uint_as_float(bpk & 0xffffffc0u), a_bound, kTruncRel);
            const int j = row_of(r) * 32 + src_lane;
            const double2 s = s64[j];
            const double qdx = fma(r00, s.x, fma(r01, s.y, tx));
            const double qdy = fma(r10, s.x, fma(r11, s.y, ty));
            double bd = kInf;
            int bi = 0x7fffffff;
            for (int i = lane; i < nt; i += 32) {
              const float2 t = target32(t32, i);
              // !(d > thr) also admits NaN so that non-finite data still yields an index
              if (!(dist2_2d(fx, fy, t.x, t.y) > thr)) {
                const double2 td = t64[i];
                const double ex = qdx - td.x, ey = qdy - td.y;
                const double d = ex * ex + ey * ey;
                if (d < bd || bi == 0x7fffffff) { bd = d; bi = i; }
              }
            }
            warp_argmin(bd, bi);
            if (lane == src_lane) corr[r] = (bi == 0x7fffffff) ? 0 : bi;
          }
        }
        PHASE(3);
        // 4. float64 sums over this thread's matched pairs
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int j = row_of(r) * 32 + lane;
          if (r < nr && j < ns) {
            const double2 s = s64[j];
            const double qdx = fma(r00, s.x, fma(r01, s.y, tx));
            const double qdy = fma(r10, s.x, fma(r11, s.y, ty));
            const double2 t = t64[corr[r]];
            const double ux = qdx - ox, uy = qdy - oy, vx = t.x - ox, vy = t.y - oy;
            acc[0] += ux; acc[1] += uy; acc[2] += vx; acc[3] += vy;
            acc[4] = fma(ux, vx, acc[4]); acc[5] = fma(ux, vy, acc[5]);
            acc[6] = fma(uy, vx, acc[6]); acc[7] = fma(uy, vy, acc[7]);
            const double ex = qdx - t.x, ey = qdy - t.y;
            acc[8] += ex * ex + ey * ey;
            if (normalize) {
              acc[9] = fmin(acc[9], qdx);   acc[10] = fmax(acc[10], qdx);
              acc[11] = fmin(acc[11], qdy); acc[12] = fmax(acc[12], qdy);
              acc[13] = fmin(acc[13], t.x); acc[14] = fmax(acc[14], t.x);
              acc[15] = fmin(acc[15], t.y); acc[16] = fmax(acc[16], t.y);
            }
            if (p.nearest_only && p.dist2_out)
              p.dist2_out[(size_t)pair * p.n_max + j] = ex * ex + ey * ey;
          }
        }
        }  // !PRUNE
#pragma unroll
        for (int k = 0; k < kSums2d; ++k) acc[k] = warp_sum(acc[k]);
        if (!PRUNE && normalize) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            acc[kSums2d + 2 * k] = warp_min(acc[kSums2d + 2 * k]);
            acc[kSums2d + 2 * k + 1] = warp_max(acc[kSums2d + 2 * k + 1]);
          }
        }
      }
      PHASE(4);
      // exchange: one shared-memory round per pass, double-buffered by pass parity
      double* slot = red + (size_t)(passes & 1) * W * kRed2d;
      if (lane == 0) {
#pragma unroll
        for (int k = 0; k < kSums2d; ++k) slot[warp * kRed2d + k] = acc[k];
        if (normalize) {
#pragma unroll
          for (int k = kSums2d; k < kRed2d; ++k) slot[warp * kRed2d + k] = acc[k];
        }
      }
      __syncthreads();
      // Column k of the W partial rows is reduced by lane k (same order as a sequential sum over
      // the warps), then every lane fetches the totals it needs by shuffle.
      double tot[kRed2d];
      {
        const int kcol = lane < kRed2d ? lane : kRed2d - 1;
        double col = slot[kcol];
#pragma unroll
        for (int w = 1; w < W; ++w) {
          const double o = slot[w * kRed2d + kcol];
          col = kcol < kSums2d ? col + o : ((kcol & 1) ? fmin(col, o) : fmax(col, o));   // 9, 11, ..: minima
        }
#pragma unroll
        for (int k = 0; k < kSums2d; ++k) tot[k] = __shfl_sync(0xffffffffu, col, k);
#pragma unroll
        for (int k = kSums2d; k < kRed2d; ++k) tot[k] = normalize ? __shfl_sync(0xffffffffu, col, k) : 0.0;
      }
      passes += 1;

      PHASE(5);
      // 5. error of this pass (before the update), solve, compose, stop tests
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
        if constexpr (PRUNE) {
          const volatile double* ws = wscr;
          r00 = ws[8]; r01 = ws[9]; r10 = ws[10]; r11 = ws[11]; tx = ws[12]; ty = ws[13];
          ox = ws[14]; oy = ws[15];
        }
        const double qbx = ox + mqx, qby = oy + mqy, pbx = ox + mpx, pby = oy + mpy;
        const double dtx = pbx - (d00 * qbx + d01 * qby);
        const double dty = pby - (d10 * qbx + d11 * qby);
        const double n00 = d00 * r00 + d01 * r10, n01 = d00 * r01 + d01 * r11;
        const double n10 = d10 * r00 + d11 * r10, n11 = d10 * r01 + d11 * r11;
        const double ntx = d00 * tx + d01 * ty + dtx, nty = d10 * tx + d11 * ty + dty;
        r00 = n00; r01 = n01; r10 = n10; r11 = n11; tx = ntx; ty = nty; zs *= dz;
        if constexpr (PRUNE) {
          __syncwarp();            // every lane has read the old pose
          if (lane == 0) {
            wscr[8] = r00; wscr[9] = r01; wscr[10] = r10; wscr[11] = r11; wscr[12] = tx; wscr[13] = ty;
          }
          __syncwarp();
        }
        status = stop_test(st, err, epsilon, p.max_iters, p.stopping_thresh);
      }
      PHASE(6);
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
          int* out = p.corr_out + (size_t)pair * p.n_max;
#pragma unroll
          for (int r = 0; r < R; ++r) {
            const int rr = PRUNE ? r - (R - nr) : r;       // pruned variant: register queue, see above
            const int j = row_of(rr) * 32 + lane;
            if (rr >= 0 && rr < nr && j < ns) out[j] = corr[r];
          }
          for (int j = ns + tid; j < p.n_max; j += W * 32) out[j] = -1;
        }
        break;
      }
    }
  }
  PHASE_FLUSH;
}

}  // namespace b200icp
