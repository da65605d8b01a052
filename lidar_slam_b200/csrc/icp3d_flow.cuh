// General ICP (3-D clouds, planar scans too large for the 2-D kernel, non-planar inits) as a
// DATAFLOW over the whole GPU.  Replaces modules/icp.py:123-189 for the call shape of
// code/icp_warm_up/test_icp.py:78-88 (3-D clouds, 24 yaw initialisations per target).
//
// Round 1 gave every pair a thread-block cluster that swept every candidate; its profile showed
// 28 % of the time at the per-pass cluster barrier, SMs 70 % active (96 problems with 2..26
// passes), the fp32 target copy rebuilt per pair although 24 pairs share a target, and no pruning.
// This version has three kernels per launch:
//
//   flow_setup     (1 CTA)  finds the distinct scans the pair list references, initialises the
//                           per-pair state and queues pass 0 of every pair.
//   flow_prepare   (1 CTA per distinct scan)  sorts the scan along a Morton curve (stable LSD
//                           radix sort, deterministic) and writes, ONCE per scan however many
//                           pairs use it: the sorted float64 points with their original index,
//                           the fp32 copy in [x4 y4 z4] groups, the bounding box of every
//                           32-point block, max |coordinate| and the centroid.
//   icp3d_flow     persistent CTAs pulling work items (pair, chunk of 512 source points) from a
//                           ring in global memory.  A warp takes a row of 64 Morton-consecutive
//                           source points (two per lane), bounds every target block against the
//                           row box with the candidate's own monotone arithmetic (exact; see
//                           icp2d.cuh), sweeps the blocks that overlap the row, then those within
//                           the largest candidate threshold -- each one only if at least one lane's
//                           own point-to-box bound admits it -- reading target blocks straight
//                           from L1/L2 (they are shared by every pair on that target).  Exact
//                           float64 resolution as in the planar kernel (lowest ORIGINAL index on
//                           exact ties).  The chunk's partial sums go to global memory; the CTA
//                           that retires the last chunk of a pass adds the partials in chunk
//                           order, solves the 3x3 Kabsch problem (kabsch3.cuh), applies the
//                           reference's stop tests and queues the next pass of the pair.
//
// No cluster barriers, no idle SMs while any pair has work, and the summation order is a
// function of the pair alone (chunks of 512 sorted points), not of the batch shape: results are
// batch-invariant.
#pragma once
#include "icp2d.cuh"
#include "kabsch3.cuh"

namespace b200icp {

constexpr int kFlowWarps = 8;                      // warps per CTA = rows per work item
#ifndef B200ICP_FLOW_Q
#define B200ICP_FLOW_Q 2
#endif
#ifndef B200ICP_FLOW_MINB
#define B200ICP_FLOW_MINB 2
#endif
constexpr int kFlowQ = B200ICP_FLOW_Q;             // source points per lane and row
constexpr int kFlowMinBlocks = B200ICP_FLOW_MINB;  // resident CTAs per SM the registers are bounded for
constexpr int kFlowRowQ = 32 * kFlowQ;             // source points per row
constexpr int kFlowChunkPts = kFlowWarps * kFlowRowQ;
constexpr int kFlowMaxPairs = 1024;                // pairs per launch; longer lists are cut by the host side
constexpr int kSums3d = 16;                        // sum u(3) v(3) u v^T(9) sse
constexpr int kRed3d = 22;                         // + min/max per axis over moved source and matched target
constexpr unsigned kRingEmpty = 0xffffffffu;
constexpr int kPrepThreads = 512;
constexpr unsigned long long kFlowSpinLimitNs = 4000000000ull;

// control words (uint32) at the start of the workspace
enum { kCtlHead = 0, kCtlTail = 1, kCtlTilesLo = 2, kCtlTilesHi = 3, kCtlPairsDone = 4, kCtlAbort = 5 };

struct __align__(16) FlowPair {
  double T[12];          // rows 0..2 of the current transform (mutable)
  double bottom[4];      // its 4th row as the caller passed it
  double prev_err;       // stop-test state (mutable)
  double epsilon;
  int have_prev, count, passes;   // (mutable)
  unsigned done;         // chunks of the current pass retired (mutable)
  int ns, nt, n_chunks, src_slot, tgt_slot, normalize;
  int pad[2];
};
static_assert(sizeof(FlowPair) == 192, "FlowPair layout");

struct FlowSlotLayout { size_t off_f64, off_f32, off_box, off_perm, bytes; int np32, nblk_pad; };
__host__ __device__ inline FlowSlotLayout flow_slot_layout(int n_max) {
  FlowSlotLayout s;
  s.np32 = ((n_max + 31) / 32) * 32;
  s.nblk_pad = ((s.np32 / 32 + 31) / 32) * 32;
  size_t o = 256;                                          // header: n, nblk, tmax, origin
  s.off_f64 = o; o += (size_t)s.np32 * 32;                 // double4 (x, y, z, original index); sort scratch before
  s.off_f32 = o; o += (size_t)s.np32 * 12;                 // [x4 y4 z4] groups
  s.off_box = o; o += (size_t)s.nblk_pad * 32;             // per 32-point block: (lo.xyz, 0), (hi.xyz, 0)
  s.off_perm = o; o += (size_t)s.np32 * 4;                 // sorted position -> original index
  s.bytes = (o + 255) / 256 * 256;
  return s;
}
struct FlowSlotHeader { int n, nblk; float tmax; int pad; double ox, oy, oz; };

struct FlowLayout {
  size_t off_first, off_pairs, off_ring, off_part, off_slots, total;
  int ring_cap, max_chunks, n_pairs;
  FlowSlotLayout slot;
};
__host__ __device__ inline FlowLayout flow_layout(int n_pairs, int n_max) {
  FlowLayout L;
  L.n_pairs = n_pairs;
  L.max_chunks = (n_max + kFlowChunkPts - 1) / kFlowChunkPts;
  long long want = 2LL * n_pairs * L.max_chunks;
  int cap = 64;
  while (cap < want) cap <<= 1;
  L.ring_cap = cap;
  L.slot = flow_slot_layout(n_max);
  size_t o = 256;
  L.off_first = o; o += ((size_t)2 * n_pairs * 4 + 255) / 256 * 256;
  L.off_pairs = o; o += ((size_t)n_pairs * sizeof(FlowPair) + 255) / 256 * 256;
  L.off_ring = o; o += (size_t)cap * 4;
  L.off_part = o; o += ((size_t)n_pairs * L.max_chunks * kRed3d * 8 + 255) / 256 * 256;
  L.off_slots = o; o += (size_t)2 * n_pairs * L.slot.bytes;
  L.total = o;
  return L;
}

__device__ __forceinline__ void load_point(const double* base, int dim, int i, double& x, double& y, double& z) {
  if (dim == 3) {
    x = base[3 * (size_t)i]; y = base[3 * (size_t)i + 1]; z = base[3 * (size_t)i + 2];
  } else {
    const double2 v = reinterpret_cast<const double2*>(base)[i];
    x = v.x; y = v.y; z = 0.0;
  }
}

__device__ __forceinline__ unsigned ld_volatile_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// ---------------------------------------------------------------------------------------------
// setup: distinct scans, per-pair state, pass 0 of every pair queued
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) flow_setup_kernel(const RunParams p, const int dim, const FlowLayout L, char* ws) {
  unsigned* ctl = reinterpret_cast<unsigned*>(ws);
  int* first = reinterpret_cast<int*>(ws + L.off_first);
  FlowPair* ps = reinterpret_cast<FlowPair*>(ws + L.off_pairs);
  unsigned* ring = reinterpret_cast<unsigned*>(ws + L.off_ring);
  __shared__ int s_scan[32];
  __shared__ int s_early;
  const int tid = threadIdx.x, B = p.n_pairs;
  const double kNaN = __longlong_as_double(0x7ff8000000000000LL);
  if (tid == 0) s_early = 0;
  for (int i = tid; i < L.ring_cap; i += blockDim.x) ring[i] = kRingEmpty;
  // first reference of every referenced scan (2b: source of pair b, 2b + 1: its target)
  for (int k = tid; k < 2 * B; k += blockDim.x) {
    const int sc = p.pairs[k];
    int f = -1;
    if (sc >= 0 && sc < p.n_scans && p.counts[sc] > 0) {
      f = k;
      for (int j = 0; j < k; ++j)
        if (p.pairs[j] == sc) { f = j; break; }
    }
    first[k] = f;
  }
  __syncthreads();
  int my_chunks = 0;
  for (int b0 = 0; b0 < B; b0 += blockDim.x) {      // B <= kFlowMaxPairs = blockDim.x: one round
    const int b = b0 + tid;
    int chunks = 0;
    if (b < B) {
      const int si = p.pairs[2 * b], ti = p.pairs[2 * b + 1];
      int early = -1;
      if (si < 0 || ti < 0 || si >= p.n_scans || ti >= p.n_scans) early = B200ICP_STATUS_SKIPPED;
      else if (p.counts[si] <= 0 || p.counts[ti] <= 0) early = B200ICP_STATUS_EMPTY;
      FlowPair& P = ps[b];
      if (early >= 0) {
        if (p.status_out) p.status_out[b] = early;
        if (p.iters_out) p.iters_out[b] = 0;
        if (p.err_out) p.err_out[b] = kNaN;
        if (p.T_out)
          for (int k = 0; k < 16; ++k) p.T_out[(size_t)b * 16 + k] = p.init ? p.init[(size_t)b * 16 + k] : 0.0;
        if (p.init) {
          const double* T0 = p.init + (size_t)b * 16;
          store_pose_to_peers(p, b, T0[3], T0[7], T0[0], T0[4], kNaN);
        }
        if (p.corr_out)
          for (int j = 0; j < p.n_max; ++j) p.corr_out[(size_t)b * p.n_max + j] = -1;
        P.n_chunks = 0;
        atomicAdd(&s_early, 1);
      } else {
        for (int k = 0; k < 12; ++k) P.T[k] = p.nearest_only ? ((k % 5 == 0) ? 1.0 : 0.0) : p.init[(size_t)b * 16 + k];
        for (int k = 0; k < 4; ++k) P.bottom[k] = p.nearest_only ? (k == 3 ? 1.0 : 0.0) : p.init[(size_t)b * 16 + 12 + k];
        P.prev_err = 0.0; P.have_prev = 0; P.count = 0; P.passes = 0; P.done = 0u;
        P.epsilon = p.nearest_only ? 0.0 : p.epsilon[b];
        P.normalize = p.nearest_only ? 0 : (p.normalize[b] != 0);
        P.ns = min(p.counts[si], p.n_max);
        P.nt = min(p.counts[ti], p.n_max);
        P.src_slot = first[2 * b];
        P.tgt_slot = first[2 * b + 1];
        chunks = (P.ns + kFlowChunkPts - 1) / kFlowChunkPts;
        P.n_chunks = chunks;
      }
    }
    my_chunks = chunks;
  }
  // exclusive prefix of the chunk counts over the pairs -> ring positions of pass 0
  const int lane = tid & 31, warp = tid >> 5;
  int incl = my_chunks;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += up;
  }
  if (lane == 31) s_scan[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int w = s_scan[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int up = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += up;
    }
    s_scan[lane] = w;     // inclusive over warps
  }
  __syncthreads();
  const int base = (warp ? s_scan[warp - 1] : 0) + incl - my_chunks;
  if (tid < B)
    for (int c = 0; c < my_chunks; ++c) ring[(base + c) & (L.ring_cap - 1)] = ((unsigned)tid << 21) | (unsigned)c;
  if (tid == 0) {
    ctl[kCtlHead] = 0u;
    ctl[kCtlTail] = (unsigned)s_scan[31];
    ctl[kCtlPairsDone] = (unsigned)s_early;
    if (p.n_peers > 0 && p.flag_off != 0 && s_early == B) raise_arrival_flags(p);   // nothing to run
  }
}

// ---------------------------------------------------------------------------------------------
// prepare: one CTA per distinct scan
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned expand10(unsigned v) {   // 10 bits -> every third bit
  v = (v | (v << 16)) & 0x030000FFu;
  v = (v | (v << 8)) & 0x0300F00Fu;
  v = (v | (v << 4)) & 0x030C30C3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}

__device__ __forceinline__ double block_reduce(double v, int op, double* sh) {   // op 0 sum, 1 min, 2 max; result to all
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double u = __shfl_xor_sync(0xffffffffu, v, o);
    v = op == 0 ? v + u : (op == 1 ? fmin(v, u) : fmax(v, u));
  }
  __syncthreads();
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  double r = sh[0];
  for (int w = 1; w < nw; ++w) r = op == 0 ? r + sh[w] : (op == 1 ? fmin(r, sh[w]) : fmax(r, sh[w]));
  return r;
}

__global__ void __launch_bounds__(kPrepThreads) flow_prepare_kernel(const RunParams p, const int dim, const FlowLayout L, char* ws) {
  extern __shared__ unsigned s_cnt[];          // [16][kPrepThreads] digit counters, then offsets
  __shared__ double s_red[kPrepThreads / 32];
  __shared__ unsigned s_wtot[16 * (kPrepThreads / 32)];
  const int k = blockIdx.x;
  const int* first = reinterpret_cast<const int*>(ws + L.off_first);
  if (first[k] != k) return;                   // not the first reference of a valid scan
  const int sc = p.pairs[k];
  const int n = min(p.counts[sc], p.n_max);
  const double* pts = p.scans + (size_t)sc * p.n_max * dim;
  char* slot = ws + L.off_slots + (size_t)k * L.slot.bytes;
  FlowSlotHeader* hdr = reinterpret_cast<FlowSlotHeader*>(slot);
  double4* f64 = reinterpret_cast<double4*>(slot + L.slot.off_f64);
  float* f32 = reinterpret_cast<float*>(slot + L.slot.off_f32);
  float4* box = reinterpret_cast<float4*>(slot + L.slot.off_box);
  int* perm = reinterpret_cast<int*>(slot + L.slot.off_perm);
  uint2* bufA = reinterpret_cast<uint2*>(slot + L.slot.off_f64);           // (key, index) ping-pong inside the f64 region
  uint2* bufB = bufA + L.slot.np32;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double kInf = __longlong_as_double(0x7ff0000000000000LL);

  // statistics: bounding box (for the Morton grid), max |coord| in fp32, centroid
  double lo[3] = {kInf, kInf, kInf}, hi[3] = {-kInf, -kInf, -kInf}, sum[3] = {0.0, 0.0, 0.0};
  float amax = 0.f;
  for (int i = tid; i < n; i += kPrepThreads) {
    double x, y, z;
    load_point(pts, dim, i, x, y, z);
    lo[0] = fmin(lo[0], x); hi[0] = fmax(hi[0], x); sum[0] += x;
    lo[1] = fmin(lo[1], y); hi[1] = fmax(hi[1], y); sum[1] += y;
    lo[2] = fmin(lo[2], z); hi[2] = fmax(hi[2], z); sum[2] += z;
    amax = fmaxf(amax, fmaxf(fabsf((float)x), fmaxf(fabsf((float)y), fabsf((float)z))));
  }
  double inv[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    lo[a] = block_reduce(lo[a], 1, s_red);
    hi[a] = block_reduce(hi[a], 2, s_red);
    sum[a] = block_reduce(sum[a], 0, s_red);
    const double ext = hi[a] - lo[a];
    inv[a] = (ext > 0.0 && ext < kInf) ? 1024.0 / ext : 0.0;
  }
  const double amax_d = block_reduce((double)amax, 2, s_red);
  if (tid == 0) {
    hdr->n = n; hdr->nblk = (n + 31) >> 5;
    hdr->tmax = (float)amax_d * (1.0f + 4.0f * kU);
    hdr->ox = sum[0] / (double)n; hdr->oy = sum[1] / (double)n; hdr->oz = sum[2] / (double)n;
  }
  // Morton keys
  for (int i = tid; i < n; i += kPrepThreads) {
    double x, y, z;
    load_point(pts, dim, i, x, y, z);
    const int gx = min(max((int)((x - lo[0]) * inv[0]), 0), 1023);
    const int gy = min(max((int)((y - lo[1]) * inv[1]), 0), 1023);
    const int gz = min(max((int)((z - lo[2]) * inv[2]), 0), 1023);
    bufA[i] = make_uint2(expand10((unsigned)gx) | (expand10((unsigned)gy) << 1) | (expand10((unsigned)gz) << 2), (unsigned)i);
  }
  __syncthreads();
  // stable LSD radix sort, 4 bits per pass: thread t owns the contiguous items [t*ipt, (t+1)*ipt)
  const int ipt = (n + kPrepThreads - 1) / kPrepThreads;
  const int i_lo = min(tid * ipt, n), i_hi = min(i_lo + ipt, n);
  uint2* src = bufA;
  uint2* dst = bufB;
  for (int shift = 0; shift < 30; shift += 4) {
#pragma unroll
    for (int d = 0; d < 16; ++d) s_cnt[d * kPrepThreads + tid] = 0u;
    for (int i = i_lo; i < i_hi; ++i) s_cnt[((src[i].x >> shift) & 15u) * kPrepThreads + tid] += 1u;
    // exclusive scan in (digit, thread) order: warp scans per digit, then the 16 x 16 warp totals
    unsigned cnt[16], inc[16];
#pragma unroll
    for (int d = 0; d < 16; ++d) {
      cnt[d] = s_cnt[d * kPrepThreads + tid];
      unsigned v = cnt[d];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned up = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += up;
      }
      inc[d] = v;
      if (lane == 31) s_wtot[d * (kPrepThreads / 32) + warp] = v;
    }
    __syncthreads();
    if (warp == 0) {        // 256 totals in (digit, warp) order, 8 per lane
      unsigned loc[8], run = 0u;
#pragma unroll
      for (int e = 0; e < 8; ++e) { loc[e] = s_wtot[lane * 8 + e]; run += loc[e]; }
      unsigned v = run;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned up = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += up;
      }
      unsigned excl = v - run;
#pragma unroll
      for (int e = 0; e < 8; ++e) { s_wtot[lane * 8 + e] = excl; excl += loc[e]; }
    }
    __syncthreads();
#pragma unroll
    for (int d = 0; d < 16; ++d)
      s_cnt[d * kPrepThreads + tid] = s_wtot[d * (kPrepThreads / 32) + warp] + inc[d] - cnt[d];
    for (int i = i_lo; i < i_hi; ++i) {
      const uint2 e = src[i];
      const unsigned at = s_cnt[((e.x >> shift) & 15u) * kPrepThreads + tid]++;
      dst[at] = e;
    }
    __syncthreads();
    uint2* t = src; src = dst; dst = t;
  }
  // the permutation leaves the scratch (which the float64 copy is about to overwrite)
  for (int i = tid; i < L.slot.np32; i += kPrepThreads) perm[i] = i < n ? (int)src[i].y : 0x7fffffff;
  __syncthreads();
  // sorted copies (+ sentinels up to the next multiple of 32), block boxes
  for (int i = tid; i < L.slot.np32; i += kPrepThreads) {
    double x = 1.0e300, y = 1.0e300, z = 1.0e300;       // float64 sentinel: distance +inf
    float fx = kBigCoord, fy = kBigCoord, fz = kBigCoord;
    const int oi = perm[i];
    if (i < n) {
      load_point(pts, dim, oi, x, y, z);
      fx = (float)x; fy = (float)y; fz = (float)z;
    }
    f64[i] = make_double4(x, y, z, __longlong_as_double((long long)oi));
    float* g = f32 + (size_t)(i >> 2) * 12 + (i & 3);
    g[0] = fx; g[4] = fy; g[8] = fz;
  }
  __syncthreads();
  const float kInfF = __int_as_float(0x7f800000);
  const int nblk = (n + 31) >> 5;
  for (int b = warp; b < L.slot.nblk_pad; b += kPrepThreads / 32) {
    float4 blo = make_float4(kInfF, kInfF, kInfF, 0.f), bhi = make_float4(kInfF, kInfF, kInfF, 0.f);
    if (b < nblk) {
      const int i = b * 32 + lane;
      const bool real = i < n;
      const float* g = f32 + (size_t)(i >> 2) * 12 + (i & 3);
      const float x = g[0], y = g[4], z = g[8];
      blo = make_float4(redux_minf(real ? x : kInfF), redux_minf(real ? y : kInfF), redux_minf(real ? z : kInfF), 0.f);
      bhi = make_float4(redux_maxf(real ? x : -kInfF), redux_maxf(real ? y : -kInfF), redux_maxf(real ? z : -kInfF), 0.f);
    }
    if (lane == 0) { box[2 * b] = blo; box[2 * b + 1] = bhi; }
  }
}

// ---------------------------------------------------------------------------------------------
// the dataflow ICP kernel
// ---------------------------------------------------------------------------------------------
struct FlowTarget {
  const float* t32;
  const double4* t64;
  const float4* box;
  int nt, nblk;
  float tmax;
};

// lower bound of the fp32 candidate distance between a box of queries (qlo, qhi) -- or a single
// query when qlo == qhi -- and every target of a block box; same operations, same rounding, same
// order as dist2_3d, so it never exceeds a computed candidate distance (monotone rounding)
__device__ __forceinline__ float box_lb3(float qlx, float qly, float qlz, float qhx, float qhy, float qhz,
                                         const float4 blo, const float4 bhi) {
  const float ex = fmaxf(fmaxf(__fsub_rn(blo.x, qhx), __fsub_rn(qlx, bhi.x)), 0.f);
  const float ey = fmaxf(fmaxf(__fsub_rn(blo.y, qhy), __fsub_rn(qly, bhi.y)), 0.f);
  const float ez = fmaxf(fmaxf(__fsub_rn(blo.z, qhz), __fsub_rn(qlz, bhi.z)), 0.f);
  return __fmaf_rn(ex, ex, __fmaf_rn(ey, ey, __fmul_rn(ez, ez)));
}

// One 32-target block against the two queries of every lane; targets come from L1/L2 (uniform
// addresses: every load is a warp-wide broadcast).  Tracks, per query, the best block (value, id,
// minima of its eight groups of four) and the smallest minimum of any OTHER swept block.
__device__ __forceinline__ bool flow_sweep_block(const float* t32, int b, const f32x2 (&qxx)[kFlowQ], const f32x2 (&qyy)[kFlowQ],
                                                 const f32x2 (&qzz)[kFlowQ], float (&bestv)[kFlowQ], float (&secondv)[kFlowQ],
                                                 int (&bestb)[kFlowQ], float (&gbest)[kFlowQ][8]) {
  const ulonglong2* tb = reinterpret_cast<const ulonglong2*>(t32 + (size_t)b * 96);
  float gm[kFlowQ][8];
#pragma unroll
  for (int g = 0; g < 8; ++g) {
    const ulonglong2 X = __ldg(tb + 3 * g), Y = __ldg(tb + 3 * g + 1), Z = __ldg(tb + 3 * g + 2);
#pragma unroll
    for (int q = 0; q < kFlowQ; ++q) {
      f32x2 dx = sub2(qxx[q], X.x), dy = sub2(qyy[q], Y.x), dz = sub2(qzz[q], Z.x);
      const f32x2 d01 = fma2(dx, dx, fma2(dy, dy, mul2(dz, dz)));
      dx = sub2(qxx[q], X.y); dy = sub2(qyy[q], Y.y); dz = sub2(qzz[q], Z.y);
      const f32x2 d23 = fma2(dx, dx, fma2(dy, dy, mul2(dz, dz)));
      float d0, d1, d2, d3;
      unpack2(d01, d0, d1);
      unpack2(d23, d2, d3);
      gm[q][g] = fminf(fmin3(d0, d1, d2), d3);
    }
  }
  bool improved = false;
#pragma unroll
  for (int q = 0; q < kFlowQ; ++q) {
    const float bm = fminf(fmin3(fmin3(fmin3(gm[q][0], gm[q][1], gm[q][2]), gm[q][3], gm[q][4]), gm[q][5], gm[q][6]),
                           gm[q][7]);
    if (bm < bestv[q]) {        // strict: on equal minima the earlier (lower) block stays the best one
      secondv[q] = bestv[q];
      bestv[q] = bm;
      bestb[q] = b;
#pragma unroll
      for (int g = 0; g < 8; ++g) gbest[q][g] = gm[q][g];
      improved = true;
    } else {
      secondv[q] = fminf(secondv[q], bm);
    }
  }
  return improved;
}

// min / max of a double over the warp with two 32-bit REDUX steps on its order-preserving key
__device__ __forceinline__ double warp_max_f64(double v) {
  long long b = __double_as_longlong(v);
  unsigned long long k = b < 0 ? ~(unsigned long long)b : ((unsigned long long)b | 0x8000000000000000ull);
  const unsigned hi = (unsigned)(k >> 32), lo = (unsigned)k;
  const unsigned mh = __reduce_max_sync(0xffffffffu, hi);
  const unsigned ml = __reduce_max_sync(0xffffffffu, hi == mh ? lo : 0u);
  k = ((unsigned long long)mh << 32) | ml;
  b = (k & 0x8000000000000000ull) ? (long long)(k & 0x7fffffffffffffffull) : (long long)~k;
  return __longlong_as_double(b);
}
__device__ __forceinline__ double warp_min_f64(double v) { return -warp_max_f64(-v); }

#ifdef B200ICP_FLOW_DEBUG
#define FLOW_DBG(code) do { if (p.dist2_out && !p.nearest_only && threadIdx.x == 0) { \
    volatile unsigned* dbg_ = reinterpret_cast<volatile unsigned*>(p.dist2_out); dbg_[64 + blockIdx.x] = (code); } } while (0)
#define FLOW_DBG_W0(code) do { if (p.dist2_out && !p.nearest_only && threadIdx.x == 0) { \
    volatile unsigned* dbg_ = reinterpret_cast<volatile unsigned*>(p.dist2_out); dbg_[64 + blockIdx.x] = (code); } } while (0)
#define FLOW_CNT(i) do { if (p.dist2_out && !p.nearest_only && threadIdx.x == 0) \
    atomicAdd_system(reinterpret_cast<unsigned*>(p.dist2_out) + (i), 1u); } while (0)
#else
#define FLOW_DBG(code) do { } while (0)
#define FLOW_DBG_W0(code) do { } while (0)
#define FLOW_CNT(i) do { } while (0)
#endif

template <bool PRUNE>
__global__ void __launch_bounds__(kFlowWarps * 32, kFlowMinBlocks) icp3d_flow_kernel(const RunParams p, const FlowLayout L, char* ws) {
  __shared__ unsigned s_item;
  __shared__ double s_part[kFlowWarps][kRed3d];
  unsigned* ctl = reinterpret_cast<unsigned*>(ws);
  FlowPair* ps = reinterpret_cast<FlowPair*>(ws + L.off_pairs);
  unsigned* ring = reinterpret_cast<unsigned*>(ws + L.off_ring);
  double* part = reinterpret_cast<double*>(ws + L.off_part);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned full = 0xffffffffu;
  const float kInfF = __int_as_float(0x7f800000);
  const double kInf = __longlong_as_double(0x7ff0000000000000LL);
  unsigned tiles = 0;

  for (;;) {
    // ---- pop a work item ----------------------------------------------------------------------
    FLOW_DBG(1);
    if (tid == 0) {
      unsigned item = kRingEmpty;
      const unsigned ticket = atomicAdd(&ctl[kCtlHead], 1u);
      unsigned* slot = ring + (ticket & (unsigned)(L.ring_cap - 1));
      unsigned long long t0 = 0;
      for (unsigned spins = 0;; ++spins) {
        const unsigned v = ld_volatile_u32(slot);
        if (v != kRingEmpty && atomicCAS(slot, v, kRingEmpty) == v) { item = v; break; }
        if (ld_volatile_u32(&ctl[kCtlPairsDone]) >= (unsigned)L.n_pairs || ld_volatile_u32(&ctl[kCtlAbort]) != 0u) break;
        if (spins > 16u) {
          __nanosleep(200);
          if (t0 == 0) t0 = global_timer_ns();
          else if (global_timer_ns() - t0 > kFlowSpinLimitNs) { atomicExch(&ctl[kCtlAbort], 1u); break; }
        }
      }
      __threadfence();      // acquire: the state the producer published before the item
      s_item = item;
    }
    __syncthreads();
    const unsigned item = s_item;
    if (item == kRingEmpty) break;
    FLOW_DBG(2); FLOW_CNT(0);
    const int pair = (int)(item >> 21), chunk = (int)(item & 0x1fffffu);
    FlowPair* P = ps + pair;
    const int ns = P->ns, n_chunks = P->n_chunks;
    const bool normalize = P->normalize != 0;
    double T[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) T[k] = __ldcg(&P->T[k]);
    const char* sslot = ws + L.off_slots + (size_t)P->src_slot * L.slot.bytes;
    const char* tslot = ws + L.off_slots + (size_t)P->tgt_slot * L.slot.bytes;
    const double4* src64 = reinterpret_cast<const double4*>(sslot + L.slot.off_f64);
    const FlowSlotHeader* th = reinterpret_cast<const FlowSlotHeader*>(tslot);
    FlowTarget tg;
    tg.t32 = reinterpret_cast<const float*>(tslot + L.slot.off_f32);
    tg.t64 = reinterpret_cast<const double4*>(tslot + L.slot.off_f64);
    tg.box = reinterpret_cast<const float4*>(tslot + L.slot.off_box);
    tg.nt = th->n; tg.nblk = th->nblk; tg.tmax = th->tmax;
    const double ox = th->ox, oy = th->oy, oz = th->oz;

    double acc[kRed3d];
#pragma unroll
    for (int k = 0; k < kSums3d; ++k) acc[k] = 0.0;
#pragma unroll
    for (int k = 0; k < 3; ++k) { acc[kSums3d + 2 * k] = kInf; acc[kSums3d + 2 * k + 1] = -kInf; }

    const int row = chunk * kFlowWarps + warp;
    if (row * kFlowRowQ < ns) {
      // ---- this warp's row: two queries per lane -------------------------------------------
      const int j0 = row * kFlowRowQ + lane;
      double qd[kFlowQ][3];
      float qf[kFlowQ][3];
      int sorig[kFlowQ];
#pragma unroll
      for (int q = 0; q < kFlowQ; ++q) {
        const double2* s2 = reinterpret_cast<const double2*>(src64 + min(j0 + 32 * q, ns - 1));   // ragged tail: repeats
        const double2 a = __ldg(s2), b = __ldg(s2 + 1);
        sorig[q] = (int)__double_as_longlong(b.y);
        qd[q][0] = fma(T[0], a.x, fma(T[1], a.y, fma(T[2], b.x, T[3])));
        qd[q][1] = fma(T[4], a.x, fma(T[5], a.y, fma(T[6], b.x, T[7])));
        qd[q][2] = fma(T[8], a.x, fma(T[9], a.y, fma(T[10], b.x, T[11])));
#pragma unroll
        for (int a3 = 0; a3 < 3; ++a3) qf[q][a3] = (float)qd[q][a3];
      }
      float lo3[3], hi3[3];
#pragma unroll
      for (int a3 = 0; a3 < 3; ++a3) {
        lo3[a3] = hi3[a3] = qf[0][a3];
#pragma unroll
        for (int q = 1; q < kFlowQ; ++q) { lo3[a3] = fminf(lo3[a3], qf[q][a3]); hi3[a3] = fmaxf(hi3[a3], qf[q][a3]); }
      }
      const float rlx = redux_minf(lo3[0]), rhx = redux_maxf(hi3[0]);
      const float rly = redux_minf(lo3[1]), rhy = redux_maxf(hi3[1]);
      const float rlz = redux_minf(lo3[2]), rhz = redux_maxf(hi3[2]);

      f32x2 qxx[kFlowQ], qyy[kFlowQ], qzz[kFlowQ];
      float bestv[kFlowQ], secondv[kFlowQ], thr[kFlowQ], a_bound[kFlowQ], gbest[kFlowQ][8];
      int bestb[kFlowQ];
#pragma unroll
      for (int q = 0; q < kFlowQ; ++q) {
        qxx[q] = pack2(qf[q][0], qf[q][0]); qyy[q] = pack2(qf[q][1], qf[q][1]); qzz[q] = pack2(qf[q][2], qf[q][2]);
        bestv[q] = secondv[q] = thr[q] = kInfF;
        bestb[q] = -1;
#pragma unroll
        for (int g = 0; g < 8; ++g) gbest[q][g] = kInfF;
        a_bound[q] = fmaxf(fabsf(qf[q][0]), fmaxf(fabsf(qf[q][1]), fabsf(qf[q][2]))) * (1.0f + 4.0f * kU) + tg.tmax;
      }
      // sweep block b unless no lane's own bound admits it
      auto consider = [&](int b) {
        if (PRUNE) {
          const float4 blo = __ldg(tg.box + 2 * b), bhi = __ldg(tg.box + 2 * b + 1);
          bool need = false;
#pragma unroll
          for (int q = 0; q < kFlowQ; ++q)
            need = need || !(box_lb3(qf[q][0], qf[q][1], qf[q][2], qf[q][0], qf[q][1], qf[q][2], blo, bhi) > thr[q]);
          if (!__any_sync(full, need)) return;
        }
        tiles += (unsigned)kFlowQ;
        const bool imp = flow_sweep_block(tg.t32, b, qxx, qyy, qzz, bestv, secondv, bestb, gbest);
        if (__any_sync(full, imp)) {
#pragma unroll
          for (int q = 0; q < kFlowQ; ++q) thr[q] = candidate_threshold(bestv[q], a_bound[q], 0.f);
        }
      };
      // phase A: the blocks whose box overlaps the row box (bound 0: they pass any threshold)
      float seed_lb = kInfF;
      int seed_b = 0;
      bool swept_any = false;
      for (int c0 = 0; c0 < tg.nblk; c0 += 32) {
        const int b = c0 + lane;                 // boxes are padded with +inf boxes to a multiple of 32 blocks
        const float lb = box_lb3(rlx, rly, rlz, rhx, rhy, rhz, __ldg(tg.box + 2 * b), __ldg(tg.box + 2 * b + 1));
        if (lb < seed_lb) { seed_lb = lb; seed_b = b; }
        unsigned m = PRUNE ? __ballot_sync(full, lb == 0.f) : __ballot_sync(full, b < tg.nblk);
        swept_any = swept_any || m != 0u;
        while (m) {
          const int bb = c0 + __ffs((int)m) - 1;
          m &= m - 1u;
          consider(bb);
        }
      }
      if (PRUNE) {
        int seed_block = -1;
        if (!swept_any) {      // the row overlaps no block: start from the block of smallest bound
          const unsigned mn = __reduce_min_sync(full, __float_as_uint(seed_lb));
          const unsigned who = __ballot_sync(full, __float_as_uint(seed_lb) == mn);
          seed_block = __shfl_sync(full, seed_b, __ffs((int)who) - 1);
          if (seed_block < tg.nblk) consider(seed_block);
        }
        // phase B: every other block within the largest candidate threshold of the row
        auto thr_max = [&]() { float m = thr[0];
#pragma unroll
          for (int q = 1; q < kFlowQ; ++q) m = fmaxf(m, thr[q]);
          return m; };
        float bound = redux_maxf(thr_max());
        for (int c0 = 0; c0 < tg.nblk; c0 += 32) {
          const int b = c0 + lane;
          const float lb = box_lb3(rlx, rly, rlz, rhx, rhy, rhz, __ldg(tg.box + 2 * b), __ldg(tg.box + 2 * b + 1));
          unsigned m = __ballot_sync(full, lb > 0.f && !(lb > bound) && b != seed_block);
          if (m == 0u) continue;
          while (m) {
            const int bb = c0 + __ffs((int)m) - 1;
            m &= m - 1u;
            consider(bb);
          }
          bound = redux_maxf(thr_max());
        }
      }

      // ---- exact resolution in float64 ----------------------------------------------------
      int cpos[kFlowQ], corig[kFlowQ];
      bool amb[kFlowQ];
#pragma unroll
      for (int q = 0; q < kFlowQ; ++q) {
        cpos[q] = -1; corig[q] = 0x7fffffff;
        amb[q] = !(secondv[q] > thr[q]);
        if (!amb[q]) {
          unsigned gmask = 0;
#pragma unroll
          for (int g = 0; g < 8; ++g)
            if (gbest[q][g] <= thr[q]) gmask |= 1u << g;
          double bd = kInf;
          while (gmask) {
            const int g = __ffs((int)gmask) - 1;
            gmask &= gmask - 1u;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int pos = bestb[q] * 32 + 4 * g + u;
              const double2* m2 = reinterpret_cast<const double2*>(tg.t64 + pos);
              const double2 a = __ldg(m2), b = __ldg(m2 + 1);
              const int oi = (int)__double_as_longlong(b.y);
              const double ex = qd[q][0] - a.x, ey = qd[q][1] - a.y, ez = qd[q][2] - b.x;
              const double d = ex * ex + ey * ey + ez * ez;
              if (d < bd || (d == bd && oi < corig[q])) { bd = d; corig[q] = oi; cpos[q] = pos; }
            }
          }
          if (cpos[q] < 0) amb[q] = true;   // non-finite data only
        }
      }
      // rare: a second block is within the slack -> the warp scans all targets
#pragma unroll
      for (int q = 0; q < kFlowQ; ++q) {
        unsigned vote = __ballot_sync(full, amb[q]);
        while (vote) {
          const int src_lane = __ffs((int)vote) - 1;
          vote &= vote - 1u;
          const float fx = __shfl_sync(full, qf[q][0], src_lane);
          const float fy = __shfl_sync(full, qf[q][1], src_lane);
          const float fz = __shfl_sync(full, qf[q][2], src_lane);
          const float thr_s = __shfl_sync(full, thr[q], src_lane);
          const double sx = __shfl_sync(full, qd[q][0], src_lane);
          const double sy = __shfl_sync(full, qd[q][1], src_lane);
          const double sz = __shfl_sync(full, qd[q][2], src_lane);
          double bd = kInf;
          int bo = 0x7fffffff, bp = -1;
          for (int i = lane; i < tg.nt; i += 32) {
            const float* g = tg.t32 + (size_t)(i >> 2) * 12 + (i & 3);
            // !(d > thr) also admits NaN so that non-finite data still yields an index
            if (!(dist2_3d(fx, fy, fz, __ldg(g), __ldg(g + 4), __ldg(g + 8)) > thr_s)) {
              const double2* m2 = reinterpret_cast<const double2*>(tg.t64 + i);
              const double2 a = __ldg(m2), b = __ldg(m2 + 1);
              const int oi = (int)__double_as_longlong(b.y);
              const double ex = sx - a.x, ey = sy - a.y, ez = sz - b.x;
              const double d = ex * ex + ey * ey + ez * ez;
              if (bp < 0 || d < bd || (d == bd && oi < bo)) { bd = d; bo = oi; bp = i; }
            }
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            const double od = __shfl_xor_sync(full, bd, o);
            const int oo = __shfl_xor_sync(full, bo, o);
            const int op = __shfl_xor_sync(full, bp, o);
            if (op >= 0 && (bp < 0 || od < bd || (od == bd && oo < bo))) { bd = od; bo = oo; bp = op; }
          }
          if (lane == src_lane) { cpos[q] = bp < 0 ? 0 : bp; corig[q] = bp < 0 ? (int)__double_as_longlong(__ldg(reinterpret_cast<const double2*>(tg.t64) + 1).y) : bo; }
        }
      }
      // ---- float64 sums -------------------------------------------------------------------
#pragma unroll
      for (int q = 0; q < kFlowQ; ++q) {
        if (j0 + 32 * q < ns) {
          const double2* m2 = reinterpret_cast<const double2*>(tg.t64 + cpos[q]);
          const double2 a = __ldg(m2), b = __ldg(m2 + 1);
          const double x = qd[q][0], y = qd[q][1], z = qd[q][2], ax = a.x, ay = a.y, az = b.x;
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
          if (p.corr_out) p.corr_out[(size_t)pair * p.n_max + sorig[q]] = corig[q];
          if (p.nearest_only && p.dist2_out) p.dist2_out[(size_t)pair * p.n_max + sorig[q]] = d;
        }
      }
#pragma unroll
      for (int k = 0; k < kSums3d; ++k) acc[k] = warp_sum(acc[k]);
      if (normalize) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          acc[kSums3d + 2 * k] = warp_min_f64(acc[kSums3d + 2 * k]);
          acc[kSums3d + 2 * k + 1] = warp_max_f64(acc[kSums3d + 2 * k + 1]);
        }
      }
    }
    if (lane == 0) {
#pragma unroll
      for (int k = 0; k < kRed3d; ++k) s_part[warp][k] = acc[k];
    }
    FLOW_DBG(3);
    __syncthreads();
    if (warp != 0) continue;          // warps 1..7 wait for the next item at the barrier above
    FLOW_DBG_W0(4);

    // ---- warp 0: this chunk's partial sums to global memory; retire the chunk -------------------
    const bool is_min = lane >= kSums3d && ((lane - kSums3d) & 1) == 0;
    const bool is_max = lane >= kSums3d && ((lane - kSums3d) & 1) == 1;
    if (lane < kRed3d) {
      double v = s_part[0][lane];
#pragma unroll
      for (int w = 1; w < kFlowWarps; ++w) {
        const double o = s_part[w][lane];
        v = is_min ? fmin(v, o) : (is_max ? fmax(v, o) : v + o);
      }
      __stcg(part + ((size_t)pair * L.max_chunks + chunk) * kRed3d + lane, v);
    }
    __threadfence();
    __syncwarp();
    unsigned old = 0;
    if (lane == 0) old = atomicAdd(&P->done, 1u);
    old = __shfl_sync(full, old, 0);
    FLOW_CNT(1);
    if (old != (unsigned)(n_chunks - 1)) continue;
    FLOW_DBG_W0(5); FLOW_CNT(2);

    // ---- last chunk of the pass: totals in chunk order, solve, stop tests ------------------------
    __threadfence();
    double v = 0.0;
    if (lane < kRed3d) {
      v = __ldcg(part + (size_t)pair * L.max_chunks * kRed3d + lane);
      for (int c = 1; c < n_chunks; ++c) {
        const double o = __ldcg(part + ((size_t)pair * L.max_chunks + c) * kRed3d + lane);
        v = is_min ? fmin(v, o) : (is_max ? fmax(v, o) : v + o);
      }
    }
    double tot[kRed3d];
#pragma unroll
    for (int k = 0; k < kRed3d; ++k) tot[k] = __shfl_sync(full, v, k);
    // every lane reads the pair's mutable state BEFORE lane 0 rewrites it below: without the barrier a
    // lane that runs late would see the updated stop state, take the other branch and leave the warp
    // waiting for it forever (seen at >= 24 chunks per pass)
    const int passes = __ldcg(&P->passes) + 1;
    StopState st{__ldcg(&P->prev_err), __ldcg(&P->have_prev), __ldcg(&P->count)};
    __syncwarp();
    double err = tot[15];
    if (normalize) {
      const double wx = tot[17] - tot[16], wy = tot[19] - tot[18], wz = tot[21] - tot[20];
      const double diag = sqrt(wx * wx + wy * wy + wz * wz);
      err = err / (diag * diag * (double)ns);
    }
    double bottom[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) bottom[k] = P->bottom[k];
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
      double nT[12];
#pragma unroll
      for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int c = 0; c < 4; ++c)
          nT[a * 4 + c] = Rm[a * 3] * T[c] + Rm[a * 3 + 1] * T[4 + c] + Rm[a * 3 + 2] * T[8 + c] + dt[a] * bottom[c];
#pragma unroll
      for (int k = 0; k < 12; ++k) T[k] = nT[k];
      status = stop_test(st, err, P->epsilon, p.max_iters, p.stopping_thresh);
      status = __shfl_sync(full, status, 0);          // one decision for the warp
      if (status < 0 && lane == 0) {
        P->prev_err = st.prev_err; P->have_prev = st.have_prev; P->count = st.count;
      }
    }
    FLOW_DBG_W0(6);
    if (status < 0) {
      // ---- next pass: publish the pose, then the items -----------------------------------------
      if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 12; ++k) __stcg(&P->T[k], T[k]);
        P->passes = passes;
        P->done = 0u;
      }
      __threadfence();
      __syncwarp();
      FLOW_DBG_W0(7);
      unsigned pos = 0;
      if (lane == 0) pos = atomicAdd(&ctl[kCtlTail], (unsigned)n_chunks);
      FLOW_DBG_W0(70);
      pos = __shfl_sync(full, pos, 0);
      FLOW_DBG_W0(71);
#ifdef B200ICP_FLOW_DEBUG
      if (p.dist2_out && lane == 0) { reinterpret_cast<volatile unsigned*>(p.dist2_out)[32] = pos; reinterpret_cast<volatile unsigned*>(p.dist2_out)[33] = (unsigned)n_chunks; reinterpret_cast<volatile unsigned*>(p.dist2_out)[34] = (unsigned)L.ring_cap; }
#endif
      // the ring never fills up (capacity 2x the items that can be in flight): every slot written here is empty
      for (int c = lane; c < n_chunks; c += 32) {
        unsigned* slot = ring + ((pos + (unsigned)c) & (unsigned)(L.ring_cap - 1));
        const unsigned it = ((unsigned)pair << 21) | (unsigned)c;
        const unsigned was = atomicCAS(slot, kRingEmpty, it);
#ifdef B200ICP_FLOW_DEBUG
        if (p.dist2_out) reinterpret_cast<volatile unsigned*>(p.dist2_out)[128 + c] = was;
#endif
        if (was != kRingEmpty) atomicExch(&ctl[kCtlAbort], 2u);       // cannot happen; never spin on it
      }
      FLOW_DBG_W0(72);
      FLOW_DBG_W0(8); FLOW_CNT(3);
    } else {
      // ---- the pair is finished ------------------------------------------------------------------
      if (lane == 0) {
        if (p.T_out) {
          double* To = p.T_out + (size_t)pair * 16;
#pragma unroll
          for (int k = 0; k < 12; ++k) To[k] = T[k];
#pragma unroll
          for (int k = 0; k < 4; ++k) To[12 + k] = bottom[k];
        }
        if (p.err_out) p.err_out[pair] = err;
        if (p.iters_out) p.iters_out[pair] = passes;
        if (p.status_out) p.status_out[pair] = status;
        store_pose_to_peers(p, pair, T[3], T[7], T[0], T[4], err);
      }
      if (p.corr_out)
        for (int j = ns + lane; j < p.n_max; j += 32) p.corr_out[(size_t)pair * p.n_max + j] = -1;
      if (p.n_peers > 0) __threadfence_system(); else __threadfence();
      __syncwarp();
      if (lane == 0) {
        const unsigned was = atomicAdd(&ctl[kCtlPairsDone], 1u);
        if (p.n_peers > 0 && p.flag_off != 0 && was == (unsigned)(L.n_pairs - 1)) raise_arrival_flags(p);
      }
    }
  }
  FLOW_DBG(9);
  if (lane == 0 && tiles)
    atomicAdd(reinterpret_cast<unsigned long long*>(ctl) + 1, (unsigned long long)tiles);
}

}  // namespace b200icp
