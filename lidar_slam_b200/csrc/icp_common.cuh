// Device helpers shared by the 2-D and 3-D ICP kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200icp.h"

namespace b200icp {

unsigned env_flags();   // developer overrides from the environment, read once (api.cu)

constexpr int kBlockTargets = 32;          // targets per bookkeeping block in the NN sweep
constexpr float kU = 5.9604644775390625e-8f;  // 2^-24, fp32 unit round-off
constexpr float kBigCoord = 1.0e18f;       // sentinel coordinate for padded targets

// ---------------------------------------------------------------------------
// fp32 candidate distance.  Written with explicit round-to-nearest intrinsics so
// that the sweep and the rescan produce bit-identical values (no re-association,
// no surprise contraction).  2 FADD + FMUL + FFMA = the 4 FP32-pipe
// lane-instructions SURVEY.md section 8d counts per 2-D candidate.
// ---------------------------------------------------------------------------
__device__ __forceinline__ float dist2_2d(float qx, float qy, float tx, float ty) {
  float dx = __fsub_rn(qx, tx);
  float dy = __fsub_rn(qy, ty);
  return __fmaf_rn(dx, dx, __fmul_rn(dy, dy));
}
__device__ __forceinline__ float dist2_3d(float qx, float qy, float qz, float tx, float ty, float tz) {
  float dx = __fsub_rn(qx, tx);
  float dy = __fsub_rn(qy, ty);
  float dz = __fsub_rn(qz, tz);
  return __fmaf_rn(dx, dx, __fmaf_rn(dy, dy, __fmul_rn(dz, dz)));
}

// Packed fp32x2 arithmetic (Blackwell FADD2 / FMUL2 / FFMA2): two IEEE round-to-nearest
// operations per lane per issue slot, bit-identical to the scalar intrinsics above, so the
// sweep (packed, two targets at a time) and the rescans (scalar) agree exactly.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}

// Three-input fp32 minimum: one FMNMX3 on sm_100 (PTX ISA 8.6+).
__device__ __forceinline__ float fmin3(float a, float b, float c) {
  float r;
  asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}

// ---------------------------------------------------------------------------
// Rigorous slack between an fp32 candidate distance and the float64 one.
//
// Let q, t be the float64 points, q32 = fl(q), t32 = fl(t), u = 2^-24, and let
// A >= max|q_c| + max|t_c| over the coordinates c.  Per coordinate the computed
// difference d_c = fl(q32_c - t32_c) obeys |d_c - (q_c - t_c)| <= u*A + u*|d_c|.
// Squaring, summing n_dim terms and adding the rounding of the (at most 3)
// multiply/FMA steps gives, with s = sqrt(d32),
//     |d32 - d64| <= 2*sqrt(n_dim)*u*(A + s)*s + n_dim*u^2*(A + s)^2 + 3*u*d32 .
// slack() returns a bound that dominates this for n_dim <= 3 with more than a
// factor 2 to spare.  A candidate set {t : d32(t) <= m + 2*slack} around the
// fp32 minimum m therefore always contains the float64 nearest neighbour; the
// callers use 4*slack (and evaluate at an upper bound of m).
// tests/test_host.py::test_fp32_slack_bounds_the_fp64_distance checks the inequality on random clouds,
// tests/test_margin.py on near-ties, offsets of kilometres, subnormal separations and fp32 collisions.
// ---------------------------------------------------------------------------
__device__ __forceinline__ float slack(float d_hi, float a_bound) {
  float s = sqrtf(d_hi);
  float as = a_bound + s;
  // + FLT_MIN: where the squares are subnormal in fp32 (separations below ~1e-19) the relative bound
  // underflows to 0 while every product still errs by up to 2^-150 (tests/test_margin.py)
  return 8.0f * kU * as * s + 8.0f * kU * d_hi + 8.0f * kU * kU * as * as + 1.1754944e-38f;
}

// Upper bound of the true block minimum given the truncated packed value
// (low 6..10 mantissa bits replaced by the block id), plus the candidate threshold.
__device__ __forceinline__ float candidate_threshold(float trunc_min, float a_bound, float trunc_rel) {
  float hi = __fmaf_rn(trunc_min, trunc_rel, trunc_min);  // >= every value that truncates to trunc_min
  float thr = hi + 4.0f * slack(hi, a_bound);
  return __fmaf_rn(thr, 4.0f * kU, thr);                  // round the threshold up, never down
}

// ---------------------------------------------------------------------------
// warp reductions
// ---------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_maxf(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_minf(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// (distance, index) minimum; exact ties go to the lower index.
__device__ __forceinline__ void warp_argmin(double& d, int& i) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    double od = __shfl_xor_sync(0xffffffffu, d, o);
    int oi = __shfl_xor_sync(0xffffffffu, i, o);
    if (od < d || (od == d && oi < i)) { d = od; i = oi; }
  }
}

// ---------------------------------------------------------------------------
// TMA bulk copy (1-D, no tensor map) + mbarrier, for staging scans into smem.
// SASS: UBLKCP / SYNCS.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_addr(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_addr(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async_global() {
  asm volatile("fence.proxy.async.global;" ::: "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// system-scope acquire load: pairs with a flag written by a host-to-device copy that follows the
// data it announces in stream order
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned int* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// bytes must be a multiple of 16; both addresses 16-byte aligned.
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gmem_src, uint32_t bytes,
                                            uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_addr(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_addr(bar))
      : "memory");
}

// ---------------------------------------------------------------------------
// Stop logic of the reference driver (modules/icp.py:172-181), one pass.
// Returns -1 to continue, otherwise the B200ICP_STATUS_* that ends the pair.
// ---------------------------------------------------------------------------
struct StopState {
  double prev_err;
  int have_prev;
  int count;  // the reference's `iteration`
};
__device__ __forceinline__ int stop_test(StopState& st, double err, double epsilon, int max_iters,
                                         double stopping_thresh) {
  if (!(fabs(err) <= 1.7976931348623157e308)) return B200ICP_STATUS_NONFINITE;
  if (err < epsilon) return B200ICP_STATUS_EPSILON;
  if (st.count >= max_iters) return B200ICP_STATUS_MAX_ITERS;
  if (st.have_prev && fabs(st.prev_err - err) < stopping_thresh) return B200ICP_STATUS_STALLED;
  st.prev_err = err;
  st.have_prev = 1;
  st.count += 1;
  return -1;
}

constexpr int kMaxPeers = 16;

struct RunParams {
  const double* scans;
  const int32_t* counts;
  int n_scans, n_max;
  const int32_t* pairs;
  const double* init;
  const double* epsilon;
  const uint8_t* normalize;
  int n_pairs, max_iters;
  double stopping_thresh;
  double* T_out;
  double* err_out;
  int32_t* iters_out;
  int32_t* status_out;
  int32_t* corr_out;
  double* dist2_out;      // nearest-only mode
  int nearest_only;       // 1: single NN pass on untransformed queries, no solve
  unsigned int* queue;    // work counter in the workspace (zeroed before launch)
  // fused pose exchange (multi-GPU): when n_peers > 0 the thread that finishes pair b also
  // stores its packed pose row (x, y, theta, error) into row peer_row0 + b of every rank's
  // gather buffer -- peer-mapped device pointers, i.e. plain stores that travel over NVLink.
  double* peer_out[kMaxPeers];
  int n_peers;
  long long peer_row0;
  // arrival flags (b200icp_exchange): flag_off != 0 -> the thread that retires the launch's last pair
  // stores flag_val into the uint32 at byte flag_off + 4 * flag_idx of every peer buffer
  long long flag_off;
  int flag_idx;
  unsigned flag_val;
  // gated launch (b200icp_run_gated): scans [0, *ready_scans) of the pool are resident; the value
  // is raised by copies on another stream while the kernel runs.  Null: the whole pool is resident.
  const unsigned int* ready_scans;
  unsigned long long gate_timeout_ns;   // gated launch: give a pair up after waiting this long for its scans
  // RANGES mode (b200icp_run_ranges): the pool is [n_scans, n_beams] ranges (float64, or float32 when
  // ranges_f32), scans == counts == nullptr; the planar kernel builds the two scans of a pair itself.
  const void* ranges;
  int ranges_f32, n_beams;
  const double* cos_tab;
  const double* sin_tab;
  double range_min, range_max, off_x, off_y;
  int32_t* counts_out;                  // [n_scans] or nullptr: valid beams per scan
};

// kept out of line: atan2's slow path must not widen the register footprint of the ICP loops
__device__ __noinline__ double pose_angle(double r10, double r00) { return atan2(r10, r00); }

// pose_from_T(TSE2_from_TSE3(T)) (modules/utils.py:196-225) + error, to every peer buffer
__device__ __forceinline__ void store_pose_to_peers(const RunParams& p, int pair, double x, double y,
                                                    double r00, double r10, double err) {
  if (p.n_peers <= 0) return;
  double4 row;
  row.x = x; row.y = y; row.z = pose_angle(r10, r00); row.w = err;
  for (int k = 0; k < p.n_peers; ++k)
    reinterpret_cast<double4*>(p.peer_out[k])[p.peer_row0 + pair] = row;
}

__device__ __forceinline__ void raise_arrival_flags(const RunParams& p) {
  __threadfence_system();
  for (int k = 0; k < p.n_peers; ++k) {
    unsigned* f = reinterpret_cast<unsigned*>(reinterpret_cast<char*>(p.peer_out[k]) + p.flag_off) + p.flag_idx;
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f), "r"(p.flag_val) : "memory");
  }
}
// Pairs retired by ONE thread of a CTA (the thread that wrote their results and stored their rows to the
// peers), announced together when the CTA leaves the work queue: one system-scope fence per CTA instead of
// one per pair (a fence waits for the acknowledgement of every peer store over NVLink; per pair it cost
// 29 us per 4955-pair launch at 2 GPUs, profiles/r2c_*).  The CTA whose count completes the launch
// announces it to every rank.
__device__ __forceinline__ void pairs_retired(const RunParams& p, unsigned* counter, unsigned count) {
  if (p.n_peers <= 0 || p.flag_off == 0 || count == 0u) return;
  __threadfence_system();
  if (atomicAdd(counter, count) + count == (unsigned)p.n_pairs) raise_arrival_flags(p);
}

}  // namespace b200icp
