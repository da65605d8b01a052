// Occupancy-grid ray casting on the device: the consumer of the
// poses the ICP path produces.
//
//   ogm_build_kernel   OccupancyGridMap.update_map / build_map   modules/ogm.py:149-195, :46-60
//                      OccupancyGridMap.bresenham2D              modules/ogm.py:197-231
//                      OccupancyGridMap.world2grid               modules/ogm.py:104-126
//   ogm_binarize       the pmf thresholding of build_map         modules/ogm.py:62-65
//
// The reference updates a float32 log-odds grid one ray after another:
//     cell -= log 4  for every in-bounds cell of the ray but the last,  cell += log 4 for the last,
// each update being  f32(f64(cell) -/+ log 4)  (numpy in-place op with a float64 scalar), and
// clips the grid to [-20, 20] after every scan.  Rounding makes the result depend on the ORDER
// of the updates a cell receives, so the kernel reproduces that order exactly:
//   phase 1  all rays of the scan are walked in parallel (one thread per ray; the in-bounds part
//            of a ray is entered through the closed form of the reference's Bresenham variant);
//            every visited cell gets its decrement / increment COUNT with a fire-and-forget
//            atomic on an int32 scratch grid;
//   phase 2  the scan's bounding window of the scratch grid is swept: a cell that only saw
//            decrements (or only increments) replays that many rounded updates -- the order
//            among identical updates is immaterial; a cell that saw both is queued;
//   phase 3  rays are walked again and append (ray, sign) events to the queued cells, each
//            cell then sorts its events by ray index and replays them in the reference's order.
// Scans are order-dependent (clip + rounding), so one map is a sequential chain of scans.  A
// thread-block CLUSTER of 16 CTAs (16 SMs) owns one map: the per-scan work is spread over the
// cluster's threads (the count atomics are bound by the per-SM atomic issue rate, so SMs are
// what it needs) and the phases are separated by hardware cluster barriers instead of kernel
// launches.  Several maps (odometry / scan-matching / optimised trajectories) run side by
// side, one cluster each.
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200icp.h"

namespace {

constexpr int kThreads = 256;     // per CTA
constexpr int kCluster = 16;      // CTAs (= SMs) that share one map; opt-in cluster size on sm_100
constexpr int kMixedCap = 16384;
constexpr double kLogOdds = 1.3862943611198906;   // np.log(4.)  (modules/ogm.py:44)
constexpr int kSaturate = 30;                     // 30 * log 4 = 41.6 > 40 = the width of the clip interval

struct OgmParams {
  float* log_odds;        // [n_maps][width*height]
  int width, height;
  double res, min_x, min_y;
  const double* poses;    // [n_maps][n_scans][3]  (x, y, yaw)
  const double* cs;       // [n_maps][n_scans][2]  (cos yaw, sin yaw) as numpy evaluates them
  const double* scans;    // [n_scans][n_max][2] shared by all maps
  const int32_t* counts;  // [n_scans]
  int n_scans, n_max, n_maps;
  int32_t* workspace;     // per map: scratch[w*h] | ctl[16] | mixed_cell/cnt/off/fill[cap] | events[ev_cap]
  long long ws_stride;    // int32 elements per map
  int ev_cap;
};

// modules/ogm.py:104-126: ceil((x - min)/res) - 1 (float64 arithmetic, then integer)
__device__ __forceinline__ int world2grid(double v, double vmin, double res) {
  return (int)ceil(__ddiv_rn(__dsub_rn(v, vmin), res)) - 1;
}

struct Ray {
  int sx, sy, ex, ey;   // start / end cell
  int dx, dy;           // major / minor extents (after the steep swap)
  int steep, c0;        // c0 = dx - 1 - floor(dx/2): numerator offset of the closed form
  int klo, khi;         // in-bounds range of major steps (empty if klo > khi)
};

// Cell k of the reference's bresenham2D (modules/ogm.py:197-231): the major coordinate moves by
// k, the minor one by  m_k = floor((k*dy + dx - 1 - floor(dx/2)) / dx),  the count of
// wrap-arounds of the error term  mod(floor(dx/2) - k*dy, dx)  that the reference detects with
// diff(...) >= 0 (checked against the reference's known answers in tests/test_ogm_oracle.py).
__device__ __forceinline__ Ray make_ray(int sx, int sy, int ex, int ey, int width, int height) {
  Ray r;
  r.sx = sx; r.sy = sy; r.ex = ex; r.ey = ey;
  int dx = abs(ex - sx), dy = abs(ey - sy);
  r.steep = dy > dx;
  if (r.steep) { const int t = dx; dx = dy; dy = t; }
  r.dx = dx; r.dy = dy;
  r.c0 = dx - 1 - dx / 2;
  // in-bounds steps form one interval (a line meets a box in a segment): bisect both ends
  // on the predicate "cell k is inside", which is monotone on either side of the interval.
  // cell 0 is the start cell and cell dx the end cell (m_0 = 0, m_dx = dy: c0 < dx)
  const bool in0 = sx >= 0 && sx < width && sy >= 0 && sy < height;
  const bool in1 = ex >= 0 && ex < width && ey >= 0 && ey < height;
  if (in0 && in1) { r.klo = 0; r.khi = dx; return r; }
  // find any inside step by clipping the major coordinate range first
  int lo = 0, hi = dx;
  {
    // major coordinate is monotone in k: restrict to where the major coordinate is in range
    const int s = r.steep ? sy : sx, e = r.steep ? ey : ex, lim = r.steep ? height : width;
    // coordinate at k: s + sign*k
    if (s <= e) { lo = max(lo, -s); hi = min(hi, lim - 1 - s); }
    else        { lo = max(lo, s - (lim - 1)); hi = min(hi, s); }
  }
  if (lo > hi) { r.klo = 1; r.khi = 0; return r; }
  // minor coordinate is monotone (non-strict) in k as well: bisect its entry and exit
  const int ms = r.steep ? sx : sy, me = r.steep ? ex : ey, mlim = r.steep ? width : height;
  auto minor = [&](int k) {
    const int m = dx > 0 ? (int)(((unsigned)k * (unsigned)dy + (unsigned)r.c0) / (unsigned)dx) : 0;
    return ms + (ms <= me ? m : -m);
  };
  // first k in [lo,hi] whose minor coordinate is within range, last such k
  int a = lo, b = hi;
  if (ms <= me) {
    // minor increasing: need minor(k) >= 0 (entry) and <= mlim-1 (exit)
    int l = lo, h = hi + 1;               // first k with minor(k) >= 0
    while (l < h) { const int mid = (l + h) >> 1; if (minor(mid) >= 0) h = mid; else l = mid + 1; }
    a = l;
    l = lo - 1; h = hi;                   // last k with minor(k) <= mlim-1
    while (l < h) { const int mid = (l + h + 1) >> 1; if (minor(mid) <= mlim - 1) l = mid; else h = mid - 1; }
    b = l;
  } else {
    int l = lo, h = hi + 1;               // minor decreasing: first k with minor(k) <= mlim-1
    while (l < h) { const int mid = (l + h) >> 1; if (minor(mid) <= mlim - 1) h = mid; else l = mid + 1; }
    a = l;
    l = lo - 1; h = hi;                   // last k with minor(k) >= 0
    while (l < h) { const int mid = (l + h + 1) >> 1; if (minor(mid) >= 0) l = mid; else h = mid - 1; }
    b = l;
  }
  r.klo = a; r.khi = b;
  return r;
}

__device__ __forceinline__ float apply_updates(float a, int n, double delta) {
  for (int i = 0; i < n; ++i) a = __double2float_rn(__dadd_rn((double)a, delta));
  return a;
}
__device__ __forceinline__ float clip20(float a) { return fminf(fmaxf(a, -20.0f), 20.0f); }

// Walk every ray of the scan: MODE 0 counts visits on the scratch grid, MODE 1 appends
// (ray, sign) events to the queued (tagged) cells.
template <int MODE>
__device__ __forceinline__ void walk_rays(const int2* ends, int n_rays, int sx, int sy, int width,
                                          int height, int32_t* scratch, const int32_t* mixed_off,
                                          int32_t* mixed_fill, int32_t* events, int gtid, int gthreads) {
  // one thread per ray, classic incremental stepping: rem is the numerator of the closed form
  // modulo dx, m its quotient
  for (int ray = gtid; ray < n_rays; ray += gthreads) {
    const int2 e = ends[ray];
    const Ray r = make_ray(sx, sy, e.x, e.y, width, height);
    if (r.klo > r.khi) continue;
    const int dx = r.dx, dy = r.dy;
    int m = 0, rem = r.c0;
    if (r.klo > 0) {
      const unsigned num = (unsigned)r.klo * (unsigned)dy + (unsigned)r.c0;
      m = (int)(num / (unsigned)dx);
      rem = (int)(num - (unsigned)m * (unsigned)dx);
    }
    // cell index advances by a fixed stride per major step and another per minor step
    const int smaj = r.steep ? (r.sy <= r.ey ? 1 : -1) : (r.sx <= r.ex ? height : -height);
    const int smin = r.steep ? (r.sx <= r.ex ? height : -height) : (r.sy <= r.ey ? 1 : -1);
    int cell = sx * height + sy + r.klo * smaj + m * smin;
    if (MODE == 0) {
      for (int k = r.klo; k <= r.khi; ++k) {
        atomicAdd(&scratch[cell], k == r.khi ? 0x10000 : 1);
        rem += dy;
        cell += smaj;
        if (rem >= dx && dx > 0) { rem -= dx; cell += smin; }
      }
    } else {
      // the tag loads are independent of one another: issue eight before looking at any
      constexpr int U = 8;
      for (int k0 = r.klo; k0 <= r.khi; k0 += U) {
        int cells[U], tags[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          cells[u] = cell;
          tags[u] = (k0 + u <= r.khi) ? __ldcg(&scratch[cell]) : 0;
          rem += dy;
          cell += smaj;
          if (rem >= dx && dx > 0) { rem -= dx; cell += smin; }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (tags[u] < 0) {
            const int slot = tags[u] & 0x7fffffff;
            const int pos = atomicAdd(&mixed_fill[slot], 1);
            events[mixed_off[slot] + pos] = (ray << 1) | ((k0 + u == r.khi) ? 1 : 0);
          }
        }
      }
    }
  }
}

#ifdef B200ICP_OGM_PROFILE
__device__ __forceinline__ long long gtime() { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define PH(i) do { const long long now_ = gtime(); ph[i] += now_ - t_last; t_last = now_; } while (0)
#else
#define PH(i) do { } while (0)
#endif

__global__ void __launch_bounds__(kThreads, 1) ogm_build_kernel(const OgmParams p) {
#ifdef B200ICP_OGM_PROFILE
  long long ph[6] = {0, 0, 0, 0, 0, 0};
  long long t_last = gtime();
#endif
  extern __shared__ __align__(16) unsigned char smem_raw[];
  int2* ends = reinterpret_cast<int2*>(smem_raw);                  // [n_max] end cells of the rays
  __shared__ int s_bbox[4];
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  const int map = blockIdx.x / kCluster;
  const int tid = threadIdx.x;
  const int gtid = (int)cluster.block_rank() * kThreads + tid;      // thread index within the cluster
  constexpr int gthreads = kCluster * kThreads;
  float* grid = p.log_odds + (size_t)map * p.width * p.height;
  int32_t* ws = p.workspace + (size_t)map * p.ws_stride;
  int32_t* scratch = ws;
  int32_t* ctl = ws + (size_t)p.width * p.height;                  // [0] n_mixed, [1] ev_total, [2] again
  int32_t* mixed_cell = ctl + 16;
  int32_t* mixed_cnt = mixed_cell + kMixedCap;
  int32_t* mixed_off = mixed_cnt + kMixedCap;
  int32_t* mixed_fill = mixed_off + kMixedCap;
  int32_t* events = mixed_fill + kMixedCap;
  const int W = p.width, H = p.height;

  for (int s = 0; s < p.n_scans; ++s) {
    const double* pose = p.poses + ((size_t)map * p.n_scans + s) * 3;
    const double* cs = p.cs + ((size_t)map * p.n_scans + s) * 2;
    const double x = pose[0], y = pose[1], c = cs[0], sn = cs[1];
    const int n_rays = min(p.counts[s], p.n_max);
    // ray start: pose + lidar offset, NOT rotated (modules/ogm.py:165-166)
    const int sx = world2grid(__dadd_rn(x, 0.13323), p.min_x, p.res);
    const int sy = world2grid(__dadd_rn(y, 0.0), p.min_y, p.res);
    if (tid == 0) { s_bbox[0] = sx; s_bbox[1] = sx; s_bbox[2] = sy; s_bbox[3] = sy; }
    __syncthreads();
    // end points to the world frame (modules/ogm.py:160-162) and to cells (:170)
    const double2* z = reinterpret_cast<const double2*>(p.scans + (size_t)s * p.n_max * 2);
    int bx0 = sx, bx1 = sx, by0 = sy, by1 = sy;
    for (int i = tid; i < n_rays; i += kThreads) {
      const double2 q = z[i];
      // np.dot(z_t, R_wr.T) + [x, y]: row . column with R = [[c, -s], [s, c]]
      const double wx = __dadd_rn(__dadd_rn(__dmul_rn(q.x, c), __dmul_rn(q.y, -sn)), x);
      const double wy = __dadd_rn(__dadd_rn(__dmul_rn(q.x, sn), __dmul_rn(q.y, c)), y);
      const int ex = world2grid(wx, p.min_x, p.res), ey = world2grid(wy, p.min_y, p.res);
      ends[i] = make_int2(ex, ey);
      bx0 = min(bx0, ex); bx1 = max(bx1, ex); by0 = min(by0, ey); by1 = max(by1, ey);
    }
    atomicMin(&s_bbox[0], bx0); atomicMax(&s_bbox[1], bx1);
    atomicMin(&s_bbox[2], by0); atomicMax(&s_bbox[3], by1);
    __syncthreads();
    PH(0);
    // ---- phase 1: visit counts --------------------------------------------------------
    walk_rays<0>(ends, n_rays, sx, sy, W, H, scratch, nullptr, nullptr, nullptr, gtid, gthreads);
    __threadfence();
    cluster.sync();
    PH(1);
    const int wx0 = max(s_bbox[0], 0), wx1 = min(s_bbox[1], W - 1);
    const int wy0 = max(s_bbox[2], 0), wy1 = min(s_bbox[3], H - 1);
    const int wh = wy1 - wy0 + 1, ww = wx1 - wx0 + 1;
    const long long wcells = (ww > 0 && wh > 0) ? (long long)ww * wh : 0;
    for (;;) {
      if (gtid == 0) { ctl[0] = 0; ctl[1] = 0; ctl[2] = 0; }
      __threadfence();
      cluster.sync();
      // ---- phase 2: replay single-sign cells, queue mixed ones -------------------------
      auto handle = [&](int cell, int v) {
        if (v == 0) return;
        const int dec = v & 0xffff, inc = (v >> 16) & 0x7fff;
        if (dec == 0 || inc == 0) {
          float a = __ldcg(&grid[cell]);
          // the grid enters a scan inside [-20, 20] and every update moves by log 4 (up to an ulp):
          // after 30 same-sign updates the value is beyond the clip whatever it started from
          if (dec >= kSaturate) a = -20.0f;
          else if (inc >= kSaturate) a = 20.0f;
          else a = dec ? apply_updates(a, dec, -kLogOdds) : apply_updates(a, inc, kLogOdds);
          grid[cell] = clip20(a);                                   // modules/ogm.py:191
          scratch[cell] = 0;
        } else {
          const int n = dec + inc;
          const int off = atomicAdd(&ctl[1], n);
          if (off + n > p.ev_cap) { ctl[2] = 1; return; }
          const int slot = atomicAdd(&ctl[0], 1);
          if (slot >= kMixedCap) { ctl[2] = 1; return; }
          mixed_cell[slot] = cell; mixed_cnt[slot] = n; mixed_off[slot] = off; mixed_fill[slot] = 0;
          scratch[cell] = (int)(0x80000000u | (unsigned)slot);
        }
      };
      // four independent loads in flight per thread (the sweep is latency-bound otherwise)
      for (int t0 = gtid; t0 < (int)wcells; t0 += 4 * gthreads) {
        int cell[4], v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int t = t0 + u * gthreads;
          cell[u] = -1; v[u] = 0;
          if (t < (int)wcells) {
            cell[u] = (wx0 + t / wh) * H + wy0 + t % wh;
            v[u] = __ldcg(&scratch[cell[u]]);
          }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) handle(cell[u], v[u]);
      }
      __threadfence();
      cluster.sync();
      PH(2);
      const int n_mixed = min(__ldcg(&ctl[0]), kMixedCap);
      if (n_mixed > 0) {
        // ---- phase 3: ordered replay of the mixed cells -------------------------------
        walk_rays<1>(ends, n_rays, sx, sy, W, H, scratch, mixed_off, mixed_fill, events, gtid, gthreads);
        __threadfence();
        cluster.sync();
        PH(3);
        for (int slot = gtid; slot < n_mixed; slot += gthreads) {
          const int32_t* ev = events + mixed_off[slot];
          const int n = mixed_cnt[slot];
          const int cell = mixed_cell[slot];
          float a = __ldcg(&grid[cell]);
          if (n <= 16) {
            int e[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) e[i] = i < n ? __ldcg(&ev[i]) : 0x7fffffff;
            // rank sort: ray indices are distinct, so the rank of an event is its position
            for (int want = 0; want < n; ++want) {
              int pick = 0;
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                int rank = 0;
#pragma unroll
                for (int j = 0; j < 16; ++j) rank += (e[j] < e[i]) ? 1 : 0;
                if (rank == want) pick = e[i];
              }
              a = apply_updates(a, 1, (pick & 1) ? kLogOdds : -kLogOdds);
            }
          } else {
            int32_t* evw = events + mixed_off[slot];
            for (int i = 1; i < n; ++i) {                           // insertion sort by ray index
              const int key = evw[i];
              int j = i - 1;
              while (j >= 0 && evw[j] > key) { evw[j + 1] = evw[j]; --j; }
              evw[j + 1] = key;
            }
            for (int i = 0; i < n; ++i) a = apply_updates(a, 1, (evw[i] & 1) ? kLogOdds : -kLogOdds);
          }
          grid[cell] = clip20(a);
          scratch[cell] = 0;
        }
        __threadfence();
      }
      cluster.sync();                                               // scan fully applied, scratch clean again
      PH(4);
      const int again = __ldcg(&ctl[2]);
      if (!again) break;
      cluster.sync();                                               // ctl[] is reset at the loop top
    }
  }
#ifdef B200ICP_OGM_PROFILE
  if (gtid == 0) for (int i = 0; i < 6; ++i) ctl[4 + i] = (int)(ph[i] / 1000);   // microseconds per phase
#endif
}

__global__ void ogm_binarize_kernel(const float* __restrict__ log_odds, uint8_t* __restrict__ grid,
                                    long long n, float one_at_or_below) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) grid[i] = log_odds[i] <= one_at_or_below ? 1 : 0;
}

}  // namespace

extern "C" {

int b200icp_ogm_workspace_bytes(int width, int height, int n_max, int n_maps, size_t* bytes_out) {
  if (!bytes_out || width <= 0 || height <= 0 || n_max <= 0 || n_maps <= 0) return B200ICP_ERR_BAD_ARG;
  const size_t ev_cap = (size_t)n_max * 64 + 65536;
  const size_t per_map = (size_t)width * height + 16 + 4 * (size_t)kMixedCap + ev_cap;
  *bytes_out = per_map * sizeof(int32_t) * (size_t)n_maps;
  return B200ICP_OK;
}

int b200icp_ogm_build(float* log_odds, int width, int height, double res, double min_x, double min_y,
                      const double* poses, const double* cos_sin, const double* scans,
                      const int32_t* counts, int n_scans, int n_max, int n_maps, void* workspace,
                      size_t workspace_bytes, void* stream) {
  if (!log_odds || !poses || !cos_sin || !scans || !counts || width <= 0 || height <= 0 ||
      !(res > 0.0) || n_scans < 0 || n_max <= 0 || n_maps <= 0 || n_max > 32767)
    return B200ICP_ERR_BAD_ARG;
  if ((long long)width * height > 0x7fffffffLL) return B200ICP_ERR_TOO_LARGE;
  size_t need = 0;
  b200icp_ogm_workspace_bytes(width, height, n_max, n_maps, &need);
  if (!workspace || workspace_bytes < need) return B200ICP_ERR_WORKSPACE;
  if (n_scans == 0) return B200ICP_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (cudaMemsetAsync(workspace, 0, need, st) != cudaSuccess) return B200ICP_ERR_CUDA;
  OgmParams p{};
  p.log_odds = log_odds; p.width = width; p.height = height; p.res = res; p.min_x = min_x; p.min_y = min_y;
  p.poses = poses; p.cs = cos_sin; p.scans = scans; p.counts = counts;
  p.n_scans = n_scans; p.n_max = n_max; p.n_maps = n_maps;
  p.workspace = static_cast<int32_t*>(workspace);
  p.ev_cap = n_max * 64 + 65536;
  p.ws_stride = (long long)width * height + 16 + 4LL * kMixedCap + p.ev_cap;
  const size_t smem = (size_t)n_max * sizeof(int2);
  if (cudaFuncSetAttribute(ogm_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
    return B200ICP_ERR_TOO_LARGE;
  if (cudaFuncSetAttribute(ogm_build_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess)
    return B200ICP_ERR_CUDA;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)n_maps * kCluster);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr{};
  attr.id = cudaLaunchAttributeClusterDimension;
  attr.val.clusterDim.x = kCluster; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
  cfg.attrs = &attr; cfg.numAttrs = 1;
  if (cudaLaunchKernelEx(&cfg, ogm_build_kernel, p) != cudaSuccess) { cudaGetLastError(); return B200ICP_ERR_CUDA; }
  return cudaGetLastError() == cudaSuccess ? B200ICP_OK : B200ICP_ERR_CUDA;
}

int b200icp_ogm_binarize(const float* log_odds, uint8_t* grid, long long n_cells, float one_at_or_below,
                         void* stream) {
  if (!log_odds || !grid || n_cells < 0) return B200ICP_ERR_BAD_ARG;
  if (n_cells == 0) return B200ICP_OK;
  ogm_binarize_kernel<<<(unsigned)((n_cells + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      log_odds, grid, n_cells, one_at_or_below);
  return cudaGetLastError() == cudaSuccess ? B200ICP_OK : B200ICP_ERR_CUDA;
}

}  // extern "C"
