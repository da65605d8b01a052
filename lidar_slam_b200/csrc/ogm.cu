// Occupancy-grid ray casting on the device: the consumer of the
// poses the ICP path produces.
//
//   ogm_count/classify/event/apply   OccupancyGridMap.update_map / build_map   modules/ogm.py:149-195, :46-60
//                                    OccupancyGridMap.bresenham2D              modules/ogm.py:197-231
//                                    OccupancyGridMap.world2grid               modules/ogm.py:104-126
//   ogm_binarize                     the pmf thresholding of build_map         modules/ogm.py:62-65
//
// The reference updates a float32 log-odds grid one ray after another:
//     cell -= log 4  for every in-bounds cell of the ray but the last,  cell += log 4 for the last,
// each update being  f32(f64(cell) -/+ log 4)  (numpy in-place op with a float64 scalar), and
// clips the grid to [-20, 20] after every scan.  Rounding and clipping make the value of a cell
// depend on the ORDER of the updates it receives -- but only on the updates of THAT cell.  So
// the chain "scan after scan, ray after ray" is kept per cell and everything else runs in
// parallel over the whole GPU, a BATCH of up to 64 scans (of every map) at a time:
//   count     one thread per ray of every scan of the batch (the in-bounds part of a ray is
//             entered through the closed form of the reference's Bresenham variant); every
//             visited cell gets its decrement / increment COUNT with a fire-and-forget atomic on
//             the int32 count plane of its scan;
//   classify  the bounding window of every scan's plane is swept; a (scan, cell) that saw both
//             decrements and increments is given an event list;
//   event     the rays are walked again and append (ray, sign) to the tagged cells;
//   apply     one thread per cell walks the batch scan by scan: single-sign counts replay that
//             many rounded updates (the order among identical updates is immaterial), tagged
//             entries sort their events by ray index and replay them in the reference's order;
//             clip after every scan; the planes are left zeroed for the next batch.
// Event storage is sized for the usual case; an entry that does not fit is replayed by testing
// every ray of its scan against the cell with the closed form (slow, exact, rare).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include <mutex>

#include "../../include/b200icp.h"

namespace {

constexpr int kThreads = 256;
constexpr int kBatchMax = 64;                     // scans per batch (count planes per map)
constexpr size_t kPlaneBudget = (size_t)48 << 20; // bytes of count planes per map and buffer set that set the batch size
constexpr int kEventsPerRay = 16;                 // event words per ray of the batch (average capacity)
constexpr unsigned kOverflowTag = 0xffffffffu;
constexpr double kLogOdds = 1.3862943611198906;   // np.log(4.)  (modules/ogm.py:44)

struct OgmParams {
  float* log_odds;        // [n_maps][width*height]
  int width, height;
  double res, min_x, min_y;
  const double* poses;    // [n_maps][n_scans][3]  (x, y, yaw)
  const double* cs;       // [n_maps][n_scans][2]  (cos yaw, sin yaw) as numpy evaluates them
  const double* scans;    // [n_scans][n_max][2] shared by all maps
  const int32_t* counts;  // [n_scans]
  int n_scans, n_max, n_maps;
  int s0, sb;             // this batch: scans [s0, s0 + sb)
  int batch;              // planes per map (capacity)
  int32_t* planes;        // [n_maps][batch][width*height]  dec count | inc count << 16, or tag
  int32_t* bbox;          // [n_maps][batch][4]  min x, min -x, min y, min -y of the visited window
  int32_t* ctl;           // [n_maps][4]  event lists made, event words used
  int32_t* events;        // [n_maps][ev_cap + 32]  lists: [fill counter == n after the event pass][increments][n events]
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
    const int m = dx > 0 ? (int)(((unsigned long long)k * (unsigned long long)dy + (unsigned long long)r.c0) / (unsigned long long)dx) : 0;   // 64-bit: k*dy wraps 32 bits for end points far outside a large grid
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


// Pose-dependent data of one scan of one map.
struct ScanFrame {
  double x, y, c, sn;
  int sx, sy;        // start cell: pose + lidar offset, NOT rotated (modules/ogm.py:165-166)
  int n_rays;
};

__device__ __forceinline__ ScanFrame scan_frame(const OgmParams& p, int map, int s) {
  ScanFrame f;
  const double* pose = p.poses + ((size_t)map * p.n_scans + s) * 3;
  const double* cs = p.cs + ((size_t)map * p.n_scans + s) * 2;
  f.x = pose[0]; f.y = pose[1]; f.c = cs[0]; f.sn = cs[1];
  f.sx = world2grid(__dadd_rn(f.x, 0.13323), p.min_x, p.res);
  f.sy = world2grid(__dadd_rn(f.y, 0.0), p.min_y, p.res);
  f.n_rays = min(p.counts[s], p.n_max);
  return f;
}

// End cell of ray i: the end point to the world frame (modules/ogm.py:160-162:
// np.dot(z_t, R_wr.T) + [x, y], row . column with R = [[c, -s], [s, c]]) and to cells (:170).
__device__ __forceinline__ int2 ray_end(const OgmParams& p, const ScanFrame& f, int s, int i) {
  const double2 q = reinterpret_cast<const double2*>(p.scans + (size_t)s * p.n_max * 2)[i];
  const double wx = __dadd_rn(__dadd_rn(__dmul_rn(q.x, f.c), __dmul_rn(q.y, -f.sn)), f.x);
  const double wy = __dadd_rn(__dadd_rn(__dmul_rn(q.x, f.sn), __dmul_rn(q.y, f.c)), f.y);
  return make_int2(world2grid(wx, p.min_x, p.res), world2grid(wy, p.min_y, p.res));
}

// Walk one ray, classic incremental stepping (rem is the numerator of the closed form modulo
// dx, m its quotient).  MODE 0 counts visits on the plane, MODE 1 appends (ray, sign) events to
// the tagged cells.
template <int MODE>
__device__ __forceinline__ void walk_ray(const Ray& r, int ray, int height, int32_t* plane, int32_t* events) {
  if (r.klo > r.khi) return;
  const int dx = r.dx, dy = r.dy;
  int m = 0, rem = r.c0;
  if (r.klo > 0) {
    const unsigned long long num = (unsigned long long)r.klo * (unsigned long long)dy + (unsigned long long)r.c0;
    m = (int)(num / (unsigned long long)dx);
    rem = (int)(num - (unsigned long long)m * (unsigned long long)dx);
  }
  // the cell index advances by a fixed stride per major step and another per minor step
  const int smaj = r.steep ? (r.sy <= r.ey ? 1 : -1) : (r.sx <= r.ex ? height : -height);
  const int smin = r.steep ? (r.sx <= r.ex ? height : -height) : (r.sy <= r.ey ? 1 : -1);
  int cell = r.sx * height + r.sy + r.klo * smaj + m * smin;
  if (MODE == 0) {
    for (int k = r.klo; k <= r.khi; ++k) {
      atomicAdd(&plane[cell], k == r.khi ? 0x10000 : 1);
      rem += dy;
      cell += smaj;
      if (rem >= dx && dx > 0) { rem -= dx; cell += smin; }
    }
  } else {
    // the tag loads are independent of one another: issue eight before looking at any
    constexpr int U = 8;
    for (int k0 = r.klo; k0 <= r.khi; k0 += U) {
      int tags[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        tags[u] = (k0 + u <= r.khi) ? __ldcg(&plane[cell]) : 0;
        rem += dy;
        cell += smaj;
        if (rem >= dx && dx > 0) { rem -= dx; cell += smin; }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (tags[u] < 0 && (unsigned)tags[u] != kOverflowTag) {
          int32_t* list = events + (tags[u] & 0x7fffffff);
          const int pos = atomicAdd(&list[0], 1);
          list[2 + pos] = (ray << 1) | ((k0 + u == r.khi) ? 1 : 0);
        }
      }
    }
  }
}

// Does ray r visit cell (cx, cy) (a cell inside the grid)?  Returns 0 (no), -1 (decrement) or
// +1 (increment: the last in-bounds cell of the ray).  Closed form of the walk above.
__device__ __forceinline__ int ray_visit(const Ray& r, int cx, int cy) {
  if (r.klo > r.khi) return 0;
  const int k = r.steep ? (r.sy <= r.ey ? cy - r.sy : r.sy - cy) : (r.sx <= r.ex ? cx - r.sx : r.sx - cx);
  if (k < r.klo || k > r.khi) return 0;
  const int m = r.dx > 0 ? (int)(((unsigned long long)k * (unsigned long long)r.dy + (unsigned long long)r.c0) / (unsigned long long)r.dx) : 0;
  const int minor = r.steep ? r.sx + (r.sx <= r.ex ? m : -m) : r.sy + (r.sy <= r.ey ? m : -m);
  if (minor != (r.steep ? cx : cy)) return 0;
  return k == r.khi ? 1 : -1;
}

__global__ void ogm_reset_kernel(const OgmParams p) {
  const int n_bbox = p.n_maps * p.batch * 4, n_ctl = p.n_maps * 4;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_bbox + n_ctl; i += gridDim.x * blockDim.x) {
    if (i < n_bbox) p.bbox[i] = 0x7fffffff;
    else p.ctl[i - n_bbox] = 0;
  }
}

__device__ __forceinline__ int warp_min_i(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// count: grid (ray blocks, scans of the batch, maps)
__global__ void __launch_bounds__(kThreads) ogm_count_kernel(const OgmParams p) {
  const int sl = blockIdx.y, map = blockIdx.z, s = p.s0 + sl;
  const ScanFrame f = scan_frame(p, map, s);
  if (blockIdx.x * kThreads >= max(f.n_rays, 1)) return;
  const size_t cells = (size_t)p.width * p.height;
  int32_t* plane = p.planes + ((size_t)map * p.batch + sl) * cells;
  const int ray = blockIdx.x * kThreads + threadIdx.x;
  // window of the scan: start cell and every end cell (a Bresenham line stays inside the box of
  // its two ends); kept as minima of x, -x, y, -y
  int b0 = f.sx, b1 = -f.sx, b2 = f.sy, b3 = -f.sy;
  if (ray < f.n_rays) {
    const int2 e = ray_end(p, f, s, ray);
    b0 = min(b0, e.x); b1 = min(b1, -e.x); b2 = min(b2, e.y); b3 = min(b3, -e.y);
    const Ray r = make_ray(f.sx, f.sy, e.x, e.y, p.width, p.height);
    walk_ray<0>(r, ray, p.height, plane, nullptr);
  }
  b0 = warp_min_i(b0); b1 = warp_min_i(b1); b2 = warp_min_i(b2); b3 = warp_min_i(b3);
  if ((threadIdx.x & 31) == 0) {
    int32_t* bb = p.bbox + ((size_t)map * p.batch + sl) * 4;
    atomicMin(&bb[0], b0); atomicMin(&bb[1], b1); atomicMin(&bb[2], b2); atomicMin(&bb[3], b3);
  }
}

struct Window { int x0, x1, y0, y1; };
__device__ __forceinline__ Window scan_window(const OgmParams& p, int map, int sl) {
  const int32_t* bb = p.bbox + ((size_t)map * p.batch + sl) * 4;
  Window w;
  w.x0 = max(bb[0], 0); w.x1 = min(-bb[1], p.width - 1);
  w.y0 = max(bb[2], 0); w.y1 = min(-bb[3], p.height - 1);
  return w;
}

// classify: grid (cell blocks, scans of the batch, maps)
__global__ void __launch_bounds__(kThreads) ogm_classify_kernel(const OgmParams p) {
  const int sl = blockIdx.y, map = blockIdx.z;
  const Window w = scan_window(p, map, sl);
  const int wh = w.y1 - w.y0 + 1, ww = w.x1 - w.x0 + 1;
  if (wh <= 0 || ww <= 0) return;
  const int wcells = ww * wh;
  const size_t cells = (size_t)p.width * p.height;
  int32_t* plane = p.planes + ((size_t)map * p.batch + sl) * cells;
  int32_t* ctl = p.ctl + map * 4;
  int32_t* events = p.events + (size_t)map * (p.ev_cap + 32);
  // four independent loads in flight per thread; the list area is reserved with ONE atomic per
  // warp (prefix sum over the lanes that need a list): the counter is a single hot address
  const int lane = threadIdx.x & 31;
  const int stride = gridDim.x * kThreads;
  for (int t0 = blockIdx.x * kThreads + threadIdx.x; t0 - (int)threadIdx.x < wcells; t0 += 4 * stride) {
    int cell[4], v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int t = t0 + u * stride;
      cell[u] = -1; v[u] = 0;
      if (t < wcells) {
        cell[u] = (w.x0 + t / wh) * p.height + w.y0 + t % wh;
        v[u] = __ldcg(&plane[cell[u]]);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int dec = v[u] & 0xffff, inc = (v[u] >> 16) & 0x7fff;
      const bool mixed = dec != 0 && inc != 0;
      const unsigned need = __ballot_sync(0xffffffffu, mixed);
      if (need == 0) continue;
      const int words = mixed ? dec + inc + 2 : 0;       // fill counter, increment count, events
      int pre = words;                                    // inclusive prefix sum over the warp
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, pre, o);
        if (lane >= o) pre += up;
      }
      const int total = __shfl_sync(0xffffffffu, pre, 31);
      unsigned base = 0;
      if (lane == 0) { base = (unsigned)atomicAdd(&ctl[1], total); ctl[0] = 1; }
      base = __shfl_sync(0xffffffffu, base, 0);
      if (mixed) {
        const unsigned o = base + (unsigned)(pre - words);
        unsigned tag = kOverflowTag;
        if (o <= (unsigned)p.ev_cap && o + (unsigned)words <= (unsigned)p.ev_cap) {
          events[o] = 0;
          events[o + 1] = inc;
          tag = 0x80000000u | o;
        }
        plane[cell[u]] = (int)tag;
      }
    }
  }
}

// event: same grid as count
__global__ void __launch_bounds__(kThreads) ogm_event_kernel(const OgmParams p) {
  const int sl = blockIdx.y, map = blockIdx.z, s = p.s0 + sl;
  if (__ldcg(&p.ctl[map * 4]) == 0) return;          // no event list in this batch
  const ScanFrame f = scan_frame(p, map, s);
  const int ray = blockIdx.x * kThreads + threadIdx.x;
  if (ray >= f.n_rays) return;
  const size_t cells = (size_t)p.width * p.height;
  int32_t* plane = p.planes + ((size_t)map * p.batch + sl) * cells;
  const int2 e = ray_end(p, f, s, ray);
  const Ray r = make_ray(f.sx, f.sy, e.x, e.y, p.width, p.height);
  walk_ray<1>(r, ray, p.height, plane, p.events + (size_t)map * (p.ev_cap + 32));
}

// Replay the n <= N events of one (scan, cell) in ray order: a bitonic network in registers
// sorts the (ray << 1 | sign) words (ray indices are distinct; the padding sorts last).
template <int N>
__device__ __forceinline__ float replay_sorted(float a, const int32_t* ev, int n, const int (&first)[4]) {
  int e[N];
#pragma unroll
  for (int i = 0; i < N; ++i) {
    const int w = i < 4 ? first[i] : __ldcg(&ev[i]);    // the list area is padded: reads past n are harmless
    e[i] = i < n ? w : 0x7fffffff;
  }
#pragma unroll
  for (int k = 2; k <= N; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
#pragma unroll
      for (int i = 0; i < N; ++i) {
        const int l = i ^ j;
        if (l > i) {
          const int lo = min(e[i], e[l]), hi = max(e[i], e[l]);
          if ((i & k) == 0) { e[i] = lo; e[l] = hi; } else { e[i] = hi; e[l] = lo; }
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < N; ++i)
    if (i < n) a = apply_updates(a, 1, (e[i] & 1) ? kLogOdds : -kLogOdds);
  return a;
}

// Long lists with few increments: only the interleaving of the signs matters, so the increments
// (<= 16) are sorted and every decrement is counted into the gap between two of them.
__device__ __forceinline__ float replay_by_gaps(float a, const int32_t* ev, int n) {
  int incr[16], gap[17];
#pragma unroll
  for (int j = 0; j < 16; ++j) incr[j] = 0x7fffffff;
#pragma unroll
  for (int j = 0; j < 17; ++j) gap[j] = 0;
  int ni = 0;
  for (int i = 0; i < n; ++i) {
    const int w = __ldcg(&ev[i]);
    if (w & 1) {
#pragma unroll
      for (int j = 0; j < 16; ++j) if (j == ni) incr[j] = w;
      ++ni;
    }
  }
#pragma unroll
  for (int k = 2; k <= 16; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const int l = i ^ j;
        if (l > i) {
          const int lo = min(incr[i], incr[l]), hi = max(incr[i], incr[l]);
          if ((i & k) == 0) { incr[i] = lo; incr[l] = hi; } else { incr[i] = hi; incr[l] = lo; }
        }
      }
    }
  }
  for (int i = 0; i < n; ++i) {
    const int w = __ldcg(&ev[i]);
    if (!(w & 1)) {
      int rank = 0;
#pragma unroll
      for (int j = 0; j < 16; ++j) rank += (incr[j] < w) ? 1 : 0;
#pragma unroll
      for (int j = 0; j < 17; ++j) gap[j] += (rank == j) ? 1 : 0;
    }
  }
#pragma unroll
  for (int j = 0; j < 17; ++j) {
    a = apply_updates(a, gap[j], -kLogOdds);
    if (j < ni) a = apply_updates(a, 1, kLogOdds);
  }
  return a;
}

// apply: grid (cell blocks, 1, maps); one thread per cell of the union of the batch's windows
__global__ void __launch_bounds__(kThreads) ogm_apply_kernel(const OgmParams p) {
  __shared__ Window win[kBatchMax];
  __shared__ int uni[4];
  const int map = blockIdx.z, tid = threadIdx.x;
  if (tid == 0) { uni[0] = 0x7fffffff; uni[1] = -1; uni[2] = 0x7fffffff; uni[3] = -1; }
  __syncthreads();
  if (tid < p.sb) {
    const Window w = scan_window(p, map, tid);
    win[tid] = w;
    if (w.x0 <= w.x1 && w.y0 <= w.y1) {
      atomicMin(&uni[0], w.x0); atomicMax(&uni[1], w.x1);
      atomicMin(&uni[2], w.y0); atomicMax(&uni[3], w.y1);
    }
  }
  __syncthreads();
  const int ux0 = uni[0], ux1 = uni[1], uy0 = uni[2], uy1 = uni[3];
  const int uh = uy1 - uy0 + 1, uw = ux1 - ux0 + 1;
  if (uh <= 0 || uw <= 0) return;
  const int ucells = uw * uh;
  const size_t cells = (size_t)p.width * p.height;
  float* grid = p.log_odds + (size_t)map * cells;
  int32_t* planes = p.planes + (size_t)map * p.batch * cells;
  int32_t* events = p.events + (size_t)map * (p.ev_cap + 32);

  for (int t = blockIdx.x * kThreads + tid; t < ucells; t += gridDim.x * kThreads) {
    const int cx = ux0 + t / uh, cy = uy0 + t % uh;
    const int cell = cx * p.height + cy;
    float a = 0.f;
    bool loaded = false;
    for (int s8 = 0; s8 < p.sb; s8 += 8) {
      int v[8];
      int any = 0;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        v[u] = 0;
        const int sl = s8 + u;
        if (sl < p.sb) {
          const Window w = win[sl];
          if (cx >= w.x0 && cx <= w.x1 && cy >= w.y0 && cy <= w.y1) v[u] = __ldcg(&planes[(size_t)sl * cells + cell]);
        }
        any |= v[u];
      }
      if (!any) continue;
      if (!loaded) { a = __ldcg(&grid[cell]); loaded = true; }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        if (v[u] == 0) continue;
        const int sl = s8 + u;
        if (v[u] > 0) {
          // only decrements or only increments: the order among identical updates is immaterial.
          // The clip comes after the scan, so when  a -/+ k log 4  lands beyond it by more than the
          // k roundings can move it (< k * 2^-18), the result is the clip value whatever the
          // roundings were -- the usual case for the cells around the robot, which sit at -20.
          const int dec = v[u] & 0xffff, inc = (v[u] >> 16) & 0x7fff;
          const double est = dec ? (double)a - (double)dec * kLogOdds : (double)a + (double)inc * kLogOdds;
          if (est < -20.5) a = -20.0f;
          else if (est > 20.5) a = 20.0f;
          else a = dec ? apply_updates(a, dec, -kLogOdds) : apply_updates(a, inc, kLogOdds);
        } else if ((unsigned)v[u] == kOverflowTag) {
          // no room for an event list: test every ray of the scan against this cell, in ray order
          const int s = p.s0 + sl;
          const ScanFrame f = scan_frame(p, map, s);
          for (int i = 0; i < f.n_rays; ++i) {
            const int2 e = ray_end(p, f, s, i);
            const int hit = ray_visit(make_ray(f.sx, f.sy, e.x, e.y, p.width, p.height), cx, cy);
            if (hit) a = apply_updates(a, 1, hit > 0 ? kLogOdds : -kLogOdds);
          }
        } else {
          // counter, increments and the first four events in one round trip (most lists are short)
          int32_t* list = events + (v[u] & 0x7fffffff);
          const int n = __ldcg(&list[0]), inc = __ldcg(&list[1]), dec = n - inc;
          const int first[4] = {__ldcg(&list[2]), __ldcg(&list[3]), __ldcg(&list[4]), __ldcg(&list[5])};
          int32_t* ev = list + 2;
          // The clip comes after the scan: when  a + (inc - dec) log 4  lands beyond it by more than
          // the n roundings can move it (each < |value| * 2^-24, |value| <= 20 + 1.39 n), the result
          // is the clip value whatever the order was.
          const double nn = (double)n;
          const double margin = 0.5 + nn * (20.0 + 1.4 * nn) * 5.9604644775390625e-08;
          const double est = (double)a + (double)(inc - dec) * kLogOdds;
          if (est < -20.0 - margin) {
            a = -20.0f;
          } else if (est > 20.0 + margin) {
            a = 20.0f;
          } else if (n <= 4) {
            a = replay_sorted<4>(a, ev, n, first);
          } else if (n <= 8) {
            a = replay_sorted<8>(a, ev, n, first);
          } else if (n <= 16) {
            a = replay_sorted<16>(a, ev, n, first);
          } else if (inc <= 16) {
            a = replay_by_gaps(a, ev, n);
          } else {
            for (int i = 1; i < n; ++i) {                           // insertion sort by ray index
              const int key = ev[i];
              int k = i - 1;
              while (k >= 0 && ev[k] > key) { ev[k + 1] = ev[k]; --k; }
              ev[k + 1] = key;
            }
            for (int i = 0; i < n; ++i) a = apply_updates(a, 1, (ev[i] & 1) ? kLogOdds : -kLogOdds);
          }
        }
        a = clip20(a);                                              // modules/ogm.py:191, after every scan
        planes[(size_t)sl * cells + cell] = 0;                      // clean for the next batch
      }
    }
    if (loaded) grid[cell] = a;
  }
}

__global__ void ogm_binarize_kernel(const float* __restrict__ log_odds, uint8_t* __restrict__ grid,
                                    long long n, float one_at_or_below) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) grid[i] = log_odds[i] <= one_at_or_below ? 1 : 0;
}

struct OgmLayout {
  int batch, ev_cap;
  size_t off_planes, off_bbox, off_ctl, off_events, set_stride, total;   // in int32 elements
};

// Two buffer sets: while the apply pass walks batch b on the auxiliary stream, the ray passes of
// batch b + 1 fill the other set on the caller's stream (every pass alone leaves most SMs idle).
struct OgmStreams {
  cudaStream_t aux = nullptr;
  cudaEvent_t ready[2] = {nullptr, nullptr}, done[2] = {nullptr, nullptr};
};
OgmStreams* ogm_streams() {
  static OgmStreams per_device[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  OgmStreams& s = per_device[dev];
  if (!s.aux) {
    if (cudaStreamCreateWithFlags(&s.aux, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    for (int k = 0; k < 2; ++k) {
      if (cudaEventCreateWithFlags(&s.ready[k], cudaEventDisableTiming) != cudaSuccess) return nullptr;
      if (cudaEventCreateWithFlags(&s.done[k], cudaEventDisableTiming) != cudaSuccess) return nullptr;
    }
  }
  return &s;
}

int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  if (!e || !*e) return dflt;
  const int v = atoi(e);
  return v < 0 ? dflt : v;
}

OgmLayout ogm_layout(int width, int height, int n_max, int n_maps) {
  OgmLayout L;
  const size_t cells = (size_t)width * height;
  size_t b = kPlaneBudget / (cells * sizeof(int32_t));
  L.batch = (int)(b < 1 ? 1 : (b > (size_t)kBatchMax ? (size_t)kBatchMax : b));
  // capacity of the event lists (tuning / test knob: the overflow path is exact whatever it is)
  const long long cap = (long long)L.batch * n_max * env_int("B200ICP_OGM_EVENTS_PER_RAY", kEventsPerRay);
  L.ev_cap = (int)(cap < 1 ? 1 : (cap > 0x3fffffffLL ? 0x3fffffffLL : cap));
  size_t o = 0;
  auto take = [&](size_t n) { const size_t at = o; o += (n + 31) / 32 * 32; return at; };
  L.off_planes = take((size_t)n_maps * L.batch * cells);
  L.off_bbox = take((size_t)n_maps * L.batch * 4);
  L.off_ctl = take((size_t)n_maps * 4);
  L.off_events = take((size_t)n_maps * ((size_t)L.ev_cap + 32));
  L.set_stride = o;
  L.total = 2 * o;
  return L;
}

}  // namespace

extern "C" {

int b200icp_ogm_workspace_bytes(int width, int height, int n_max, int n_maps, size_t* bytes_out) {
  if (!bytes_out || width <= 0 || height <= 0 || n_max <= 0 || n_maps <= 0 || n_max > 32767)
    return B200ICP_ERR_BAD_ARG;
  if ((long long)width * height > 0x7fffffffLL) return B200ICP_ERR_TOO_LARGE;
  *bytes_out = ogm_layout(width, height, n_max, n_maps).total * sizeof(int32_t);
  return B200ICP_OK;
}

int b200icp_ogm_build(float* log_odds, int width, int height, double res, double min_x, double min_y,
                      const double* poses, const double* cos_sin, const double* scans,
                      const int32_t* counts, int n_scans, int n_max, int n_maps, void* workspace,
                      size_t workspace_bytes, void* stream) {
  if (!log_odds || !poses || !cos_sin || !scans || !counts || width <= 0 || height <= 0 ||
      !(res > 0.0) || n_scans < 0 || n_max <= 0 || n_maps <= 0 || n_max > 32767 || n_maps > 65535)
    return B200ICP_ERR_BAD_ARG;
  if ((long long)width * height > 0x7fffffffLL) return B200ICP_ERR_TOO_LARGE;
  const OgmLayout L = ogm_layout(width, height, n_max, n_maps);
  if (!workspace || workspace_bytes < L.total * sizeof(int32_t)) return B200ICP_ERR_WORKSPACE;
  if (n_scans == 0) return B200ICP_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int32_t* ws = static_cast<int32_t*>(workspace);
  const size_t cells = (size_t)width * height;
  static std::mutex enqueue_mutex;      // the auxiliary stream and its events are shared per device
  std::lock_guard<std::mutex> lock(enqueue_mutex);
  OgmStreams* streams = ogm_streams();
  if (!streams) return B200ICP_ERR_CUDA;
  // zero the count planes this call will use (the apply pass leaves them zeroed again): all of a
  // set when the scans fill a batch, otherwise only the leading planes of every map (update_map
  // with one scan then clears one plane, not sixty-four)
  const int sets_used = n_scans > L.batch ? 2 : 1;
  const int planes_used = n_scans < L.batch ? n_scans : L.batch;
  for (int k = 0; k < sets_used; ++k) {
    int32_t* planes = ws + k * L.set_stride + L.off_planes;
    if (planes_used == L.batch) {
      if (cudaMemsetAsync(planes, 0, (size_t)n_maps * L.batch * cells * sizeof(int32_t), st) != cudaSuccess) return B200ICP_ERR_CUDA;
    } else {
      for (int m = 0; m < n_maps; ++m)
        if (cudaMemsetAsync(planes + (size_t)m * L.batch * cells, 0, (size_t)planes_used * cells * sizeof(int32_t), st) != cudaSuccess)
          return B200ICP_ERR_CUDA;
    }
  }
  OgmParams p{};
  p.log_odds = log_odds; p.width = width; p.height = height; p.res = res; p.min_x = min_x; p.min_y = min_y;
  p.poses = poses; p.cs = cos_sin; p.scans = scans; p.counts = counts;
  p.n_scans = n_scans; p.n_max = n_max; p.n_maps = n_maps; p.batch = L.batch; p.ev_cap = L.ev_cap;
  const unsigned ray_blocks = (unsigned)((n_max + kThreads - 1) / kThreads);
  const size_t cell_blocks = (cells + kThreads - 1) / kThreads;
  const unsigned sweep_blocks = (unsigned)(cell_blocks < 64 ? cell_blocks : 64);
  const unsigned apply_blocks = (unsigned)(cell_blocks < 148 * 8 ? cell_blocks : 148 * 8);
  int b = 0;
  for (int s0 = 0; s0 < n_scans; s0 += L.batch, ++b) {
    const int k = b & 1;
    int32_t* set = ws + k * L.set_stride;
    p.planes = set + L.off_planes; p.bbox = set + L.off_bbox; p.ctl = set + L.off_ctl; p.events = set + L.off_events;
    p.s0 = s0;
    p.sb = n_scans - s0 < L.batch ? n_scans - s0 : L.batch;
    if (b >= 2 && cudaStreamWaitEvent(st, streams->done[k], 0) != cudaSuccess) return B200ICP_ERR_CUDA;   // set k is free again
    ogm_reset_kernel<<<1, kThreads, 0, st>>>(p);
    ogm_count_kernel<<<dim3(ray_blocks, (unsigned)p.sb, (unsigned)n_maps), kThreads, 0, st>>>(p);
    ogm_classify_kernel<<<dim3(sweep_blocks, (unsigned)p.sb, (unsigned)n_maps), kThreads, 0, st>>>(p);
    ogm_event_kernel<<<dim3(ray_blocks, (unsigned)p.sb, (unsigned)n_maps), kThreads, 0, st>>>(p);
    if (cudaEventRecord(streams->ready[k], st) != cudaSuccess) return B200ICP_ERR_CUDA;
    if (cudaStreamWaitEvent(streams->aux, streams->ready[k], 0) != cudaSuccess) return B200ICP_ERR_CUDA;
    ogm_apply_kernel<<<dim3(apply_blocks, 1, (unsigned)n_maps), kThreads, 0, streams->aux>>>(p);   // batches in order: one stream
    if (cudaEventRecord(streams->done[k], streams->aux) != cudaSuccess) return B200ICP_ERR_CUDA;
  }
  // join: the caller's stream continues after the last apply pass (the earlier ones precede it on aux)
  if (cudaStreamWaitEvent(st, streams->done[(b - 1) & 1], 0) != cudaSuccess) return B200ICP_ERR_CUDA;
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
