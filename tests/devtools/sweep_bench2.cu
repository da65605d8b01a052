// Developer microbenchmark 2: block-centred expanded-form NN sweep (2 FMA per candidate).
//   d(q,t) = |q-c|^2 + ( |t-c|^2 - 2 (q-c).(t-c) ),   c = centre of the 32/64-target block of t
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/sweep_bench2 tools/sweep_bench2.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <vector>

typedef unsigned long long u64;
__device__ __forceinline__ float fmin3(float a, float b, float c) {
  float r; asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r;
}
__device__ __forceinline__ u64 pk2(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void up2(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 sub2(u64 a, u64 b) { u64 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }

// smem layout: per group of 4 targets 12 floats [x'0..3][y'0..3][t2 0..3]; block centres
// [cx cx cy cy] per block after the targets.
// SETUP 0: packed per-block setup (sub2,sub2,mul2,fma2); 1: scalar setup + packs
template <int R, int W, int MINB, int BT, int SETUP, int MATH>
__global__ void __launch_bounds__(W * 32, MINB) sweep(const float* tg, int nt, int iters, float* out) {
  extern __shared__ __align__(16) float sm[];
  const int np = ((nt + BT - 1) / BT) * BT;
  const int nblk = np / BT;
  float* cb = sm + np * 3;
  for (int b = threadIdx.x; b < nblk; b += W * 32) {
    int i0 = min(b * BT + BT / 2, nt - 1);
    cb[4 * b] = cb[4 * b + 1] = tg[2 * i0]; cb[4 * b + 2] = cb[4 * b + 3] = tg[2 * i0 + 1];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < np; i += W * 32) {
    const int b = i / BT;
    float x = i < nt ? tg[2 * i] - cb[4 * b] : 1e18f, y = i < nt ? tg[2 * i + 1] - cb[4 * b + 2] : 1e18f;
    float* g = sm + (i / 4) * 12 + (i % 4);
    g[0] = -2.f * x; g[4] = -2.f * y; g[8] = x * x + y * y;
  }
  __syncthreads();
  float qx[R], qy[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    int j = (blockIdx.x * W * 32 + threadIdx.x) * R + r;
    qx[r] = tg[2 * (j % nt)] + 0.01f; qy[r] = tg[2 * (j % nt) + 1] - 0.01f;
  }
  unsigned accb = 0, accs = 0;
  const ulonglong2* tp = reinterpret_cast<const ulonglong2*>(sm);
  const ulonglong2* cp = reinterpret_cast<const ulonglong2*>(cb);
  for (int it = 0; it < iters; ++it) {
    unsigned best[R], second[R];
    u64 qxx[R], qyy[R];
#pragma unroll
    for (int r = 0; r < R; ++r) { best[r] = 0xffffffffu; second[r] = 0xffffffffu; qxx[r] = pk2(qx[r], qx[r]); qyy[r] = pk2(qy[r], qy[r]); }
    for (int b = 0; b < nblk; ++b) {
      float bm[R], q2[R];
      u64 dxx[R], dyy[R];
      const ulonglong2 C = cp[b];
#pragma unroll
      for (int r = 0; r < R; ++r) {
        bm[r] = __int_as_float(0x7f800000);
        if (SETUP == 0) {
          dxx[r] = sub2(qxx[r], C.x); dyy[r] = sub2(qyy[r], C.y);
          u64 qq = fma2(dxx[r], dxx[r], mul2(dyy[r], dyy[r]));
          float hi; up2(qq, q2[r], hi);
        } else {
          float cx, cy, t; up2(C.x, cx, t); up2(C.y, cy, t);
          float dx = __fsub_rn(qx[r], cx), dy = __fsub_rn(qy[r], cy);
          q2[r] = __fmaf_rn(dx, dx, __fmul_rn(dy, dy));
          dxx[r] = pk2(dx, dx); dyy[r] = pk2(dy, dy);
        }
      }
#pragma unroll
      for (int i = 0; i < BT / 4; ++i) {
        const ulonglong2 X = tp[(b * (BT / 4) + i) * 3], Y = tp[(b * (BT / 4) + i) * 3 + 1], T = tp[(b * (BT / 4) + i) * 3 + 2];
#pragma unroll
        for (int r = 0; r < R; ++r) {
          float d0, d1, d2, d3;
          if (MATH == 0) {
            const u64 a01 = fma2(dxx[r], X.x, fma2(dyy[r], Y.x, T.x));
            const u64 a23 = fma2(dxx[r], X.y, fma2(dyy[r], Y.y, T.y));
            up2(a01, d0, d1); up2(a23, d2, d3);
          } else {
            float x0, x1, x2, x3, y0, y1, y2, y3, t0, t1, t2, t3, sx, sy, tmp;
            up2(X.x, x0, x1); up2(X.y, x2, x3); up2(Y.x, y0, y1); up2(Y.y, y2, y3);
            up2(T.x, t0, t1); up2(T.y, t2, t3); up2(dxx[r], sx, tmp); up2(dyy[r], sy, tmp);
            if (MATH == 1) {
              d0 = __fmaf_rn(sx, x0, __fmaf_rn(sy, y0, t0)); d1 = __fmaf_rn(sx, x1, __fmaf_rn(sy, y1, t1));
            } else {
              const u64 a01 = fma2(dxx[r], X.x, fma2(dyy[r], Y.x, T.x)); up2(a01, d0, d1);
            }
            d2 = __fmaf_rn(sx, x2, __fmaf_rn(sy, y2, t2)); d3 = __fmaf_rn(sx, x3, __fmaf_rn(sy, y3, t3));
          }
          bm[r] = fmin3(fmin3(bm[r], d0, d1), d2, d3);
        }
      }
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const float v = fmaxf(__fadd_rn(bm[r], q2[r]), 0.f);
        const unsigned pk = (__float_as_uint(v) & 0xffffffc0u) | (unsigned)b;
        second[r] = min(second[r], max(best[r], pk));
        best[r] = min(best[r], pk);
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) { accb ^= best[r]; accs += second[r]; qx[r] += 1e-4f; qy[r] -= 1e-4f; }
  }
  out[blockIdx.x * W * 32 + threadIdx.x] = __uint_as_float(accb) + (float)accs;
}

__global__ void ffma_probe(float* sink, int iters) {
  const float a = 1.0f + 1e-7f * threadIdx.x, b = 1e-7f * (blockIdx.x + 1);
  float x[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) x[k] = (float)k;
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int rep = 0; rep < 4; ++rep)
#pragma unroll
      for (int k = 0; k < 16; ++k) x[k] = __fmaf_rn(x[k], a, b);
  float s = 0.f;
  for (int k = 0; k < 16; ++k) s += x[k];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}


// FMA2 / FFMA with all-distinct register operands (register-file bandwidth probe)
template <int PACKED>
__global__ void distinct_probe(float* sink, int iters) {
  u64 x[8], y[8], z[8];
  float xs[8], ys[8], zs[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    x[k] = pk2((float)k, k + 0.5f); y[k] = pk2(1.0f + 1e-7f * (threadIdx.x + k), 1.0f - 1e-7f * k); z[k] = pk2(1e-7f * (blockIdx.x + k), 1e-7f);
    xs[k] = (float)k; ys[k] = 1.0f + 1e-7f * (threadIdx.x + k); zs[k] = 1e-7f * (blockIdx.x + k);
  }
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int rep = 0; rep < 8; ++rep)
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (PACKED) x[k] = fma2(y[(k + rep) & 7], z[(k + 3 * rep + 1) & 7], x[k]);
        else xs[k] = __fmaf_rn(ys[(k + rep) & 7], zs[(k + 3 * rep + 1) & 7], xs[k]);
      }
  float s = 0.f;
  for (int k = 0; k < 8; ++k) { float lo, hi; up2(x[k], lo, hi); s += lo + hi + xs[k]; }
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
static float* d_out;
static void probes() {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int v = 0; v < 2; ++v) {
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
      cudaEventRecord(e0);
      if (v) distinct_probe<1><<<148 * 8, 256>>>(d_out, 20000); else distinct_probe<0><<<148 * 8, 256>>>(d_out, 20000);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep && ms < best) best = ms;
    }
    printf("distinct-operand %s: %.3e instr-lanes/s\n", v ? "FFMA2" : "FFMA", 148.0 * 8 * 256 * 20000 * 64 / (best * 1e-3));
  }
}
static float* d_t; static double g_peak;

template <int R, int W, int MINB, int BT, int SETUP, int MATH>
void run(int nt, int iters) {
  auto k = sweep<R, W, MINB, BT, SETUP, MATH>;
  const int np = ((nt + BT - 1) / BT) * BT;
  size_t smem = (size_t)np * 12 + (size_t)(np / BT) * 16;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int per = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k, W * 32, smem);
  cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, k);
  int grid = per * 148;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(e0); k<<<grid, W * 32, smem>>>(d_t, nt, iters, d_out); cudaEventRecord(e1);
    cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep && ms < best) best = ms;
  }
  double cand = (double)grid * W * 32 * R * (double)np * iters;
  double rate = cand * 4 / (best * 1e-3);
  printf("expanded math=%d R=%d W=%d BT=%d setup=%d regs=%d cta/sm=%d  %.3f ms  cand/s %.3e  alg_frac(4/cand) %.3f  exec_frac(2/cand) %.3f  err=%s\n",
         MATH, R, W, BT, SETUP, fa.numRegs, per, best, cand / (best * 1e-3), rate / g_peak, rate / 2 / g_peak,
         cudaGetErrorString(cudaGetLastError()));
}

int main() {
  const int nt = 1081;
  std::vector<float> h(2 * nt);
  for (int i = 0; i < nt; ++i) { float a = -2.356f + 4.712f * i / 1080.f; float r = 3.f + 2.f * sinf(5 * a); h[2 * i] = r * cosf(a); h[2 * i + 1] = r * sinf(a); }
  cudaMalloc(&d_t, h.size() * 4); cudaMemcpy(d_t, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
  cudaMalloc(&d_out, 148 * 32 * 1024 * 4);
  {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
      cudaEventRecord(e0); ffma_probe<<<148 * 8, 256>>>(d_out, 20000); cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep && ms < best) best = ms;
    }
    g_peak = 148.0 * 8 * 256 * 20000 * 64 / (best * 1e-3);
    printf("ffma probe: %.3e lane-FFMA/s\n", g_peak);
  }
  const int it = 40;
  run<5, 8, 3, 64, 1, 0>(nt, it);
  run<5, 8, 3, 64, 1, 1>(nt, it);
  run<5, 8, 3, 64, 1, 2>(nt, it);
  run<4, 8, 4, 64, 1, 0>(nt, it);
  run<4, 8, 4, 64, 1, 1>(nt, it);
  run<4, 8, 4, 64, 1, 2>(nt, it);
  run<6, 8, 3, 64, 1, 1>(nt, it);
  run<8, 8, 2, 64, 1, 1>(nt, it);
  run<8, 8, 3, 64, 1, 1>(nt, it);
  run<5, 4, 6, 64, 1, 1>(nt, it);
  run<5, 8, 3, 32, 1, 1>(nt, it);
  run<10, 8, 2, 64, 1, 1>(nt, it);
  probes();
  return 0;
}
