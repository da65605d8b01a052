// Float64 3x3 Kabsch solve on the device (one thread), replacing
// modules/icp.py:61-68:
//     U, s, Vt = svd(C);  R = V diag(1, 1, det(V U^T)) U^T
// with C = sum (x - xbar)(y - ybar)^T.  Because the determinant fix sits on the
// SMALLEST singular direction (LAPACK sorts singular values descending), the
// result only needs the two leading singular pairs:
//     R = v1 u1^T + v2 u2^T + (v1 x v2)(u1 x u2)^T
// which is what this computes, via a one-sided (Hestenes) Jacobi SVD: high
// relative accuracy, no library, ~5 sweeps.
#pragma once
#include <math.h>
#ifdef __CUDACC__
#include <cuda_runtime.h>
#define B200ICP_HD __host__ __device__
#else
#define B200ICP_HD
#endif

namespace b200icp {

B200ICP_HD inline void cross3(const double* a, const double* b, double* c) {
  c[0] = a[1] * b[2] - a[2] * b[1];
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
}

// C is row-major C[a*3+b] = sum X_a Y_b.  Writes the row-major rotation R (maps x onto y).
B200ICP_HD inline void kabsch_rotation_3x3(const double* C, double* R) {
  // columns of A (= C) and V, stored column-wise: a[c][row]
  double a[3][3], v[3][3];
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int r = 0; r < 3; ++r) { a[c][r] = C[r * 3 + c]; v[c][r] = (r == c) ? 1.0 : 0.0; }

  for (int sweep = 0; sweep < 20; ++sweep) {
    bool rotated = false;
#pragma unroll
    for (int pq = 0; pq < 3; ++pq) {
      const int p = (pq == 2) ? 1 : 0;
      const int q = (pq == 0) ? 1 : 2;
      const double alpha = a[p][0] * a[p][0] + a[p][1] * a[p][1] + a[p][2] * a[p][2];
      const double beta = a[q][0] * a[q][0] + a[q][1] * a[q][1] + a[q][2] * a[q][2];
      const double gamma = a[p][0] * a[q][0] + a[p][1] * a[q][1] + a[p][2] * a[q][2];
      if (gamma != 0.0 && fabs(gamma) > 1.0e-16 * sqrt(alpha * beta)) {
        rotated = true;
        const double zeta = (beta - alpha) / (2.0 * gamma);
        const double t = copysign(1.0, zeta) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double cs = 1.0 / sqrt(1.0 + t * t);
        const double sn = cs * t;
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          const double ap = a[p][r], aq = a[q][r];
          a[p][r] = cs * ap - sn * aq;
          a[q][r] = sn * ap + cs * aq;
          const double vp = v[p][r], vq = v[q][r];
          v[p][r] = cs * vp - sn * vq;
          v[q][r] = sn * vp + cs * vq;
        }
      }
    }
    if (!rotated) break;
  }
  double s[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) s[c] = a[c][0] * a[c][0] + a[c][1] * a[c][1] + a[c][2] * a[c][2];
  // indices of the two largest singular values
  int i1 = 0;
  if (s[1] > s[i1]) i1 = 1;
  if (s[2] > s[i1]) i1 = 2;
  int i2 = (i1 == 0) ? 1 : 0;
#pragma unroll
  for (int c = 0; c < 3; ++c)
    if (c != i1 && s[c] > s[i2]) i2 = c;
  double u1[3], u2[3], v1[3], v2[3], u3[3], v3[3];
  const double n1 = sqrt(s[i1]), n2 = sqrt(s[i2]);
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    // dynamic column selection without local-memory indexing
    const double a1 = (i1 == 0) ? a[0][r] : (i1 == 1) ? a[1][r] : a[2][r];
    const double a2 = (i2 == 0) ? a[0][r] : (i2 == 1) ? a[1][r] : a[2][r];
    v1[r] = (i1 == 0) ? v[0][r] : (i1 == 1) ? v[1][r] : v[2][r];
    v2[r] = (i2 == 0) ? v[0][r] : (i2 == 1) ? v[1][r] : v[2][r];
    u1[r] = a1 / n1;
    u2[r] = a2 / n2;
  }
  if (!(n1 > 0.0)) {  // zero covariance: nothing to rotate
#pragma unroll
    for (int k = 0; k < 9; ++k) R[k] = (k % 4 == 0) ? 1.0 : 0.0;
    return;
  }
  if (!(n2 > 0.0)) {
    // rank one (collinear data): the rotation about u1/v1 is undetermined.  Pick the
    // completion that maps an arbitrary unit normal of u1 to the matching normal of v1.
    double e[3] = {0.0, 0.0, 0.0};
    const int k = (fabs(u1[0]) <= fabs(u1[1]) && fabs(u1[0]) <= fabs(u1[2])) ? 0 : (fabs(u1[1]) <= fabs(u1[2]) ? 1 : 2);
    e[k] = 1.0;
    cross3(u1, e, u2);
    double nn = sqrt(u2[0] * u2[0] + u2[1] * u2[1] + u2[2] * u2[2]);
    for (int r = 0; r < 3; ++r) u2[r] /= nn;
    double f[3] = {0.0, 0.0, 0.0};
    const int m = (fabs(v1[0]) <= fabs(v1[1]) && fabs(v1[0]) <= fabs(v1[2])) ? 0 : (fabs(v1[1]) <= fabs(v1[2]) ? 1 : 2);
    f[m] = 1.0;
    cross3(v1, f, v2);
    nn = sqrt(v2[0] * v2[0] + v2[1] * v2[1] + v2[2] * v2[2]);
    for (int r = 0; r < 3; ++r) v2[r] /= nn;
  }
  cross3(u1, u2, u3);
  cross3(v1, v2, v3);
  // u1,u2 are orthogonal only to rounding; renormalise the cross products
  const double nu = sqrt(u3[0] * u3[0] + u3[1] * u3[1] + u3[2] * u3[2]);
  const double nv = sqrt(v3[0] * v3[0] + v3[1] * v3[1] + v3[2] * v3[2]);
#pragma unroll
  for (int r = 0; r < 3; ++r) { u3[r] /= nu; v3[r] /= nv; }
  // x is rotated onto y: R = sum_k (right vector k)(left vector k)^T, since C = U S V^T
  // has left vectors in x-space and right vectors in y-space.
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c) R[r * 3 + c] = v1[r] * u1[c] + v2[r] * u2[c] + v3[r] * u3[c];
}

}  // namespace b200icp
