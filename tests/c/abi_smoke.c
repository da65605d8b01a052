/* A plain C caller of the drop-in boundary (no Python, no CUDA headers): builds two planar scans
 * related by a known rigid motion, aligns them through b200icp_run_host and checks the recovered
 * transform.  Compiled and run by tests/test_gpu_hostapi.py; compiled and linked (not run) by
 * tests/test_host.py.  Exit code 0 = pass. */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "b200icp.h"

#define N 400

int main(void) {
  static double pool[2][N][2];
  int32_t counts[2] = {N, N};
  const double th = 0.002, tx = 0.003, ty = -0.002;   /* below half the sample spacing: every nearest neighbour is the true match */
  /* target: an L-shaped wall sampled every centimetre; source = target moved by the INVERSE motion,
   * so that aligning source onto target recovers (th, tx, ty) */
  for (int i = 0; i < N; ++i) {
    const double s = 0.01 * i;
    const double x = i < N / 2 ? s : 0.01 * (N / 2), y = i < N / 2 ? 1.5 : 1.5 - (s - 0.01 * (N / 2));
    pool[0][i][0] = x; pool[0][i][1] = y;
    const double dx = x - tx, dy = y - ty;                 /* R^T (p - t) */
    pool[1][i][0] = cos(th) * dx + sin(th) * dy;
    pool[1][i][1] = -sin(th) * dx + cos(th) * dy;
  }
  int32_t pairs[2] = {1, 0};                               /* source scan 1 onto target scan 0 */
  double init[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
  double eps = 1e-12;
  uint8_t normalize = 0;
  double T[16], err = -1.0;
  int32_t iters = -1, status = -1;
  printf("b200icp version %d\n", b200icp_version());
  const int rc = b200icp_run_host(&pool[0][0][0], counts, 2, N, 2, pairs, init, &eps, &normalize, 1, 200, 1e-14,
                                  T, &err, &iters, &status, NULL);
  if (rc != B200ICP_OK) { printf("b200icp_run_host: %s\n", b200icp_error_string(rc)); return 2; }
  const double got_th = atan2(T[4], T[0]);
  printf("passes %d status %d error %.3e  theta %.9f (want %.9f)  t (%.9f, %.9f) (want %.9f, %.9f)\n", iters,
         status, err, got_th, th, T[3], T[7], tx, ty);
  if (fabs(got_th - th) > 1e-6 || fabs(T[3] - tx) > 1e-6 || fabs(T[7] - ty) > 1e-6) return 1;
  if (T[15] != 1.0 || T[10] != 1.0 || iters < 1) return 1;
  b200icp_host_release();
  return 0;
}
