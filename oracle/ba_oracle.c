/*
 * CPU ORACLE -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Restates, in plain C, the reference's dense bundle adjustment (src/lib/droid_kernels.cu::ba_cuda
 * and src/geom/ba.py::BA) and DSPO stage 2 (src/geom/ba.py::BA_with_scale_shift).  See
 * ba_oracle_impl.h for the per-function citations.  Built by oracle/Makefile into
 * oracle/liboracle.so and called through ctypes by oracle/oracle.py.
 *
 * Parity pin: checked against outputs of the reference itself (tests/test_oracle_golden.py) --
 *   geom_py mode + DSPO : tests/golden/geom_*.npz, produced in the build container by importing
 *                         the reference's src/geom/ba.py on CPU (tests/golden/make_golden.py);
 *   droid_cuda mode and the geometry functions (frame_distance, depth_filter, iproj, projmap):
 *                         tests/golden/droid_*.npz and geomfn.npz, produced on a B200 by the reference's
 *                         own CUDA droid_backends rebuilt unmodified for sm_100 (oracle/build_ref.py,
 *                         tests/golden/make_golden_cuda.py).  Both sets are committed.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* Dense fp64 Cholesky solve (row-major, lower).  Stands in for Eigen::SimplicialLLT<double>
 * (droid_kernels.cu:1199-1203) and torch.linalg.cholesky (chol.py:11-12): same factorisation up to
 * the fill-reducing permutation, which does not change the solution.  Returns 0 when a pivot is
 * not positive (the reference then returns a zero update). Destroys A. */
static int oracle_chol_solve(double *A, const double *b, double *x, int n) {
  int i, j, k;
  for (j = 0; j < n; j++) {
    double d = A[(size_t)j * n + j];
    for (k = 0; k < j; k++) d -= A[(size_t)j * n + k] * A[(size_t)j * n + k];
    if (!(d > 0.0) || !isfinite(d)) return 0;
    d = sqrt(d);
    A[(size_t)j * n + j] = d;
#pragma omp parallel for schedule(static) if (n - j > 256)
    for (i = j + 1; i < n; i++) {
      double s = A[(size_t)i * n + j];
      int kk;
      for (kk = 0; kk < j; kk++) s -= A[(size_t)i * n + kk] * A[(size_t)j * n + kk];
      A[(size_t)i * n + j] = s / d;
    }
  }
  for (i = 0; i < n; i++) {
    double s = b[i];
    for (k = 0; k < i; k++) s -= A[(size_t)i * n + k] * x[k];
    x[i] = s / A[(size_t)i * n + i];
  }
  for (i = n - 1; i >= 0; i--) {
    double s = x[i];
    for (k = i + 1; k < n; k++) s -= A[(size_t)k * n + i] * x[k];
    x[i] = s / A[(size_t)i * n + i];
  }
  return 1;
}

#define REAL float
#define SUFFIX f32
#include "ba_oracle_impl.h"
#undef REAL
#undef SUFFIX

#define REAL double
#define SUFFIX f64
#include "ba_oracle_impl.h"
#undef REAL
#undef SUFFIX

/* torchrun exports OMP_NUM_THREADS=1; the CPU baseline must use the cores it reports (bench.py) */
void oracle_set_threads(int n) {
#ifdef _OPENMP
  extern void omp_set_num_threads(int);
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

int oracle_num_threads(void) {
#ifdef _OPENMP
  extern int omp_get_max_threads(void);
  return omp_get_max_threads();
#else
  return 1;
#endif
}
