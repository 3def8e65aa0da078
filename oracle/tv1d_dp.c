/* CPU restatement of the 1-D total-variation prox used by the reference.
 *
 * TEST INFRASTRUCTURE (see oracle/prox_admm_np.py).  Not linked into the product.
 *
 * Reference: TotalVariation1DProx::ApplyVector (src/epsilon/prox/total_variation_1d.cc:12-24)
 * calls glmgen's tf_dp (third_party/glmgen/c_lib/glmgen/src/tf/tf_dp.c:43-170), N. Johnson's
 * O(n) dynamic program ("A dynamic programming algorithm for the fused lasso and L0
 * segmentation", 2013) for
 *        minimise_b  1/2 sum_i (b_i - y_i)^2 + lam sum_i |b_{i+1} - b_i| .
 *
 * The derivative of the k-th message  f_k(b) = min over b_0..b_{k-1}  is a continuous
 * piecewise-linear, increasing function clipped to [-lam, lam]; it is held as a deque of
 * knots, each carrying the slope/intercept increment (da, db) that applies to its right
 * (front half) or to its left (back half).  Per sample the clip points lo_k / hi_k (where
 * the derivative crosses -lam / +lam) are found by walking in from either end, the walked
 * knots are dropped, and two new end knots are pushed.  The back pass clamps b_k to
 * [lo_k, hi_k] of b_{k+1}.
 *
 * The arithmetic (operation order of every accumulation and division) follows the
 * reference statement by statement so that results are bit-identical on the same input;
 * tests/test_oracle_golden.py pins this against outputs of the compiled reference.
 *
 * Build: gcc -O2 -fPIC -shared -ffp-contract=off -o oracle/_build/liboracle.so oracle/tv1d_dp.c
 */
#include <stdlib.h>

typedef struct {
  double pos;    /* knot position                      (reference: x[])  */
  double da;     /* slope change across the knot       (reference: a[])  */
  double db;     /* intercept change across the knot   (reference: b[])  */
} knot_t;

int oracle_tv1d(int n, const double *y, double lam, double *out) {
  if (n <= 0) return 0;
  if (n == 1 || lam == 0.0) {                       /* tf_dp.c:66-71 */
    for (int i = 0; i < n; i++) out[i] = y[i];
    return 0;
  }
  /* deque grows by at most one knot on each side per sample: 2n slots, start in the middle */
  knot_t *dq = (knot_t *)malloc(sizeof(knot_t) * 2 * (size_t)n);
  double *clip_lo = (double *)malloc(sizeof(double) * (size_t)(n - 1));
  double *clip_hi = (double *)malloc(sizeof(double) * (size_t)(n - 1));
  if (!dq || !clip_lo || !clip_hi) { free(dq); free(clip_lo); free(clip_hi); return -1; }

  /* first sample: f_0'(b) = b - y_0 clipped to [-lam, lam]            (tf_dp.c:82-95) */
  int head = n - 1, tail = n;
  clip_lo[0] = -lam + y[0];
  clip_hi[0] = lam + y[0];
  dq[head].pos = clip_lo[0]; dq[head].da = 1.0;  dq[head].db = -y[0] + lam;
  dq[tail].pos = clip_hi[0]; dq[tail].da = -1.0; dq[tail].db = y[0] + lam;
  /* line left of the first knot / right of the last knot after adding (b - y_k) */
  double a_left = 1.0, b_left = -y[1] - lam;
  double a_right = -1.0, b_right = y[1] - lam;

  for (int k = 1; k < n - 1; k++) {                                  /* tf_dp.c:98-141 */
    /* walk up from the head until the derivative exceeds -lam */
    double a_lo = a_left, b_lo = b_left;
    int lo = head;
    for (; lo <= tail; lo++) {
      if (a_lo * dq[lo].pos + b_lo > -lam) break;
      a_lo += dq[lo].da;
      b_lo += dq[lo].db;
    }
    /* walk down from the tail until the derivative drops below +lam */
    double a_hi = a_right, b_hi = b_right;
    int hi = tail;
    for (; hi >= lo; hi--) {
      if (-a_hi * dq[hi].pos - b_hi < lam) break;
      a_hi += dq[hi].da;
      b_hi += dq[hi].db;
    }
    clip_lo[k] = (-lam - b_lo) / a_lo;
    head = lo - 1;
    dq[head].pos = clip_lo[k];
    clip_hi[k] = (lam + b_hi) / (-a_hi);
    tail = hi + 1;
    dq[tail].pos = clip_hi[k];
    dq[head].da = a_lo; dq[head].db = b_lo + lam;
    dq[tail].da = a_hi; dq[tail].db = b_hi + lam;
    a_left = 1.0;  b_left = -y[k + 1] - lam;
    a_right = -1.0; b_right = y[k + 1] - lam;
  }

  /* last sample: zero of the derivative                              (tf_dp.c:145-153) */
  {
    double a_lo = a_left, b_lo = b_left;
    for (int lo = head; lo <= tail; lo++) {
      if (a_lo * dq[lo].pos + b_lo > 0.0) break;
      a_lo += dq[lo].da;
      b_lo += dq[lo].db;
    }
    out[n - 1] = -b_lo / a_lo;
  }
  /* back pass                                                         (tf_dp.c:157-162) */
  for (int k = n - 2; k >= 0; k--) {
    const double nxt = out[k + 1];
    out[k] = nxt > clip_hi[k] ? clip_hi[k] : (nxt < clip_lo[k] ? clip_lo[k] : nxt);
  }
  free(dq); free(clip_lo); free(clip_hi);
  return 0;
}
