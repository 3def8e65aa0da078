// Host-visible declarations of the fused prox kernels (prox_elem.cu).
#pragma once
#include "internal.h"

namespace eps {

// One prox function of an elementwise / block stage after the VectorProx reduction
// (reference: VectorProx::InitScalar, src/epsilon/prox/vector_prox.cc:51-70):
//   w  = u + y_old            (prox_admm.cc:142  u_ += y_[i])
//   v  = Bs*w + g             (vector_prox.cc:131 PreProcessInput, B = (beta/gamma) H A')
//   x' = prox_{lam f}(v)      (ApplyVector)
//   x  = Cs*(x' - g)          (vector_prox.cc:135 PostProcessOutput, C = (1/beta) H')
//   y  = As*x                 (prox_admm.cc:144  y_[i] = A_*x_[i])
//   u  = w - y                (prox_admm.cc:145)
struct ElemSeg {
  int64_t u_off;   // first row of the constraint block in u (and of the span of y)
  int64_t y_off;   // offset of y_i in the y arena
  int64_t x_off;   // offset of x_i in the x arena
  int64_t g_off;   // offset of g in the constant arena, -1 when absent
  int64_t q_off;   // SUM_QUANTILE: alpha at q_off, beta at q_off + len; -1 otherwise
  // VectorProx::InitDiagonal (vector_prox.cc:72-116): per-element B, C, D, lambda at d_off, d_off + len,
  // d_off + 2 len, d_off + 3 len of the constant arena; x = C (x' - g) + D w.  -1: the scalar reduction.
  int64_t d_off;
  // has_axis (vector_prox.cc:150-177): a NORM_2 segment is one column (stride 1) or one row (stride = rows) of the
  // column-major matrix argument; element k sits at offset + k * stride
  int64_t stride;
  int32_t len;
  int32_t type;    // eps_prox_type
  double lam, Bs, Cs, As;
  double zone_m, zone_alpha, zone_beta;  // scaled_zone.cc:34-76
};

// TOTAL_VARIATION_1D: the pre / post maps of VectorProx handed to the solver so that it fuses them into its own
// first and last pass (tv1d_par.cu)
struct TvFuse {
  ElemSeg seg;
  const double* consts;
  double* u;
  const double* y_old;
  double* y_new;
  double* x;
};

struct SegPair {
  int64_t a_off, b_off;
  int32_t len;
  double coef;
};

// Fused elementwise step over the segments of one stage.
void ElemProxStep(Ctx* ctx, const ElemSeg* segs_dev, const int64_t* prefix_dev, int nseg,
                  int64_t total, int type, const double* consts, double* u, const double* y_old,
                  double* y_new, double* x);
// The same for a stage of ONE segment with even offsets: 128-bit accesses, no segment search (host copy of the
// segment passed by value).
void ElemProxStepSingle(Ctx* ctx, const ElemSeg& seg, const double* consts, double* u, const double* y_old,
                        double* y_new, double* x);
// NORM_2 block soft threshold, one warp (len <= 1024) or one CTA per segment.
void Norm2ProxStep(Ctx* ctx, const ElemSeg* segs_dev, int nseg, int max_len, const double* consts,
                   double* u, const double* y_old, double* y_new, double* x);
// TOTAL_VARIATION_1D for one segment (host copy of the segment passed by value).
void Tv1dProxStep(Ctx* ctx, const ElemSeg& seg, const double* consts, double* u, const double* y_old,
                  double* y_new, double* x, double* work /* Tv1dWorkDoubles(len) */,
                  unsigned* tag_state /* two host words kept with `work`: round tag of the slot maxima, rounds of the last call */);
size_t Tv1dWorkDoubles(int64_t len);

// dst[a_off + j] += c1*src1[b_off + j] + c2*src2[b_off + j]   (src2 may be null)
void SegAxpy(Ctx* ctx, const SegPair* pairs_dev, const int64_t* prefix_dev, int nseg, int64_t total,
             double* dst, const double* src1, double c1, const double* src2, double c2);
// out[0] += sum_seg coef_seg * sum_j v[a_off + j]^2 ; out[1] = max(out[1], max_seg sum_j w[b_off+j]^2)
// (either part may be skipped by passing null)
void SegSumSq(Ctx* ctx, const SegPair* pairs_dev, const SegPair* pairs_host, int nseg,
              const double* v, double* out_sum, const double* w, double* out_max);

// KKT prox with a simple scalar constraint block A = a*I:
//   pre : w = u + y_old (stored in u), rhs = a*w + c0
//   post: y_new = a*x, u = w - y_new
void KktScalarPre(Ctx* ctx, int64_t n, double a, const double* c0, double* u, const double* y_old,
                  double* rhs);
void KktScalarPost(Ctx* ctx, int64_t n, double a, const double* x, double* u, double* y_new);
// the three steps above with a DIAGONAL solve (x = dinv .* rhs) between them, fused
void KktScalarDiagStep(Ctx* ctx, int64_t n, double a, const double* c0, const double* dinv, double* u,
                       const double* y_old, double* y_new, double* x);
// Two-block consensus ADMM (prox_admm_two_block.cc:111-113,135-146), one pass:
//   z = mask .* (v - t)   (t = H'y of the projection; mask may be null)
//   u += x - z
//   sums[0..4] = ||x - z||^2, ||z - z_prev||^2, ||x||^2, ||z||^2, ||u||^2
void TwoBlockDualStep(Ctx* ctx, int64_t n, const double* x, const double* v, const double* t,
                      const double* mask, const double* z_prev, double* z, double* u, double* sums);
// ---- multi-argument / sort-based vector operators (prox_vec.cu) ---------------------------
// The operator acts on the stacked, already reduced argument vector V = [arg0 | arg1 | ...]
// (VectorProx::PreProcessInput, vector_prox.cc:131) and overwrites it with the result.
struct VecOpDesc {
  int type = 0;
  bool epigraph = false;
  int64_t n0 = 0;        // rows of the first argument
  int64_t rows = 0;      // all rows (n0 + 1 for the epigraph forms, 2*n0 for SUM_KL_DIV)
  double lam = 0;        // VectorProx lambda (alpha * beta^2 / gamma)
  double zone_m = 0, zone_alpha = 1, zone_beta = 1;
  int k = 0;             // SUM_LARGEST: sum_largest_params.k
  int64_t soc_n = 0;     // SECOND_ORDER_CONE: columns of X (n0 = m rows)
  double soc_a = 1;      // at / |ax|
  int mat_m = 0, mat_n = 0;  // matrix functions: mat(arg0, m, n), column-major (arg_size(0))
};
struct VecOpWork {
  DevBuf<double> keys, keys2, vals, vals2, scan1, scan2, scal, ev, ent, v0;
  // matrix functions (prox_mat.cu): S / M, eigenvectors, result, spectrum, operator output, weights
  DevBuf<double> mS, mQ, mX, md, mx, mw;
  VecOpDesc inner_desc;
  std::unique_ptr<VecOpWork> inner;
  DevBuf<int> idx;
  DevBuf<char> tmp;
};
bool IsVecOpType(int type, bool epigraph);
void VecOpInit(Ctx* ctx, const VecOpDesc& d, VecOpWork* w);
void VecOpApply(Ctx* ctx, const VecOpDesc& d, double* V, VecOpWork* w);
// orthogonally invariant matrix functions (prox_mat.cu), reached through VecOpInit / VecOpApply
bool IsMatOpType(int type);
void MatOpInit(Ctx* ctx, const VecOpDesc& d, VecOpWork* w);
void MatOpApply(Ctx* ctx, const VecOpDesc& d, double* V, VecOpWork* w);

// out[i + c n] = d[i] * in[i + c n] + b * y[i + c n]  for c < k   (y may be null)
void DiagMulti(Ctx* ctx, int64_t n, int k, const double* d, const double* in, double b, const double* y, double* out);
// x = dinv .* rhs
void KktDiagSolve(Ctx* ctx, int64_t n, const double* dinv, const double* rhs, double* x);

}  // namespace eps
