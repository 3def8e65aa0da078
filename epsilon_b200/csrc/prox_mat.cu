// Orthogonally invariant matrix functions F(X) = f(spectrum(X)):
//   OrthoInvariantProx::ApplyVector   src/epsilon/prox/ortho_invariant.cc:18-73
//   LAMBDA_MAX   = MAX          on the eigenvalues of (Y + Y')/2      (lambda_max.cc)
//   NEG_LOG_DET  = SUM_NEG_LOG  on the eigenvalues of (Y + Y')/2      (neg_log_det.cc)
//   SEMIDEFINITE = NON_NEGATIVE on the eigenvalues of (Y + Y')/2, plus the skew part (semidefinite.cc)
//   NORM_NUCLEAR = NORM_1       on the singular values, from the eigenvalues of Y'Y + 1e-15 I (norm_nuclear.cc)
// The argument is mat(v, m, n), column-major.  The reference calls Eigen's SelfAdjointEigenSolver;
// here the decomposition is a cyclic Jacobi iteration run by one CTA (round-robin ordering:
// n/2 disjoint rotations per round, applied to columns, then to rows), in shared memory when the
// matrix and its eigenvectors fit (n <= 112), in global memory otherwise.  X = U f(d) V' does not
// depend on the order or the signs of the eigenvectors, so the result agrees with the reference to
// rounding (~1e-13).  The vector operator on the spectrum is the one of prox_vec.cu / the closed forms.
#include "device.cuh"
#include "prox.cuh"

namespace eps {

namespace {

constexpr int kBlock = 256;
constexpr int kJacobiThreads = 1024;

inline int Grid(int64_t n) { return (int)std::max<int64_t>(1, std::min<int64_t>((n + kBlock - 1) / kBlock, 148 * 8)); }

// S = (Y + Y')/2 (symmetric part, m == n) or S = Y'Y + 1e-15 I (n x n)
__global__ void MatPrepKernel(int sym, int m, int n, const double* __restrict__ Y, double* __restrict__ S) {
  const int64_t total = (int64_t)n * n;
  for (int64_t e = (int64_t)blockIdx.x * kBlock + threadIdx.x; e < total; e += (int64_t)gridDim.x * kBlock) {
    const int i = (int)(e % n), j = (int)(e / n);
    double s;
    if (sym) {
      s = (Y[i + (int64_t)j * m] + Y[j + (int64_t)i * m]) / 2;
    } else {
      s = 0.0;
      const double* a = Y + (int64_t)i * m;
      const double* b = Y + (int64_t)j * m;
      for (int k = 0; k < m; k++) s += a[k] * b[k];
      if (i == j) s += 1e-15;
    }
    S[e] = s;
  }
}

// Cyclic Jacobi: on exit d[i] = eigenvalue i, Q(:, i) its eigenvector (A = Q diag(d) Q').
// A and Q are n x n column-major in global memory; the CTA works on shared copies when they fit.
__global__ void __launch_bounds__(kJacobiThreads)
JacobiEigKernel(int n, double* __restrict__ Ag, double* __restrict__ Qg, double* __restrict__ d, int use_smem,
                int* __restrict__ sweeps_out) {
  extern __shared__ double dyn[];
  __shared__ double red[32];
  __shared__ double s_norm;
  __shared__ int s_rot;
  const int tid = threadIdx.x, nt = blockDim.x;
  const int nn = n * n;
  const int ne = n + (n & 1), half = ne / 2;
  // rotation tables live behind the matrices (smem) or at the start of dyn
  double* A = use_smem ? dyn : Ag;
  double* Q = use_smem ? dyn + nn : Qg;
  double* cs = use_smem ? dyn + 2 * nn : dyn;   // c[half], s[half]
  int* pq = (int*)(cs + 2 * half);              // p[half], q[half]
  for (int e = tid; e < nn; e += nt) {
    if (use_smem) A[e] = Ag[e];
    Q[e] = (e % n == e / n) ? 1.0 : 0.0;
  }
  __syncthreads();
  // Frobenius norm (invariant under the rotations)
  double loc = 0.0;
  for (int e = tid; e < nn; e += nt) loc += A[e] * A[e];
  loc = WarpSum(loc);
  if ((tid & 31) == 0) red[tid >> 5] = loc;
  __syncthreads();
  if (tid == 0) {
    double t = 0.0;
    for (int w = 0; w < (nt + 31) / 32; w++) t += red[w];
    s_norm = sqrt(t);
  }
  __syncthreads();
  const double tol = 2.2e-16 * s_norm;
  int sweep = 0;
  for (; sweep < 60; sweep++) {
    if (tid == 0) s_rot = 0;
    __syncthreads();
    for (int r = 0; r < ne - 1; r++) {
      if (tid < half) {
        int a, b;
        if (tid == 0) { a = ne - 1; b = r; }
        else { a = (r + tid) % (ne - 1); b = (r - tid + ne - 1) % (ne - 1); }
        const int p = min(a, b), q = max(a, b);
        double c = 1.0, s = 0.0;
        if (q < n) {
          const double apq = A[p + q * n];
          if (fabs(apq) > tol) {
            const double tau = (A[q + q * n] - A[p + p * n]) / (2 * apq);
            const double t = (tau >= 0 ? 1.0 : -1.0) / (fabs(tau) + sqrt(1 + tau * tau));
            c = 1 / sqrt(1 + t * t);
            s = t * c;
            s_rot = 1;
          }
        }
        cs[tid] = c; cs[half + tid] = s;
        pq[tid] = p; pq[half + tid] = q;
      }
      __syncthreads();
      // columns: A <- A J, Q <- Q J
      for (int e = tid; e < half * n; e += nt) {
        const int k = e / n, i = e % n;
        const double s = cs[half + k];
        if (s == 0.0) continue;
        const double c = cs[k];
        const int p = pq[k], q = pq[half + k];
        double ap = A[i + p * n], aq = A[i + q * n];
        A[i + p * n] = c * ap - s * aq;
        A[i + q * n] = s * ap + c * aq;
        ap = Q[i + p * n]; aq = Q[i + q * n];
        Q[i + p * n] = c * ap - s * aq;
        Q[i + q * n] = s * ap + c * aq;
      }
      __syncthreads();
      // rows: A <- J' A
      for (int e = tid; e < half * n; e += nt) {
        const int k = e / n, j = e % n;
        const double s = cs[half + k];
        if (s == 0.0) continue;
        const double c = cs[k];
        const int p = pq[k], q = pq[half + k];
        const double ap = A[p + j * n], aq = A[q + j * n];
        A[p + j * n] = c * ap - s * aq;
        A[q + j * n] = s * ap + c * aq;
      }
      __syncthreads();
      if (tid < half && cs[half + tid] != 0.0) {
        const int p = pq[tid], q = pq[half + tid];
        A[p + q * n] = 0.0;
        A[q + p * n] = 0.0;
      }
      __syncthreads();
    }
    if (s_rot == 0) break;
    __syncthreads();
  }
  for (int i = tid; i < n; i += nt) d[i] = A[i + i * n];
  if (use_smem)
    for (int e = tid; e < nn; e += nt) Qg[e] = Q[e];
  if (tid == 0 && sweeps_out) *sweeps_out = sweep;
}

// spectrum -> argument of the vector operator: singular values d = sqrt(max(d, 0)) (ortho_invariant.cc:44)
__global__ void SqrtSpectrumKernel(int n, double* __restrict__ d) {
  for (int i = blockIdx.x * kBlock + threadIdx.x; i < n; i += gridDim.x * kBlock) d[i] = sqrt(fmax(d[i], 0.0));
}

// closed-form vector operators on the spectrum: x = prox_{lam f}(d)
__global__ void SpectrumProxKernel(int type, int n, double lam, const double* __restrict__ d, double* __restrict__ x) {
  for (int i = blockIdx.x * kBlock + threadIdx.x; i < n; i += gridDim.x * kBlock) {
    const double v = d[i];
    double r;
    if (type == EPS_PROX_NON_NEGATIVE) r = fmax(v, 0.0);                                  // non_negative.cc:5-9
    else if (type == EPS_PROX_NORM_1) r = v > lam ? v - lam : (v < -lam ? v + lam : 0.0); // scaled_zone.cc:96-110, M = 0
    else {                                                                                // ApplyNegLogProx, sum_neg_log.cc:10-26
      const double z = sqrt(v * v + 4 * lam);
      r = v >= 0 ? (v + z) / 2 : 2 * lam / (-v + z);
    }
    x[i] = r;
  }
}

// w_i = x_i (U = V) or x_i / d_i with 1/0 = 0 (U = Y V diag(1/d), ortho_invariant.cc:45-52)
__global__ void SpectrumWeightKernel(int sym, int n, const double* __restrict__ d, const double* __restrict__ x,
                                     double* __restrict__ w) {
  for (int i = blockIdx.x * kBlock + threadIdx.x; i < n; i += gridDim.x * kBlock)
    w[i] = sym ? x[i] : (d[i] != 0.0 ? x[i] / d[i] : 0.0);
}

// M = Q diag(w) Q'  (n x n)
__global__ void QwQtKernel(int n, const double* __restrict__ Q, const double* __restrict__ w, double* __restrict__ M) {
  const int64_t total = (int64_t)n * n;
  for (int64_t e = (int64_t)blockIdx.x * kBlock + threadIdx.x; e < total; e += (int64_t)gridDim.x * kBlock) {
    const int i = (int)(e % n), j = (int)(e / n);
    double s = 0.0;
    for (int k = 0; k < n; k++) s += Q[i + (int64_t)k * n] * w[k] * Q[j + (int64_t)k * n];
    M[e] = s;
  }
}

// mode 0: X = Y M (m x n); 1: X = M; 2: X = (M + M')/2; 3: X = M + (Y - Y')/2
__global__ void MatFinishKernel(int mode, int m, int n, const double* __restrict__ Y, const double* __restrict__ M,
                                double* __restrict__ X) {
  const int64_t total = (int64_t)m * n;
  for (int64_t e = (int64_t)blockIdx.x * kBlock + threadIdx.x; e < total; e += (int64_t)gridDim.x * kBlock) {
    const int i = (int)(e % m), j = (int)(e / m);
    double s;
    if (mode == 0) {
      s = 0.0;
      for (int k = 0; k < n; k++) s += Y[i + (int64_t)k * m] * M[k + (int64_t)j * n];
    } else if (mode == 1) {
      s = M[e];
    } else if (mode == 2) {
      s = (M[i + (int64_t)j * n] + M[j + (int64_t)i * n]) / 2;
    } else {
      s = M[e] + (Y[i + (int64_t)j * m] - Y[j + (int64_t)i * m]) / 2;
    }
    X[e] = s;
  }
}

}  // namespace

bool IsMatOpType(int type) {
  return type == EPS_PROX_LAMBDA_MAX || type == EPS_PROX_NEG_LOG_DET || type == EPS_PROX_NORM_NUCLEAR ||
         type == EPS_PROX_SEMIDEFINITE;
}

static int SpectrumType(int type) {
  switch (type) {
    case EPS_PROX_LAMBDA_MAX: return EPS_PROX_MAX;
    case EPS_PROX_NEG_LOG_DET: return EPS_PROX_SUM_NEG_LOG;
    case EPS_PROX_NORM_NUCLEAR: return EPS_PROX_NORM_1;
    default: return EPS_PROX_NON_NEGATIVE;
  }
}

static size_t JacobiSmemBytes(int n, bool* use_smem) {
  const int half = (n + (n & 1)) / 2;
  const size_t tables = (size_t)2 * half * sizeof(double) + (size_t)2 * half * sizeof(int);
  const size_t full = (size_t)2 * n * n * sizeof(double) + tables;
  *use_smem = full <= 200 * 1024;
  return *use_smem ? full : tables;
}

void MatOpInit(Ctx* ctx, const VecOpDesc& d, VecOpWork* w) {
  const int m = d.mat_m, n = d.mat_n;
  EPS_CHECK(m > 0 && n > 0 && (int64_t)m * n == d.n0, "matrix function: arg_size %d x %d for %lld entries", m, n,
            (long long)d.n0);
  const bool sym = d.type != EPS_PROX_NORM_NUCLEAR;
  if (sym) EPS_CHECK(m == n, "matrix function of type %d needs a square argument, got %d x %d", d.type, m, n);
  // for m < n the reference sizes its nested operator min(m, n) but feeds it the n eigenvalues of Y'Y
  // (ortho_invariant.cc:77 against :41-46) and returns garbage; refuse instead of imitating it
  EPS_CHECK(m >= n, "NORM_NUCLEAR over a %d x %d argument: the reference operator is only defined for m >= n", m, n);
  EPS_CHECK(n <= 2048, "matrix function over %d x %d: the Jacobi eigen-solver runs in one CTA (n <= 2048)", n, n);
  EPS_CHECK(!d.epigraph || d.type != EPS_PROX_SEMIDEFINITE, "SEMIDEFINITE has no epigraph form");
  w->mS.alloc(ctx, (size_t)n * n);
  w->mQ.alloc(ctx, (size_t)n * n);
  w->mX.alloc(ctx, (size_t)m * n);
  w->md.alloc(ctx, (size_t)n + 1);
  w->mx.alloc(ctx, (size_t)n + 1);
  w->mw.alloc(ctx, (size_t)n);
  bool use_smem;
  const size_t smem = JacobiSmemBytes(n, &use_smem);
  CUDA_CHECK(cudaFuncSetAttribute(JacobiEigKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int st = SpectrumType(d.type);
  if (IsVecOpType(st, d.epigraph)) {
    w->inner_desc = VecOpDesc();
    w->inner_desc.type = st;
    w->inner_desc.epigraph = d.epigraph;
    w->inner_desc.n0 = n;
    w->inner_desc.rows = n + (d.epigraph ? 1 : 0);
    w->inner_desc.lam = d.lam;
    w->inner = std::make_unique<VecOpWork>();
    VecOpInit(ctx, w->inner_desc, w->inner.get());
  }
}

void MatOpApply(Ctx* ctx, const VecOpDesc& d, double* V, VecOpWork* w) {
  const int m = d.mat_m, n = d.mat_n;
  const bool sym = d.type != EPS_PROX_NORM_NUCLEAR;
  cudaStream_t st = ctx->stream;
  MatPrepKernel<<<Grid((int64_t)n * n), kBlock, 0, st>>>(sym ? 1 : 0, m, n, V, w->mS.p);
  bool use_smem;
  const size_t smem = JacobiSmemBytes(n, &use_smem);
  JacobiEigKernel<<<1, kJacobiThreads, smem, st>>>(n, w->mS.p, w->mQ.p, w->md.p, use_smem ? 1 : 0, w->idx.p + 1);
  ctx->count_launch(2);
  if (!sym) {
    SqrtSpectrumKernel<<<Grid(n), kBlock, 0, st>>>(n, w->md.p);
    ctx->count_launch();
  }
  const int stype = SpectrumType(d.type);
  if (w->inner) {
    // x <- d (and s), then the vector operator in place
    VecCopy(ctx, n, w->md.p, w->mx.p);
    if (d.epigraph) VecCopy(ctx, 1, V + (int64_t)m * n, w->mx.p + n);
    VecOpApply(ctx, w->inner_desc, w->mx.p, w->inner.get());
    if (d.epigraph) VecCopy(ctx, 1, w->mx.p + n, V + (int64_t)m * n);
  } else {
    EPS_CHECK(!d.epigraph, "matrix function %d: epigraph form without a spectrum operator", d.type);
    SpectrumProxKernel<<<Grid(n), kBlock, 0, st>>>(stype, n, d.lam, w->md.p, w->mx.p);
    ctx->count_launch();
  }
  SpectrumWeightKernel<<<Grid(n), kBlock, 0, st>>>(sym ? 1 : 0, n, w->md.p, w->mx.p, w->mw.p);
  QwQtKernel<<<Grid((int64_t)n * n), kBlock, 0, st>>>(n, w->mQ.p, w->mw.p, w->mS.p);
  int mode = 0;
  if (sym) mode = d.epigraph ? 1 : (d.type == EPS_PROX_SEMIDEFINITE ? 3 : 2);
  MatFinishKernel<<<Grid((int64_t)m * n), kBlock, 0, st>>>(mode, m, n, V, w->mS.p, w->mX.p);
  ctx->count_launch(3);
  VecCopy(ctx, (int64_t)m * n, w->mX.p, V);
}

}  // namespace eps
