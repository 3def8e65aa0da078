// Parallel exact solver for the 1-D total-variation prox
//     argmin_b 0.5*||b - v||^2 + lam * sum_k |b[k+1] - b[k]|
// (reference: TotalVariation1DProx, src/epsilon/prox/total_variation_1d.cc:12-24 -> glmgen
// tf_dp.c:43-170, a sequential O(n) dynamic program).
//
// The dynamic program is sequential by construction, so the device path solves the same
// strictly convex problem through its optimality conditions instead.  With the dual variable
// z (one per boundary k between samples k and k+1),  b = v - D'z,  |z_k| <= lam,
// z_k = lam*sign(b[k+1]-b[k]) at jumps.  For a GIVEN set of jump boundaries with their signs
// ("active" boundaries) everything is closed-form from prefix sums R[k] = sum_{i<k} v_i:
//     group [s..e] between consecutive active boundaries:  b = (R[e+1]-R[s] - (z_left - z_right)) / (e-s+1)
//     free boundary k inside it:                            z_k = z_left + (k-s+1)*b - (R[k+1]-R[s])
// The active set is found by split-and-merge rounds, every round fully data parallel:
//     split: in each group, the boundary with the largest violation z_k - lam > 0 (and the one
//            with the largest -lam - z_k > 0) becomes active  (the taut-string picture: the chord
//            leaves the tube, its farthest point is pinned to the tube wall -- QuickHull-like);
//     merge: an active boundary whose jump has the wrong sign is released.
// The loop ends when a round changes nothing; the iterate then satisfies the KKT conditions of
// a strictly convex problem, i.e. it IS the minimiser the dynamic program computes (up to
// rounding: the prefix sums are carried in double-double, so group sums are exact to ~1e-30).
// Cold starts take O(log(#groups)) rounds (splits only), warm starts (the active set of the
// previous ADMM iteration is kept) 1-3.  If the loop has not settled after kMaxRounds the
// caller reports EPS_ERR_NUMERIC (tv1d.cu); nothing is approximated and there is no second algorithm.
#include <cub/cub.cuh>
#include <thrust/iterator/reverse_iterator.h>

#include "device.cuh"
#include "prox.cuh"

namespace eps {

namespace {

constexpr int kTvpBlock = 256;
constexpr int kMaxRounds = 400;

// Per-group maxima live in slots that are never cleared: a slot holds (round tag << 40 | top 39 bits of the
// positive double), so an entry of an earlier round loses against any entry of the current one.  (Clearing two
// slot arrays cost 16n bytes per round.)  Violations that agree in their top 39 bits tie and are both activated,
// which the algorithm allows: every boundary with z_k > lam may be pinned, the merge step releases wrong pins.
__device__ __forceinline__ unsigned long long PackSlot(unsigned tag, double v) {
  return ((unsigned long long)tag << 40) | ((unsigned long long)__double_as_longlong(v) >> 24);
}

struct DD { double hi, lo; };

struct DDAdd {
  __device__ __forceinline__ DD operator()(const DD& a, const DD& b) const {
    // Knuth two-sum of the high parts, low parts folded in, renormalised
    const double s = a.hi + b.hi;
    const double bb = s - a.hi;
    double e = (a.hi - (s - bb)) + (b.hi - bb);
    e += a.lo + b.lo;
    DD r;
    r.hi = s + e;
    r.lo = e - (r.hi - s);
    return r;
  }
};

struct ToDD {
  __device__ __forceinline__ DD operator()(const double& x) const { return DD{x, 0.0}; }
};

// R[b] - R[a] rounded to double
__device__ __forceinline__ double DDDiff(const DD* __restrict__ R, int b, int a) {
  const DD x = R[b], y = R[a];
  return (x.hi - y.hi) + (x.lo - y.lo);
}

struct ActiveIndex {  // k if boundary k is active else -1
  const int8_t* act;
  __device__ __forceinline__ int operator()(const int& k) const { return act[k] != 0 ? k : -1; }
};
struct ActiveIndexRev {  // reversed order: position j <-> boundary m-1-j; value m-1-k' mapped so that max == nearest
  const int8_t* act;
  int m;
  __device__ __forceinline__ int operator()(const int& j) const { return act[m - 1 - j] != 0 ? j : -1; }
};
struct MaxOp {
  __device__ __forceinline__ int operator()(const int& a, const int& b) const { return a > b ? a : b; }
};

struct Group {
  int s, e;        // first / last element
  double zl, zr;   // dual value on the boundary before s / after e (0 at the ends)
};

// Group that contains element i.  lastb[k]: last active boundary <= k (-1: none);
// nextr[j]: as lastb on the reversed boundary order (j = m-1-k).
__device__ __forceinline__ Group GroupOf(int i, int n, int m, const int8_t* __restrict__ act,
                                         const int* __restrict__ lastb, const int* __restrict__ nextr,
                                         double lam) {
  Group g;
  const int lb = i > 0 ? lastb[i - 1] : -1;                   // last active boundary < i
  int rb = m;                                                  // first active boundary >= i
  if (i < m) { const int j = nextr[m - 1 - i]; rb = j >= 0 ? m - 1 - j : m; }
  g.s = lb + 1;
  g.e = rb == m ? n - 1 : rb;
  g.zl = lb >= 0 ? lam * (double)act[lb] : 0.0;
  g.zr = rb < m ? lam * (double)act[rb] : 0.0;
  return g;
}

__device__ __forceinline__ double GroupValue(const Group& g, const DD* __restrict__ R) {
  return (DDDiff(R, g.e + 1, g.s) - (g.zl - g.zr)) / (double)(g.e - g.s + 1);
}

// Round, part 1: z on every free boundary, and per group the largest violation of each sign
// (warp-aggregated atomicMax on the slot of the group's first element; the bit pattern of a
// positive double orders like the value).
__global__ void __launch_bounds__(kTvpBlock)
TvpViolationKernel(int n, double lam, const DD* __restrict__ R, const int8_t* __restrict__ act,
                   const int* __restrict__ lastb, const int* __restrict__ nextr, double* __restrict__ z,
                   unsigned long long* __restrict__ slot_pos, unsigned long long* __restrict__ slot_neg, unsigned tag) {
  const int m = n - 1;
  const int k = blockIdx.x * kTvpBlock + threadIdx.x;
  const int lane = threadIdx.x & 31;
  int gid = -1 - k;  // unique when not a candidate
  double vp = 0.0, vn = 0.0;
  if (k < m && act[k] == 0) {
    const Group g = GroupOf(k, n, m, act, lastb, nextr, lam);
    const double b = GroupValue(g, R);
    const double zk = g.zl + (double)(k - g.s + 1) * b - DDDiff(R, k + 1, g.s);
    z[k] = zk;
    vp = zk - lam;
    vn = -lam - zk;
    gid = g.s;
  }
  vp = vp > 0.0 ? vp : 0.0;
  vn = vn > 0.0 ? vn : 0.0;
  // segmented max over runs of equal gid inside the warp
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const double op = __shfl_down_sync(kFull, vp, off);
    const double on = __shfl_down_sync(kFull, vn, off);
    const int og = __shfl_down_sync(kFull, gid, off);
    if (lane + off < 32 && og == gid) {
      vp = op > vp ? op : vp;
      vn = on > vn ? on : vn;
    }
  }
  const int pg = __shfl_up_sync(kFull, gid, 1);
  const bool head = lane == 0 || pg != gid;
  if (head && gid >= 0) {
    if (vp > 0.0) atomicMax(&slot_pos[gid], PackSlot(tag, vp));
    if (vn > 0.0) atomicMax(&slot_neg[gid], PackSlot(tag, vn));
  }
}

// Round, part 2: the new active set (double-buffered) and the number of changes.
__global__ void __launch_bounds__(kTvpBlock)
TvpUpdateKernel(int n, double lam, const DD* __restrict__ R, const int8_t* __restrict__ act,
                int8_t* __restrict__ act_out, const int* __restrict__ lastb, const int* __restrict__ nextr,
                const double* __restrict__ z, const unsigned long long* __restrict__ slot_pos,
                const unsigned long long* __restrict__ slot_neg, unsigned tag, int* __restrict__ changed) {
  const int m = n - 1;
  const int k = blockIdx.x * kTvpBlock + threadIdx.x;
  bool diff = false;
  if (k < m) {
    const int8_t a = act[k];
    int8_t out = a;
    if (a != 0) {
      // jump across an active boundary: value of the group right of k minus the one left of it
      const Group gl = GroupOf(k, n, m, act, lastb, nextr, lam);
      const Group gr = GroupOf(k + 1, n, m, act, lastb, nextr, lam);
      const double jump = GroupValue(gr, R) - GroupValue(gl, R);
      if (jump * (double)a <= 0.0) out = 0;
    } else {
      const int lb = lastb[k];  // k is free: last active boundary < k
      const int s = lb + 1;
      const double zk = z[k];
      const double vp = zk - lam, vn = -lam - zk;
      if (vp > 0.0 && PackSlot(tag, vp) == slot_pos[s]) out = 1;
      else if (vn > 0.0 && PackSlot(tag, vn) == slot_neg[s]) out = -1;
    }
    act_out[k] = out;
    diff = out != a;
  }
  const unsigned ball = __ballot_sync(kFull, diff);
  if ((threadIdx.x & 31) == 0 && ball) atomicAdd(changed, __popc(ball));
}

__global__ void __launch_bounds__(kTvpBlock)
TvpValuesKernel(int n, double lam, const DD* __restrict__ R, const int8_t* __restrict__ act,
                const int* __restrict__ lastb, const int* __restrict__ nextr, double* __restrict__ beta) {
  const int i = blockIdx.x * kTvpBlock + threadIdx.x;
  if (i >= n) return;
  const Group g = GroupOf(i, n, n - 1, act, lastb, nextr, lam);
  beta[i] = GroupValue(g, R);
}

__global__ void TvpSetZeroDD(DD* R) { R[0] = DD{0.0, 0.0}; }

size_t Align256(size_t b) { return (b + 255) / 256 * 256; }

}  // namespace

// Bytes of scratch the parallel solver needs for a signal of n samples (beyond v and beta).
size_t Tv1dParallelBytes(int64_t n) {
  size_t b = 0;
  b += Align256(sizeof(DD) * (n + 1));                 // R
  b += 2 * Align256(sizeof(int) * n);                  // lastb, nextr
  b += 2 * Align256(n);                                // act (double-buffered; [0] persists)
  b += Align256(sizeof(double) * n);                   // z
  b += 2 * Align256(sizeof(unsigned long long) * n);   // slot pair (never cleared: entries carry a round tag)
  b += 256;                                            // changed counter
  b += Align256((size_t)(n / 64 + 4096) * 16);         // cub temp storage (scans)
  return b;
}

// Returns the number of rounds used, or -1 when the loop did not settle (the caller then runs
// the sequential dynamic program).  `scratch` keeps the active set between calls (warm start);
// it must be zeroed to start cold.
int Tv1dParallelSolve(Ctx* ctx, int n, const double* v, double lam, double* beta, char* scratch,
                      int* h_changed /* pinned */, unsigned* tag_state /* host, per solver: last round tag used */) {
  const int m = n - 1;
  char* p = scratch;
  auto take = [&](size_t bytes) { char* q = p; p += Align256(bytes); return q; };
  DD* R = reinterpret_cast<DD*>(take(sizeof(DD) * (n + 1)));
  int* lastb = reinterpret_cast<int*>(take(sizeof(int) * n));
  int* nextr = reinterpret_cast<int*>(take(sizeof(int) * n));
  int8_t* act0 = reinterpret_cast<int8_t*>(take(n));
  int8_t* act1 = reinterpret_cast<int8_t*>(take(n));
  double* z = reinterpret_cast<double*>(take(sizeof(double) * n));
  unsigned long long* slots[2];
  for (int i = 0; i < 2; i++) slots[i] = reinterpret_cast<unsigned long long*>(take(sizeof(unsigned long long) * n));
  int* changed = reinterpret_cast<int*>(take(256));
  void* cub_tmp = p;
  size_t cub_bytes = Align256((size_t)(n / 64 + 4096) * 16);
  cudaStream_t st = ctx->stream;
  const int grid_n = (n + kTvpBlock - 1) / kTvpBlock;

  // R[k] = sum_{i<k} v_i in double-double
  {
    ProfScope ps(ctx, KC_TV, 24.0 * n);
    cub::TransformInputIterator<DD, ToDD, const double*> in(v, ToDD());
    size_t need = 0;
    CUDA_CHECK(cub::DeviceScan::InclusiveScan(nullptr, need, in, R + 1, DDAdd(), n, st));
    EPS_CHECK(need <= cub_bytes, "cub scan scratch %zu > %zu", need, cub_bytes);
    CUDA_CHECK(cub::DeviceScan::InclusiveScan(cub_tmp, need, in, R + 1, DDAdd(), n, st));
    TvpSetZeroDD<<<1, 1, 0, st>>>(R);
    ctx->count_launch(2);
    if (*tag_state + kMaxRounds + 2 >= (1u << 23)) {   // tag space used up: start over on cleared slots
      CUDA_CHECK(cudaMemsetAsync(slots[0], 0, sizeof(unsigned long long) * n, st));
      CUDA_CHECK(cudaMemsetAsync(slots[1], 0, sizeof(unsigned long long) * n, st));
      *tag_state = 0;
    }
  }
  int8_t* act = act0;
  int8_t* act_other = act1;
  int rounds = -1;
  cub::CountingInputIterator<int> idx(0);
  auto scans = [&](const int8_t* a) {
    cub::TransformInputIterator<int, ActiveIndex, cub::CountingInputIterator<int>> fwd(idx, ActiveIndex{a});
    cub::TransformInputIterator<int, ActiveIndexRev, cub::CountingInputIterator<int>> rev(idx, ActiveIndexRev{a, m});
    size_t need = 0;
    CUDA_CHECK(cub::DeviceScan::InclusiveScan(nullptr, need, fwd, lastb, MaxOp(), m, st));
    EPS_CHECK(need <= cub_bytes, "cub scan scratch %zu > %zu", need, cub_bytes);
    CUDA_CHECK(cub::DeviceScan::InclusiveScan(cub_tmp, need, fwd, lastb, MaxOp(), m, st));
    CUDA_CHECK(cub::DeviceScan::InclusiveScan(cub_tmp, need, rev, nextr, MaxOp(), m, st));
    ctx->count_launch(4);
  };
  if (m > 0) {
    for (int r = 0; r < kMaxRounds; r++) {
      ProfScope ps(ctx, KC_TV, 90.0 * n);
      unsigned long long* sp = slots[0], *sn = slots[1];
      const unsigned tag = ++(*tag_state);
      scans(act);
      CUDA_CHECK(cudaMemsetAsync(changed, 0, sizeof(int), st));
      TvpViolationKernel<<<grid_n, kTvpBlock, 0, st>>>(n, lam, R, act, lastb, nextr, z, sp, sn, tag);
      TvpUpdateKernel<<<grid_n, kTvpBlock, 0, st>>>(n, lam, R, act, act_other, lastb, nextr, z, sp, sn, tag, changed);
      ctx->count_launch(2);
      CUDA_CHECK(cudaMemcpyAsync(h_changed, changed, sizeof(int), cudaMemcpyDeviceToHost, st));
      CUDA_CHECK(cudaStreamSynchronize(st));
      if (*h_changed == 0) { rounds = r; break; }   // act is a fixed point: lastb/nextr match it
      std::swap(act, act_other);
    }
    if (rounds < 0) return -1;
  } else {
    rounds = 0;
  }
  {
    ProfScope ps(ctx, KC_TV, 40.0 * n);
    TvpValuesKernel<<<grid_n, kTvpBlock, 0, st>>>(n, lam, R, act, lastb, nextr, beta);
    ctx->count_launch();
    if (act != act0) CUDA_CHECK(cudaMemcpyAsync(act0, act, m, cudaMemcpyDeviceToDevice, st));
  }
  return rounds;
}

}  // namespace eps
