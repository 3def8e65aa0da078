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
//
// Layout of a round (r02b).  The boundaries are cut into tiles of kTile; a CTA works on one tile at a time.
//   * "last active boundary <= k" / "first active boundary >= k" are never stored per boundary: a CTA rebuilds
//     them for its tile in shared memory from the 1-byte active flags and two per-tile carries (TileScans).
//     The carries come from per-tile aggregates (first / last active boundary of a tile), scanned by one CTA
//     (TvpCarryKernel) -- a few thousand entries.  (Before: two device-wide scans per round, 22n bytes.)
//   * only DIRTY tiles are visited after the first round of a call: a group whose extent did not change in the
//     previous round has the same duals and the same jumps as before, so it cannot produce a change; a change at
//     boundary k marks the tiles of the groups it touches (up to the two active boundaries that delimit them).
//     TvpCarryKernel compacts the dirty tiles into a work list that persistent grids walk.
//   * the active set is updated in place: decisions need only z (written by the violation pass, which also does
//     the jump test of the active boundaries) and the group maxima.
//   * no host round trip per round: rounds are queued in batches; every kernel returns at once when a previous
//     round settled (ctl[1] != 0), and the host reads one int per batch.  The values pass is queued behind the
//     batch before the host has that int (it returns at once if the rounds did not settle), so the host's round
//     trip overlaps with it.
//   * the double-double prefix sums are three plain passes of our own (tile sums, one CTA scanning them, tiles
//     scanned with their offset) instead of cub::DeviceScan on a 16-byte type with a user operator (337 us at
//     n = 1e7, 0.7 TB/s).
//   * group values are computed once per group segment of a tile (TileGroupValues) and z is stored only for tiles
//     in which something may change; those tiles form a second list, which is all the update pass walks.
//   * the VectorProx maps around the operator are fused in (tv1d.cu hands them over as TvFuse): the pre-map and
//     the dual update run in the tile-sum pass, the post-map in the values pass; v and beta never reach memory.
//   * the violation and values passes are software-pipelined over the tiles of a CTA (flags and carries of the next
//     tile, and the coalesced stream of the current one, are requested before the scans).
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "device.cuh"
#include "prox.cuh"

namespace eps {

namespace {

constexpr int kTvpBlock = 256;
constexpr int kItems = 8;
constexpr int kTile = kTvpBlock * kItems;   // boundaries (and samples) per tile
constexpr int kMaxRounds = 400;
constexpr int kCarryBlock = 1024;

// Per-group maxima live in slots that are never cleared: a slot holds (round tag << 40 | top 39 bits of the
// positive double), so an entry of an earlier round loses against any entry of the current one.  (Clearing two
// slot arrays cost 16n bytes per round.)  Violations that agree in their top 39 bits tie and are both activated,
// which the algorithm allows: every boundary with z_k > lam may be pinned, the merge step releases wrong pins.
__device__ __forceinline__ unsigned long long PackSlot(unsigned tag, double v) {
  return ((unsigned long long)tag << 40) | ((unsigned long long)__double_as_longlong(v) >> 24);
}

struct __align__(16) DD { double hi, lo; };   // one 128-bit access

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

// R[b] - R[a] rounded to double
__device__ __forceinline__ double DDDiff(const DD* __restrict__ R, int b, int a) {
  const DD x = R[b], y = R[a];
  return (x.hi - y.hi) + (x.lo - y.lo);
}


// Everything a round needs that is not per boundary.
struct TvState {
  int8_t* act;        // [ntiles*kTile] +1 / -1 / 0 per boundary; entries >= m stay 0.  Persists between calls.
  int* tlast1;        // [ntiles] 1 + last active boundary of the tile, 0: none.            Persists.
  int* tfirst1;       // [ntiles] m - first active boundary of the tile (> 0), 0: none.     Persists.
  int* carry_l;       // [ntiles] last active boundary before the tile, -1: none
  int* carry_r;       // [ntiles] first active boundary after the tile, m: none
  uint8_t* dirty[2];  // [ntiles] tiles to visit, by round parity
  int* list;          // [ntiles] the dirty tiles of the running round, compacted
  int* clist;         // [ntiles] tiles in which the violation pass found a boundary that may change (z stored there)
  int* ctl;           // [0] changes of the running round, [1] != 0: settled (number of rounds run), [2] tiles on the list,
                      // [3] tiles on the candidate list
  double* z;          // free boundary: its dual; active boundary: jump * sign (<= 0: release)
  unsigned long long* slot_pos;
  unsigned long long* slot_neg;
  const DD* R;
};

// padded index into the shared lb / rb arrays: the blocked writes of TileScans (stride kItems) and the striped
// reads of the round kernels are both free of bank conflicts
__device__ __forceinline__ int Pad(int i) { return i + (i >> 5); }
constexpr int kPadded = kTile + kTile / 32;

struct TileShared {
  int lb[kPadded];     // last active boundary <= k  (-1: none)
  int rb[kPadded];     // first active boundary >= k (m: none)
  int8_t a[kTile];
  int wf[kTvpBlock / 32], wr[kTvpBlock / 32];
};

// lb / rb / a of the tile that starts at boundary t0, given the carries of the tile.  All kTvpBlock threads.
// The 8 flags this thread owns in tile T (blocked order)
__device__ __forceinline__ unsigned long long TileFlags(const int8_t* __restrict__ act, int T) {
  static_assert(kItems == 8, "one 8-byte load per thread");
  return *reinterpret_cast<const unsigned long long*>(act + (size_t)T * kTile + threadIdx.x * kItems);
}

__device__ __forceinline__ void TileScans(TileShared& sh, int t0, int m, unsigned long long flags, int cl, int cr) {
  const int t = threadIdx.x, lane = t & 31, w = t >> 5;
  const int base = t * kItems;
  union { unsigned long long q; int8_t b[kItems]; } ld;
  ld.q = flags;
  int f[kItems], r[kItems];
  int run = -1;
#pragma unroll
  for (int j = 0; j < kItems; j++) {
    if (ld.b[j] != 0) run = t0 + base + j;
    f[j] = run;
  }
  int xf = run;
  run = m;
#pragma unroll
  for (int j = kItems - 1; j >= 0; j--) {
    if (ld.b[j] != 0) run = t0 + base + j;
    r[j] = run;
  }
  int xr = run;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int yf = __shfl_up_sync(kFull, xf, off);
    const int yr = __shfl_down_sync(kFull, xr, off);
    if (lane >= off) xf = max(xf, yf);
    if (lane + off < 32) xr = min(xr, yr);
  }
  if (lane == 31) sh.wf[w] = xf;
  if (lane == 0) sh.wr[w] = xr;
  __syncthreads();
  int pf = cl, pr = cr;
#pragma unroll
  for (int i = 0; i < kTvpBlock / 32; i++) {
    if (i < w) pf = max(pf, sh.wf[i]);
    if (i > w) pr = min(pr, sh.wr[i]);
  }
  int ef = __shfl_up_sync(kFull, xf, 1);
  int er = __shfl_down_sync(kFull, xr, 1);
  ef = lane == 0 ? pf : max(ef, pf);
  er = lane == 31 ? pr : min(er, pr);
#pragma unroll
  for (int j = 0; j < kItems; j++) {
    sh.lb[Pad(base + j)] = max(f[j], ef);
    sh.rb[Pad(base + j)] = min(r[j], er);
    sh.a[base + j] = ld.b[j];
  }
  __syncthreads();
}

// One CTA per round: decides whether the previous round settled, compacts the dirty tiles into the work list of
// this round, and scans the tile aggregates into the carries.
__global__ void __launch_bounds__(kCarryBlock)
TvpCarryKernel(TvState S, int ntiles, int m, int round, int first) {
  __shared__ int wl[kCarryBlock / 32], wr[kCarryBlock / 32], wc[kCarryBlock / 32];
  __shared__ int stop;
  const int t = threadIdx.x, lane = t & 31, w = t >> 5;
  if (t == 0) {
    if (first) S.ctl[1] = 0;
    stop = 0;
    if (S.ctl[1] != 0) stop = 1;
    else if (!first && S.ctl[0] == 0) { S.ctl[1] = round; stop = 1; }   // `round` rounds ran, the last changed nothing
    S.ctl[0] = 0;
    S.ctl[3] = 0;
    if (stop) S.ctl[2] = 0;
  }
  __syncthreads();
  if (stop) return;
  const uint8_t* cur = S.dirty[round & 1];
  uint8_t* nxt = S.dirty[(round + 1) & 1];
  const int per = (ntiles + kCarryBlock - 1) / kCarryBlock;
  const int lo = min(t * per, ntiles), hi = min(lo + per, ntiles);
  int al = -1, ar = m, ac = 0;
  for (int T = lo; T < hi; T++) {
    const int l1 = S.tlast1[T], f1 = S.tfirst1[T];
    if (l1 > 0) al = l1 - 1;               // tiles ascend: the last one seen is the largest
    if (f1 > 0 && ar == m) ar = m - f1;    // the first one seen is the smallest
    ac += first ? 1 : (cur[T] != 0);
    nxt[T] = 0;
  }
  // inclusive scans over the threads: prefix max (al), suffix min (ar), prefix sum (ac)
  int xl = al, xr = ar, xc = ac;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int yl = __shfl_up_sync(kFull, xl, off);
    const int yr = __shfl_down_sync(kFull, xr, off);
    const int yc = __shfl_up_sync(kFull, xc, off);
    if (lane >= off) { xl = max(xl, yl); xc += yc; }
    if (lane + off < 32) xr = min(xr, yr);
  }
  if (lane == 31) { wl[w] = xl; wc[w] = xc; }
  if (lane == 0) wr[w] = xr;
  __syncthreads();
  int pl = -1, pr = m, pc = 0;
  for (int i = 0; i < kCarryBlock / 32; i++) {
    if (i < w) { pl = max(pl, wl[i]); pc += wc[i]; }
    if (i > w) pr = min(pr, wr[i]);
  }
  int cl = __shfl_up_sync(kFull, xl, 1);
  int cr = __shfl_down_sync(kFull, xr, 1);
  cl = lane == 0 ? pl : max(cl, pl);
  cr = lane == 31 ? pr : min(cr, pr);
  int pos = pc + xc - ac;
  for (int T = lo; T < hi; T++) {
    S.carry_l[T] = cl;
    const int l1 = S.tlast1[T];
    if (l1 > 0) cl = l1 - 1;
    if (first || cur[T] != 0) S.list[pos++] = T;
  }
  for (int T = hi - 1; T >= lo; T--) {
    S.carry_r[T] = cr;
    const int f1 = S.tfirst1[T];
    if (f1 > 0) cr = m - f1;
  }
  if (t == kCarryBlock - 1) S.ctl[2] = pc + xc;   // tiles on the list
}

// Closes a batch: the last queued round is judged here instead of by the next TvpCarryKernel.
__global__ void TvpCheckKernel(TvState S, int rounds_run) {
  if (S.ctl[1] == 0 && S.ctl[0] == 0) S.ctl[1] = rounds_run;
}

struct Group {
  int s, e;        // first / last element
  double zl, zr;   // dual value on the boundary before s / after e (0 at the ends)
};

__device__ __forceinline__ double GroupValue(const Group& g, const DD* __restrict__ R) {
  return (DDDiff(R, g.e + 1, g.s) - (g.zl - g.zr)) / (double)(g.e - g.s + 1);
}

// Value of every group that has an element in the tile, computed ONCE per group by the first of its elements in the
// tile (an element is a "head" when it opens the tile or follows an active boundary) and left in bval[] at the
// head's index.  Warps without a head skip the division and the two R loads altogether -- with groups of hundreds
// of samples that is most warps.  Element i of the tile finds its group's value at HeadSlot(lb) where lb is the
// last active boundary before it.
__device__ __forceinline__ int HeadSlot(int lb, int t0) { return max(lb + 1 - t0, 0); }

template <typename SignFn>
__device__ __forceinline__ void TileGroupValues(const TileShared& sh, double* __restrict__ bval, int t0, int n, int m,
                                                int cl, double lam, const DD* __restrict__ R, SignFn sign,
                                                int8_t* __restrict__ zlsign = nullptr /* sign of the dual left of the group */) {
#pragma unroll 2
  for (int jj = 0; jj < kItems; jj++) {
    const int i = jj * kTvpBlock + threadIdx.x;
    const int el = t0 + i;
    if (el >= n) continue;
    if (i != 0 && sh.a[i - 1] == 0) continue;
    const int lb = i == 0 ? cl : el - 1;               // last active boundary before the element
    const int rb = el < m ? sh.rb[Pad(i)] : m;         // first active boundary at or after it
    Group g;
    g.s = lb + 1;
    g.e = rb == m ? n - 1 : rb;
    g.zl = lb >= 0 ? lam * sign(lb) : 0.0;
    g.zr = rb < m ? lam * sign(rb) : 0.0;
    bval[i] = GroupValue(g, R);
    if (zlsign != nullptr) zlsign[i] = lb >= 0 ? (int8_t)sign(lb) : (int8_t)0;
  }
  __syncthreads();
}

// Round, part 1 (dirty tiles): z on every free boundary and per group the largest violation of each sign
// (warp-aggregated atomicMax on the slot of the group's first element; the bit pattern of a positive double
// orders like the value); on every active boundary the jump across it times its sign.
__global__ void __launch_bounds__(kTvpBlock, 3)
TvpViolationKernel(TvState S, int n, double lam, unsigned tag) {
  __shared__ TileShared sh;
  __shared__ double bval[kTile];
  __shared__ int8_t zlsign[kTile];
  if (S.ctl[1] != 0) return;
  const int cnt = S.ctl[2];
  const int m = n - 1;
  const int lane = threadIdx.x & 31;
  const DD* __restrict__ R = S.R;
  // software pipeline over the tiles of this CTA: the flags and carries of the NEXT tile are requested while the
  // current one is processed, and the coalesced R[k + 1] stream of a tile is requested before its scans
  int idx = blockIdx.x;
  int T = 0, cl = 0, cr = 0;
  unsigned long long flags = 0;
  if (idx < cnt) {
    T = S.list[idx];
    flags = TileFlags(S.act, T);
    cl = S.carry_l[T];
    cr = S.carry_r[T];
  }
  for (; idx < cnt; idx += gridDim.x) {
    const int t0 = T * kTile;
    DD rk[kItems];
#pragma unroll
    for (int jj = 0; jj < kItems; jj++) {
      const int k = t0 + jj * kTvpBlock + threadIdx.x;
      rk[jj] = k < m ? R[k + 1] : DD{0.0, 0.0};
    }
    int Tn = 0, cln = 0, crn = 0;
    unsigned long long flagsn = 0;
    if (idx + gridDim.x < cnt) {
      Tn = S.list[idx + gridDim.x];
      flagsn = TileFlags(S.act, Tn);
      cln = S.carry_l[Tn];
      crn = S.carry_r[Tn];
    }
    TileScans(sh, t0, m, flags, cl, cr);
    // sign of an active boundary: from the tile when it lies inside
    auto sign = [&](int b) -> double { return (double)((unsigned)(b - t0) < (unsigned)kTile ? sh.a[b - t0] : S.act[b]); };
    TileGroupValues(sh, bval, t0, n, m, cl, lam, R, sign, zlsign);
    double zz[kItems];
    bool cand = false;
#pragma unroll
    for (int jj = 0; jj < kItems; jj++) {
      const int i = jj * kTvpBlock + threadIdx.x;
      const int k = t0 + i;
      int gid = -1 - k;  // unique when not a candidate
      double vp = 0.0, vn = 0.0;
      zz[jj] = 0.0;
      const int a = k < m ? (int)sh.a[i] : 1;
      if (__all_sync(kFull, a == 0)) {
        // the common case, kept free of divergence: the whole warp sits on free boundaries
        const int lb = sh.lb[Pad(i)];
        const int s0 = lb + 1;
        const int slot = HeadSlot(lb, t0);
        const DD rs = R[s0];
        const double zk = lam * (double)zlsign[slot] + (double)(k - s0 + 1) * bval[slot] - ((rk[jj].hi - rs.hi) + (rk[jj].lo - rs.lo));
        zz[jj] = zk;
        vp = zk - lam;
        vn = -lam - zk;
        gid = s0;
      } else if (k < m) {
        if (a == 0) {
          const int lb = sh.lb[Pad(i)];
          const int s0 = lb + 1;
          const int slot = HeadSlot(lb, t0);
          const DD rs = R[s0];
          const double zk = lam * (double)zlsign[slot] + (double)(k - s0 + 1) * bval[slot] - ((rk[jj].hi - rs.hi) + (rk[jj].lo - rs.lo));
          zz[jj] = zk;
          vp = zk - lam;
          vn = -lam - zk;
          gid = s0;
        } else {
          // jump across an active boundary: value of the group right of k minus the one left of it
          const int lbx = i > 0 ? sh.lb[Pad(i - 1)] : cl;
          double br;
          if (i + 1 < kTile) {
            br = bval[i + 1];           // element k + 1 follows an active boundary: a head
          } else {
            Group gr;
            gr.s = k + 1; gr.e = cr == m ? n - 1 : cr;
            gr.zl = lam * (double)a;
            gr.zr = cr < m ? lam * sign(cr) : 0.0;
            br = GroupValue(gr, R);
          }
          zz[jj] = (br - bval[HeadSlot(lbx, t0)]) * (double)a;
          cand |= zz[jj] <= 0.0;
        }
      }
      vp = vp > 0.0 ? vp : 0.0;
      vn = vn > 0.0 ? vn : 0.0;
      if (__any_sync(kFull, vp > 0.0 || vn > 0.0)) {
        cand |= vp > 0.0 || vn > 0.0;
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
          if (vp > 0.0) atomicMax(&S.slot_pos[gid], PackSlot(tag, vp));
          if (vn > 0.0) atomicMax(&S.slot_neg[gid], PackSlot(tag, vn));
        }
      }
    }
    // z is only read back by the update pass, and only where something can change
    const int any = __syncthreads_or(cand);   // also: the shared tile is rebuilt by the next iteration
    if (any && threadIdx.x == 0) S.clist[atomicAdd(&S.ctl[3], 1)] = T;
    if (any) {
#pragma unroll
      for (int jj = 0; jj < kItems; jj++) {
        const int k = t0 + jj * kTvpBlock + threadIdx.x;
        if (k < m) S.z[k] = zz[jj];
      }
    }
    T = Tn; flags = flagsn; cl = cln; cr = crn;
  }
}

// Round, part 2 (dirty tiles): the new active set, in place; the tile aggregates; the tiles the next round has
// to visit; the number of changes.  Items are blocked here (thread t owns kItems consecutive boundaries): one
// 8-byte load of the flags, 64 contiguous bytes of z.
__global__ void __launch_bounds__(kTvpBlock, 4)
TvpUpdateKernel(TvState S, int n, double lam, int round, unsigned tag) {
  __shared__ TileShared sh;
  __shared__ int red_l[kTvpBlock / 32], red_f[kTvpBlock / 32];
  if (S.ctl[1] != 0) return;
  const int cnt = S.ctl[3];   // tiles with a candidate; elsewhere nothing can change: aggregates and flags stay
  const int m = n - 1;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int base = threadIdx.x * kItems;
  uint8_t* nxt = S.dirty[(round + 1) & 1];
  int nchanged = 0;
  for (int idx = blockIdx.x; idx < cnt; idx += gridDim.x) {
    const int T = S.clist[idx];
    const int t0 = T * kTile;
    union { unsigned long long q; int8_t b[kItems]; } a;
    a.q = *reinterpret_cast<const unsigned long long*>(S.act + t0 + base);
    double zk[kItems];
    if (t0 + base + kItems <= m) {
#pragma unroll
      for (int j = 0; j < kItems; j += 2) {
        const double2 zz = *reinterpret_cast<const double2*>(S.z + t0 + base + j);
        zk[j] = zz.x;
        zk[j + 1] = zz.y;
      }
    } else {
#pragma unroll
      for (int j = 0; j < kItems; j++) zk[j] = t0 + base + j < m ? S.z[t0 + base + j] : 0.0;
    }
    const int cl = S.carry_l[T], cr = S.carry_r[T];
    TileScans(sh, t0, m, a.q, cl, cr);
    int last = -1, first = m;
    bool wrote = false;
#pragma unroll
    for (int j = 0; j < kItems; j++) {
      const int i = base + j;
      const int k = t0 + i;
      if (k >= m) continue;
      int8_t out = a.b[j];
      int lo = 0, hi = 0;   // boundaries whose tiles the change touches
      if (out != 0) {
        if (zk[j] <= 0.0) {   // wrong-signed (or vanished) jump: release; the two groups beside k merge
          out = 0;
          lo = i > 0 ? sh.lb[Pad(i - 1)] : cl;
          hi = i + 1 < kTile ? sh.rb[Pad(i + 1)] : cr;
        }
      } else {
        const double vp = zk[j] - lam, vn = -lam - zk[j];
        if (vp > 0.0 || vn > 0.0) {
          lo = sh.lb[Pad(i)];
          hi = sh.rb[Pad(i)];
          const int s = lo + 1;
          if (vp > 0.0 && PackSlot(tag, vp) == S.slot_pos[s]) out = 1;
          else if (vn > 0.0 && PackSlot(tag, vn) == S.slot_neg[s]) out = -1;
        }
      }
      if (out != a.b[j]) {
        a.b[j] = out;
        wrote = true;
        nchanged++;
        const int T0 = max(lo, 0) / kTile, T1 = min(hi, m - 1) / kTile;
        for (int TT = T0; TT <= T1; TT++) nxt[TT] = 1;
      }
      if (out != 0) {
        last = max(last, k);
        first = min(first, k);
      }
    }
    if (wrote) *reinterpret_cast<unsigned long long*>(S.act + t0 + base) = a.q;
    // the new aggregates of the tile
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      last = max(last, __shfl_xor_sync(kFull, last, off));
      first = min(first, __shfl_xor_sync(kFull, first, off));
    }
    if (lane == 0) {
      red_l[w] = last;
      red_f[w] = first;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int i = 1; i < kTvpBlock / 32; i++) {
        last = max(last, red_l[i]);
        first = min(first, red_f[i]);
      }
      S.tlast1[T] = last + 1;
      S.tfirst1[T] = m - first;
    }
    __syncthreads();
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) nchanged += __shfl_xor_sync(kFull, nchanged, off);
  if (lane == 0 && nchanged) atomicAdd(&S.ctl[0], nchanged);
}

// beta of every sample, from the settled active set.  POST: the VectorProx post-map and the dual update are
// applied here (x = C (beta - g), y = A x, u -= y; tv1d.cu TvPostKernel) and beta itself is never stored.
template <bool POST>
__global__ void __launch_bounds__(kTvpBlock, 4)
TvpValuesKernel(TvState S, int n, int ntiles, double lam, double* __restrict__ beta, ElemSeg sg,
                const double* __restrict__ consts, double* __restrict__ u, double* __restrict__ y_new,
                double* __restrict__ x, int speculative) {
  __shared__ TileShared sh;
  __shared__ double bval[kTile];
  // queued behind a batch of rounds before the host knows whether they settled: a no-op if they did not
  if (speculative && S.ctl[1] == 0) return;
  const int m = n - 1;
  // software pipeline as in TvpViolationKernel: next tile's flags / carries, and this tile's u (and g) stream,
  // are in flight while the scans run
  int T = blockIdx.x, cl = 0, cr = 0;
  unsigned long long flags = 0;
  if (T < ntiles) {
    flags = TileFlags(S.act, T);
    cl = S.carry_l[T];
    cr = S.carry_r[T];
  }
  for (; T < ntiles; T += gridDim.x) {
    const int t0 = T * kTile;
    double uu[kItems], gg[kItems];
    if (POST) {
#pragma unroll
      for (int jj = 0; jj < kItems; jj++) {
        const int i = t0 + jj * kTvpBlock + threadIdx.x;
        uu[jj] = i < n ? u[sg.u_off + i] : 0.0;
        gg[jj] = (i < n && sg.g_off >= 0) ? consts[sg.g_off + i] : 0.0;
      }
    }
    const int Tn = T + gridDim.x;
    int cln = 0, crn = 0;
    unsigned long long flagsn = 0;
    if (Tn < ntiles) {
      flagsn = TileFlags(S.act, Tn);
      cln = S.carry_l[Tn];
      crn = S.carry_r[Tn];
    }
    TileScans(sh, t0, m, flags, cl, cr);
    auto sign = [&](int b) -> double { return (double)((unsigned)(b - t0) < (unsigned)kTile ? sh.a[b - t0] : S.act[b]); };
    TileGroupValues(sh, bval, t0, n, m, cl, lam, S.R, sign);
#pragma unroll
    for (int jj = 0; jj < kItems; jj++) {
      const int ii = jj * kTvpBlock + threadIdx.x;
      const int i = t0 + ii;
      if (i >= n) continue;
      const int lb = ii > 0 ? sh.lb[Pad(ii - 1)] : cl;   // last active boundary < i
      const double bv = bval[HeadSlot(lb, t0)];
      if (POST) {
        const double xv = sg.Cs * (bv - gg[jj]);
        const double yn = sg.As * xv;
        x[sg.x_off + i] = xv;
        y_new[sg.y_off + i] = yn;
        u[sg.u_off + i] = uu[jj] - yn;
      } else {
        beta[i] = bv;
      }
    }
    __syncthreads();
    flags = flagsn; cl = cln; cr = crn;
  }
}

// ---- R[k] = sum_{i<k} v_i in double-double: tile sums, one CTA scanning them, tiles scanned with their offset.
__device__ __forceinline__ DD WarpSumDD(DD x) {
  const DDAdd add;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    DD y;
    y.hi = __shfl_xor_sync(kFull, x.hi, off);
    y.lo = __shfl_xor_sync(kFull, x.lo, off);
    x = add(x, y);
  }
  return x;
}

// The VectorProx pre-map fused with the dual update (tv1d.cu TvPreKernel: w = u + y_old, u = w, v = B w + g) and
// the tile sums of v for the prefix pass.
__global__ void __launch_bounds__(kTvpBlock)
TvpPreSumKernel(int n, int ntiles, ElemSeg sg, const double* __restrict__ consts, double* __restrict__ u,
                const double* __restrict__ y_old, double* __restrict__ v, DD* __restrict__ tsum) {
  __shared__ DD ws[kTvpBlock / 32];
  const DDAdd add;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int T = blockIdx.x; T < ntiles; T += gridDim.x) {
    const int t0 = T * kTile;
    double xv[kItems];
#pragma unroll
    for (int jj = 0; jj < kItems; jj++) {
      const int i = t0 + jj * kTvpBlock + threadIdx.x;
      xv[jj] = 0.0;
      if (i < n) {
        const double ww = u[sg.u_off + i] + y_old[sg.y_off + i];
        const double gc = sg.g_off >= 0 ? consts[sg.g_off + i] : 0.0;
        u[sg.u_off + i] = ww;
        xv[jj] = sg.Bs * ww + gc;   // v: not stored, TvpPrefixKernel<true> forms it again from u
      }
    }
    DD acc{0.0, 0.0};
#pragma unroll
    for (int jj = 0; jj < kItems; jj++) acc = add(acc, DD{xv[jj], 0.0});
    acc = WarpSumDD(acc);
    if (lane == 0) ws[w] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int i = 1; i < kTvpBlock / 32; i++) acc = add(acc, ws[i]);
      tsum[T] = acc;
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kTvpBlock)
TvpTileSumKernel(int n, int ntiles, const double* __restrict__ v, DD* __restrict__ tsum) {
  __shared__ DD ws[kTvpBlock / 32];
  const DDAdd add;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int T = blockIdx.x; T < ntiles; T += gridDim.x) {
    const int t0 = T * kTile;
    DD acc{0.0, 0.0};
#pragma unroll
    for (int jj = 0; jj < kItems; jj += 2) {
      const int i = t0 + (jj / 2) * (2 * kTvpBlock) + 2 * threadIdx.x;
      double2 x = make_double2(0.0, 0.0);
      if (i + 1 < n) x = *reinterpret_cast<const double2*>(v + i);
      else if (i < n) x.x = v[i];
      acc = add(acc, DD{x.x, 0.0});
      acc = add(acc, DD{x.y, 0.0});
    }
    acc = WarpSumDD(acc);
    if (lane == 0) ws[w] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int i = 1; i < kTvpBlock / 32; i++) acc = add(acc, ws[i]);
      tsum[T] = acc;
    }
    __syncthreads();
  }
}

// exclusive scan of the tile sums, in place (one CTA)
__global__ void __launch_bounds__(kCarryBlock)
TvpTileOffsetKernel(int ntiles, DD* __restrict__ tsum) {
  __shared__ DD ws[kCarryBlock / 32];
  const DDAdd add;
  const int t = threadIdx.x, lane = t & 31, w = t >> 5;
  const int per = (ntiles + kCarryBlock - 1) / kCarryBlock;
  const int lo = min(t * per, ntiles), hi = min(lo + per, ntiles);
  DD acc{0.0, 0.0};
  for (int T = lo; T < hi; T++) acc = add(acc, tsum[T]);
  DD x = acc;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    DD y;
    y.hi = __shfl_up_sync(kFull, x.hi, off);
    y.lo = __shfl_up_sync(kFull, x.lo, off);
    if (lane >= off) x = add(y, x);
  }
  if (lane == 31) ws[w] = x;
  __syncthreads();
  DD pre{0.0, 0.0};
  for (int i = 0; i < w; i++) pre = add(pre, ws[i]);
  DD ex;
  ex.hi = __shfl_up_sync(kFull, x.hi, 1);
  ex.lo = __shfl_up_sync(kFull, x.lo, 1);
  if (lane == 0) ex = DD{0.0, 0.0};
  DD run = add(pre, ex);
  for (int T = lo; T < hi; T++) {
    const DD own = tsum[T];
    tsum[T] = run;
    run = add(run, own);
  }
}

// R[t0 + i + 1] for the samples of a tile; tile 0 also writes R[0] = 0
// FUSED: v[i] = B u[i] + g[i] (u already holds w = u + y_old, TvpPreSumKernel)
template <bool FUSED>
__global__ void __launch_bounds__(kTvpBlock, 4)
TvpPrefixKernel(int n, int ntiles, const double* __restrict__ v, const DD* __restrict__ toff, DD* __restrict__ R,
                ElemSeg sg, const double* __restrict__ consts, const double* __restrict__ u) {
  __shared__ DD stage[kTile + kTile / 8];   // blocked writes (stride kItems DDs) and striped reads, padded
  __shared__ DD ws[kTvpBlock / 32];
  const DDAdd add;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int base = threadIdx.x * kItems;
  for (int T = blockIdx.x; T < ntiles; T += gridDim.x) {
    const int t0 = T * kTile;
    // coalesced (striped) loads staged through shared memory into the blocked order the scan wants
    double* xs = reinterpret_cast<double*>(stage);
#pragma unroll
    for (int jj = 0; jj < kItems; jj++) {
      const int ii = jj * kTvpBlock + threadIdx.x;
      const int i = t0 + ii;
      double xv = 0.0;
      if (i < n) xv = FUSED ? sg.Bs * u[sg.u_off + i] + (sg.g_off >= 0 ? consts[sg.g_off + i] : 0.0) : v[i];
      xs[ii + (ii >> 3)] = xv;
    }
    __syncthreads();
    double x[kItems];
#pragma unroll
    for (int j = 0; j < kItems; j++) x[j] = xs[base + j + ((base + j) >> 3)];
    __syncthreads();   // `stage` is written again below
    DD loc[kItems];
    DD acc{0.0, 0.0};
#pragma unroll
    for (int j = 0; j < kItems; j++) {
      acc = add(acc, DD{x[j], 0.0});
      loc[j] = acc;
    }
    DD sc = acc;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      DD y;
      y.hi = __shfl_up_sync(kFull, sc.hi, off);
      y.lo = __shfl_up_sync(kFull, sc.lo, off);
      if (lane >= off) sc = add(y, sc);
    }
    if (lane == 31) ws[w] = sc;
    __syncthreads();
    DD pre = toff[T];
    for (int i = 0; i < w; i++) pre = add(pre, ws[i]);
    DD ex;
    ex.hi = __shfl_up_sync(kFull, sc.hi, 1);
    ex.lo = __shfl_up_sync(kFull, sc.lo, 1);
    if (lane != 0) pre = add(pre, ex);
#pragma unroll
    for (int j = 0; j < kItems; j++) stage[base + j + ((base + j) >> 3)] = add(pre, loc[j]);
    __syncthreads();
#pragma unroll
    for (int jj = 0; jj < kItems; jj++) {
      const int i = jj * kTvpBlock + threadIdx.x;
      if (t0 + i < n) R[t0 + i + 1] = stage[i + (i >> 3)];
    }
    if (T == 0 && threadIdx.x == 0) R[0] = DD{0.0, 0.0};
    __syncthreads();
  }
}

size_t Align256(size_t b) { return (b + 255) / 256 * 256; }
int64_t TilesFor(int64_t n) { return (n + kTile - 1) / kTile; }

}  // namespace

// Bytes of scratch the parallel solver needs for a signal of n samples (beyond v and beta).
size_t Tv1dParallelBytes(int64_t n) {
  const int64_t nt = TilesFor(n);
  size_t b = 0;
  b += Align256(sizeof(DD) * (n + 1));                 // R
  b += Align256((size_t)nt * kTile);                   // act (persists: warm start)
  b += 6 * Align256(sizeof(int) * nt);                 // tlast1, tfirst1 (persist), carry_l, carry_r, list, clist
  b += 2 * Align256(nt);                               // dirty flags by round parity
  b += 256;                                            // ctl
  b += Align256(sizeof(DD) * nt);                      // tile sums / offsets of the prefix pass
  b += Align256(sizeof(double) * n);                   // z
  b += 2 * Align256(sizeof(unsigned long long) * n);   // slot pair (never cleared: entries carry a round tag)
  return b;
}

// Returns the number of rounds used, or -1 when the loop did not settle (the caller reports EPS_ERR_NUMERIC).
// `scratch` keeps the active set between calls (warm start); it must be zeroed to start cold.
// host_state[0]: last round tag used; host_state[1]: rounds the previous call needed (sizes the first batch).
// fuse != nullptr: v is produced here from (u, y_old) and the result goes straight to (x, y_new, u); `beta` is unused.
int Tv1dParallelSolve(Ctx* ctx, int n, double* v, double lam, double* beta, char* scratch,
                      int* h_flag /* pinned */, unsigned* host_state /* per solver */, const TvFuse* fuse) {
  const int m = n - 1;
  const int nt = (int)TilesFor(n);
  unsigned* tag_state = host_state;
  char* p = scratch;
  auto take = [&](size_t bytes) { char* q = p; p += Align256(bytes); return q; };
  TvState S;
  DD* R = reinterpret_cast<DD*>(take(sizeof(DD) * (n + 1)));
  S.R = R;
  S.act = reinterpret_cast<int8_t*>(take((size_t)nt * kTile));
  S.tlast1 = reinterpret_cast<int*>(take(sizeof(int) * nt));
  S.tfirst1 = reinterpret_cast<int*>(take(sizeof(int) * nt));
  S.carry_l = reinterpret_cast<int*>(take(sizeof(int) * nt));
  S.carry_r = reinterpret_cast<int*>(take(sizeof(int) * nt));
  S.list = reinterpret_cast<int*>(take(sizeof(int) * nt));
  S.dirty[0] = reinterpret_cast<uint8_t*>(take(nt));
  S.dirty[1] = reinterpret_cast<uint8_t*>(take(nt));
  S.clist = reinterpret_cast<int*>(take(sizeof(int) * nt));
  S.ctl = reinterpret_cast<int*>(take(256));
  DD* tsum = reinterpret_cast<DD*>(take(sizeof(DD) * nt));
  S.z = reinterpret_cast<double*>(take(sizeof(double) * n));
  S.slot_pos = reinterpret_cast<unsigned long long*>(take(sizeof(unsigned long long) * n));
  S.slot_neg = reinterpret_cast<unsigned long long*>(take(sizeof(unsigned long long) * n));
  cudaStream_t st = ctx->stream;
  // persistent grids sized to the resident CTA slots of each kernel: a CTA walks tiles, so an idle round costs
  // one wave of CTAs that read one int
  static int slots[6] = {0, 0, 0, 0, 0, 0};
  if (slots[0] == 0) {
    CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&slots[1], TvpViolationKernel, kTvpBlock, 0));
    CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&slots[2], TvpUpdateKernel, kTvpBlock, 0));
    CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&slots[3], TvpValuesKernel<true>, kTvpBlock, 0));
    CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&slots[4], TvpPrefixKernel<true>, kTvpBlock, 0));
    CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&slots[5], TvpPreSumKernel, kTvpBlock, 0));
    for (int i = 1; i < 6; i++) slots[i] = std::max(1, slots[i]);
    slots[0] = 1;
  }
  auto grid_for = [&](int k) { return std::min(nt, ctx->sm_count * slots[k]); };

  // R[k] = sum_{i<k} v_i in double-double: v is read twice (tile sums, then the tiles), R written once
  {
    ProfScope ps(ctx, KC_TV, (fuse ? 48.0 : 32.0) * n);
    if (fuse)
      TvpPreSumKernel<<<grid_for(5), kTvpBlock, 0, st>>>(n, nt, fuse->seg, fuse->consts, fuse->u, fuse->y_old, v, tsum);
    else
      TvpTileSumKernel<<<grid_for(5), kTvpBlock, 0, st>>>(n, nt, v, tsum);
    TvpTileOffsetKernel<<<1, kCarryBlock, 0, st>>>(nt, tsum);
    if (fuse)
      TvpPrefixKernel<true><<<grid_for(4), kTvpBlock, 0, st>>>(n, nt, nullptr, tsum, R, fuse->seg, fuse->consts, fuse->u);
    else
      TvpPrefixKernel<false><<<grid_for(4), kTvpBlock, 0, st>>>(n, nt, v, tsum, R, ElemSeg{}, nullptr, nullptr);
    ctx->count_launch(3);
    if (*tag_state + kMaxRounds + 64 >= (1u << 23)) {   // tag space used up: start over on cleared slots
      CUDA_CHECK(cudaMemsetAsync(S.slot_pos, 0, sizeof(unsigned long long) * n, st));
      CUDA_CHECK(cudaMemsetAsync(S.slot_neg, 0, sizeof(unsigned long long) * n, st));
      *tag_state = 0;
    }
  }
  auto launch_values = [&](int speculative) {
    if (fuse)
      TvpValuesKernel<true><<<grid_for(3), kTvpBlock, 0, st>>>(S, n, nt, lam, nullptr, fuse->seg, fuse->consts, fuse->u,
                                                                fuse->y_new, fuse->x, speculative);
    else
      TvpValuesKernel<false><<<grid_for(3), kTvpBlock, 0, st>>>(S, n, nt, lam, beta, ElemSeg{}, nullptr, nullptr, nullptr,
                                                                 nullptr, speculative);
    ctx->count_launch();
  };
  int rounds = -1;
  if (m > 0) {
    // Rounds are queued in batches and judged on the device; a settled solve turns the rest of its batch into
    // idle launches.  The first round of a call visits every tile (34n bytes: flags, R, z), later ones only the
    // dirty tiles.  The first batch is sized by what the previous call needed (consecutive ADMM iterations need
    // about the same number of rounds, slowly falling to 1-2).
    int r = 0;
    int batch = host_state[1] == 0 ? 4 : std::min<int>(24, std::max<int>(2, (int)host_state[1] + 1));
    static const bool trace = getenv("EPS_TV_TRACE") != nullptr;
    // The values pass is queued right behind each batch, before the host has read the verdict (it returns at once
    // when the rounds did not settle): the host's round trip for the flag overlaps with it instead of idling the GPU.
    static cudaEvent_t evs[64] = {};
    cudaEvent_t& ev = evs[ctx->device & 63];
    if (ev == nullptr) CUDA_CHECK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    bool first_batch = true;
    while (r < kMaxRounds && rounds < 0) {
      {
        ProfScope ps(ctx, KC_TV, r == 0 ? 34.0 * n : 0.0);
        for (int b = 0; b < batch; b++, r++) {
          const unsigned tag = ++(*tag_state);
          TvpCarryKernel<<<1, kCarryBlock, 0, st>>>(S, nt, m, r, r == 0);
          TvpViolationKernel<<<grid_for(1), kTvpBlock, 0, st>>>(S, n, lam, tag);
          TvpUpdateKernel<<<grid_for(2), kTvpBlock, 0, st>>>(S, n, lam, r, tag);
        }
        TvpCheckKernel<<<1, 1, 0, st>>>(S, r);
        ctx->count_launch(3 * batch + 1);
      }
      CUDA_CHECK(cudaMemcpyAsync(h_flag, S.ctl + 1, sizeof(int), cudaMemcpyDeviceToHost, st));
      CUDA_CHECK(cudaEventRecord(ev, st));
      {
        ProfScope ps(ctx, KC_TV, first_batch ? (fuse ? 57.0 : 25.0) * n : 0.0);
        launch_values(1);
      }
      first_batch = false;
      CUDA_CHECK(cudaEventSynchronize(ev));
      if (*h_flag != 0) rounds = *h_flag - 1;   // index of the round that changed nothing
      if (trace) fprintf(stderr, "[tv] n=%d queued=%d settled_after=%d\n", n, r, *h_flag);
      batch = std::min(24, batch * 2);
    }
    if (rounds < 0) return -1;
    host_state[1] = (unsigned)(rounds + 1);
  } else {
    rounds = 0;
    ProfScope ps(ctx, KC_TV, (fuse ? 57.0 : 25.0) * n);
    launch_values(0);
  }
  return rounds;
}

}  // namespace eps
