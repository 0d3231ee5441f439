// Neighbourhood drivers on the device: ONE thread block owns one search
// chain - its RNG, proposal window, lower-bound pruning, evaluation, first-accept selection, move
// application and DP re-scan - so an accepted move costs an in-CTA re-scan, not a host round trip.
// Semantics follow vns_sop_solver.cpp line by line; see the citations at each step.
#pragma once
#include "common.cuh"
#include "path.cuh"

#define VNS_MAX_DRAWS (1 << 22)  // guard of the rejection loops (:510-519, :726-730) against hangs

// ---- CTA-wide vector helpers (all threads must call) ---------------------------------------------
// move arr[from, from+count) by delta positions (memmove semantics)
template <typename Tp>
__device__ void cta_shift(Tp *arr, int from, int count, int delta)
{
   if (count <= 0 || delta == 0) { __syncthreads(); return; }
   const int nt = blockDim.x, tid = threadIdx.x;
   if (delta > 0) {
      for (int hi = from + count; hi > from; hi -= nt) {  // chunks from the top
         const int k = hi - 1 - tid;
         Tp v = Tp();
         if (k >= from) v = arr[k];
         __syncthreads();
         if (k >= from) arr[k + delta] = v;
         __syncthreads();
      }
   } else {
      for (int lo = from; lo < from + count; lo += nt) {
         const int k = lo + tid;
         Tp v = Tp();
         if (k < from + count) v = arr[k];
         __syncthreads();
         if (k < from + count) arr[k + delta] = v;
         __syncthreads();
      }
   }
}

// rows [from_row, from_row+nrows) of B / minB move by delta rows
static __device__ void cta_shift_B(const InstDev &I, const PathDev &P, int from_row, int nrows, int delta)
{
   cta_shift(P.B, from_row * I.mp, nrows * I.mp, delta * I.mp);
   cta_shift(P.minB, from_row, nrows, delta);
}

// forward layers [lf, S] on warp 0 and backward layers [lb, 0] on warp 1, concurrently; then the
// consistency check of VnsSopPath::addPoint (vns_sop_path.cpp:80-88, tolerance 0.01)
static __device__ void cta_rescan(const InstDev &I, const PathDev &P, int T, int lf, int lb)
{
   __syncthreads();
   const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
   if (warp == 0) warp_fwd_scan(I, P, T, lf, lane);
   else if (warp == 1) warp_bwd_scan(I, P, T, lb, lane);
   __syncthreads();
   if (threadIdx.x == 0) {
      P.hdr[HDR_T] = T;
      if (fabs(P.hd[HD_LEN] - P.hd[HD_LEN_REAR]) > 0.01) P.hdr[HDR_ERR] = 4;
   }
   __syncthreads();
}

// VnsSopPath::update (vns_sop_path.cpp:1321-1404) + getReward
static __device__ void cta_update(const InstDev &I, const PathDev &P, int T)
{
   cta_rescan(I, P, T, 0, T + 1);
   if (threadIdx.x == 0) P.hd[HD_REWARD] = path_reward(I, P.tg, T);
   __syncthreads();
}

// ---- applying an accepted move: only the final state matters -----------------
// insertRandom accept (vns_sop_solver.cpp:621-646): VNS[from] is re-inserted immediately before the
// element that was at index `to`.  Returns the new T.
static __device__ int cta_apply_insert(const InstDev &I, const PathDev &P, int T, int nvns, int from, int to)
{
   __shared__ int s_c;
   if (threadIdx.x == 0) s_c = P.vns[from];
   __syncthreads();
   const int c = s_c;
   if (from < T) {
      if (to > from) {       // element travels up to index to-1; layers from+1..to change
         cta_shift(P.vns, from + 1, to - 1 - from, -1);
         cta_shift(P.tg, from + 1, to - 1 - from, -1);
         if (threadIdx.x == 0) { P.vns[to - 1] = c; P.tg[to - 1] = c; }
         cta_rescan(I, P, T, from + 1, to);
      } else {               // travels down to index `to`; layers to+1..from+1 change
         cta_shift(P.vns, to, from - to, +1);
         cta_shift(P.tg, to, from - to, +1);
         if (threadIdx.x == 0) { P.vns[to] = c; P.tg[to] = c; }
         cta_rescan(I, P, T, to + 1, from + 1);
      }
      return T;
   }
   // from outside the path (from >= T >= to): the path grows by one target at index `to`
   cta_shift(P.vns, to, from - to, +1);
   cta_shift(P.tg, to, T - to, +1);
   cta_shift_B(I, P, to + 1, T + 2 - (to + 1), +1);   // old layers to+1..S keep their B rows
   if (threadIdx.x == 0) { P.vns[to] = c; P.tg[to] = c; }
   cta_rescan(I, P, T + 1, to + 1, to + 1);
   return T + 1;
}

// exchangeRandom accept (vns_sop_solver.cpp:807-832): VNS[from] <-> VNS[to]; `to` is always in the path
static __device__ void cta_apply_exchange(const InstDev &I, const PathDev &P, int T, int from, int to)
{
   if (threadIdx.x == 0) {
      const int cf = P.vns[from], ct = P.vns[to];
      P.vns[to] = cf; P.vns[from] = ct;
      P.tg[to] = cf;
      if (from < T) P.tg[from] = ct;
   }
   const int lo = (from < T && from < to) ? from : to;
   const int hi = (from < T && from > to) ? from : to;
   cta_rescan(I, P, T, lo + 1, hi + 1);
}

// ---- the local-search chain ---------------------------------------------------------------------

struct LsShared {
   RngState rng;                    // glibc rand() replay (math_common.cpp:21-28)
   RngState rng0;                   // state at the start of the current window
   int from[VNS_CHAIN_WARPS], to[VNS_CHAIN_WARPS];
   int draws[VNS_CHAIN_WARPS];      // cumulative rand() calls up to and including proposal w
   int skip[VNS_CHAIN_WARPS], acc[VNS_CHAIN_WARPS];
   int first, err;
   LsStats st;
};

// insertRandom (kind 1, vns_sop_solver.cpp:469-684) / exchangeRandom (kind 2, :686-864).
// The next VNS_CHAIN_WARPS proposals are drawn against the CURRENT state and evaluated in parallel,
// one warp each; the first accepted one in draw order is applied, the RNG is rewound to just after
// its draws, and everything behind it is discarded - exactly the state sequence of the reference's
// sequential loop.  vbuf_all: VNS_CHAIN_WARPS * 2 * I.mp doubles of shared memory.
static __device__ void cta_local_search(const InstDev &I, const PathDev &P, int kind, int num_changes, double budget, int skipflag,
                                 LsShared &sh, double *vbuf_all)
{
   const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
   double *vbuf = vbuf_all + (size_t)warp * 2 * I.mp;
   int T = P.hdr[HDR_T];
   const int nvns = P.hdr[HDR_NVNS];
   __syncthreads();
   // guards :498 and :711
   const bool enabled = kind == 1 ? (nvns >= 3) : (nvns >= 4 && T > 0);
   int var = 0;
   while (enabled && var < num_changes) {
      const int W = min(VNS_CHAIN_WARPS, num_changes - var);
      if (threadIdx.x == 0) {
         sh.rng0 = sh.rng;
         int cnt = 0;
         for (int w = 0; w < W; ++w) {
            int from, to;
            if (kind == 1) {                                   // :507-519
               from = rng_int(sh.rng, 0, nvns - 1);
               to = rng_int(sh.rng, 0, T);
               cnt += 2;
               while (abs(from - to) <= 1 && cnt < VNS_MAX_DRAWS) {
                  from = rng_int(sh.rng, 0, nvns - 1);
                  to = rng_int(sh.rng, 0, T);
                  cnt += 2;
               }
            } else {                                           // :723-730
               from = rng_int(sh.rng, 0, nvns - 1);
               to = rng_int(sh.rng, 0, T - 1);
               cnt += 2;
               while (abs(from - to) <= 2 && cnt < VNS_MAX_DRAWS) {
                  from = rng_int(sh.rng, 0, nvns - 1);
                  to = rng_int(sh.rng, 0, T - 1);
                  cnt += 2;
               }
            }
            sh.from[w] = from; sh.to[w] = to; sh.draws[w] = cnt;
         }
         if (cnt >= VNS_MAX_DRAWS) sh.err = 3;
         sh.first = W;
      }
      __syncthreads();
      if (sh.err) break;
      if (warp < W) {
         const int from = sh.from[warp], to = sh.to[warp];
         const double L = P.hd[HD_LEN];
         const bool in = from < T;
         const int cf = P.vns[from];
         CandPlan pl;
         double dR = 0.0;
         if (kind == 1) {
            pl = in ? make_plan(T, 3, from, to, I.C) : make_plan(T, 0, cf, to, I.C);
            if (!in) dR = __ldg(I.reward + cf);                                      // :581
         } else {
            pl = in ? make_plan(T, 4, from, to, I.C) : make_plan(T, 1, cf, to, I.C);
            if (!in) dR = 0.0 + (__ldg(I.reward + cf) - __ldg(I.reward + P.vns[to])); // :784
         }
         const double lb = plan_lower_bound(I, P, T, pl);       // :536,:570-571 (- 0.0),:745,:776
         int skip = skipflag && lb > budget;                    // :539,:574,:749,:779
         int acc = 0;
         if (!skip) {
            double len = warp_eval(I, P, T, pl, vbuf, lane);
            if (kind == 1 && !in) len = (L - 0.0) + (len - L);  // :580,:583 with lengthRemoved = L - L
            acc = len <= budget && (dR > 0 || (L - len) > VNS_MIN_CHANGE_EPS);       // :619,:805-806
         }
         if (lane == 0) {
            sh.skip[warp] = skip; sh.acc[warp] = acc;
            if (acc) atomicMin(&sh.first, warp);
         }
      }
      __syncthreads();
      const int first = sh.first;
      const int used = first < W ? first + 1 : W;               // proposals the reference consumed
      if (threadIdx.x == 0) {
         int sk = 0;
         for (int w = 0; w < used; ++w) sk += sh.skip[w];
         sh.st.proposals += used; sh.st.lb_skips += sk; sh.st.evals += used - sk; sh.st.rounds += 1;
         if (first < W) {
            sh.st.accepts += 1;
            RngState g = sh.rng0;                               // rewind to just after proposal `first`
            for (int d = 0; d < sh.draws[first]; ++d) rng_next(g);
            sh.rng = g;
         }
      }
      var += used;
      if (first < W) {
         const int from = sh.from[first], to = sh.to[first];
         __syncthreads();
         if (kind == 1) T = cta_apply_insert(I, P, T, nvns, from, to);
         else cta_apply_exchange(I, P, T, from, to);
         if (P.hdr[HDR_ERR]) break;
      }
      __syncthreads();
   }
   __syncthreads();
   if (threadIdx.x == 0) {
      P.hdr[HDR_T] = T;
      P.hd[HD_REWARD] = path_reward(I, P.tg, T);                // :650,:835
      if (sh.err && !P.hdr[HDR_ERR]) P.hdr[HDR_ERR] = sh.err;
   }
   __syncthreads();
}

// ---- shake + fit-to-budget ------------------------------------------------------------------------
static __device__ void vec_insert1(int *v, int &n, int pos, const int *src, int k)
{
   for (int x = n - 1; x >= pos; --x) v[x + k] = v[x];
   for (int x = 0; x < k; ++x) v[pos + x] = src[x];
   n += k;
}
static __device__ void vec_erase1(int *v, int &n, int a, int b_incl)
{
   const int k = b_incl - a + 1;
   for (int x = b_incl + 1; x < n; ++x) v[x - k] = v[x];
   n -= k;
}
// k distinct draws in [0, hi] (unordered_set insert loop :936-939 / :1007-1010), sorted (:945,:1015)
static __device__ void draw_distinct_sorted(RngState &g, int hi, int k, int *out)
{
   int n = 0, guard = 0;
   while (n < k && guard++ < VNS_MAX_DRAWS) {
      const int v = rng_int(g, 0, hi);
      bool dup = false;
      for (int i = 0; i < n; ++i) dup |= out[i] == v;
      if (!dup) out[n++] = v;
   }
   for (int a = 1; a < k; ++a)
      for (int b = a; b > 0 && out[b - 1] > out[b]; --b) { int t = out[b]; out[b] = out[b - 1]; out[b - 1] = t; }
}

// fitDOPtoBudget vns_sop_solver.cpp:364-385.  Needs F valid for layers <= T (forward DP) on entry;
// runs on warp 0 only (other warps wait at the barrier), leaves a fully re-scanned path.
static __device__ void cta_fit(const InstDev &I, const PathDev &P, int &T_io, double budget, double *vbuf)
{
   __shared__ int s_T;
   const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
   __syncthreads();
   if (warp == 0) {
      int T = T_io;
      const int nvns = P.hdr[HDR_NVNS];
      const int mS = __ldg(I.m + I.goal);
      // :366-368  pop the last target while the path is over budget
      while (P.minF[T + 1] > budget && T > 0) {
         --T;
         warp_fwd_layer(I, P, T, T + 1, lane);
      }
      // :372-383  append VNS[T] while tryToAdd(VNS[T], T) fits
      for (;;) {
         if (T >= nvns) break;
         const int c = P.vns[T];
         for (int j = lane; j < mS; j += 32) P.B[(size_t)(T + 1) * I.mp + j] = 0.0;  // goal row of this T
         __syncwarp();
         const CandPlan pl = make_plan(T, 0, c, T, I.C);
         const double d = warp_eval(I, P, T, pl, vbuf, lane);
         if (!(d <= budget)) break;
         if (lane == 0) P.tg[T] = c;
         __syncwarp();
         ++T;
         warp_fwd_layer(I, P, T, T, lane);
         warp_fwd_layer(I, P, T, T + 1, lane);
      }
      if (lane == 0) { s_T = T; P.hd[HD_LEN] = P.minF[T + 1]; }
   }
   __syncthreads();
   T_io = s_T;
   cta_rescan(I, P, T_io, T_io + 2, T_io + 1);   // forward already valid; full backward scan
}

// pathInsert(.., 1) (kind 1, vns_sop_solver.cpp:931-999) / pathExchange(.., 1) (kind 2, :1001-1089):
// the same vector splices on the VNS vector and (for the parts inside the path) on targets, then
// fitDOPtoBudget.  The splices are serial on thread 0 (O(|VNS|) integer moves, once per VNS step).
// seg: scratch of 2*rows ints in global memory.
static __device__ void cta_shake(const InstDev &I, const PathDev &P, int kind, double budget, RngState &rng, int *seg, double *vbuf)
{
   __shared__ int s_T2, s_first;
   __syncthreads();
   if (threadIdx.x == 0) {
      int T = P.hdr[HDR_T], nvns = P.hdr[HDR_NVNS];
      int *vns = P.vns, *tg = P.tg;
      const int sizeV = nvns - 1;
      int first = T + 2;   // lowest changed target index (for the forward re-scan)
      if (kind == 1 && sizeV >= 2) {
         int r[3];
         draw_distinct_sorted(rng, sizeV, 3, r);
         int fs = r[0], fe = r[1], to = r[2], up = 1;
         if (rng_int(rng, 0, 1) == 1) { to = r[0]; fs = r[1]; fe = r[2]; up = 0; }        // :951-956
         const int n = fe - fs + 1;
         for (int x = 0; x < n; ++x) seg[x] = vns[fs + x];
         first = min(fs, to);
         if (up) {                                                                      // :958-972
            vec_insert1(vns, nvns, to, seg, n);
            vec_erase1(vns, nvns, fs, fe);
            if (to < T) vec_insert1(tg, T, to, seg, n);
            if (fs < T) vec_erase1(tg, T, fs, min(fe, T - 1));
         } else {                                                                       // :974-990
            vec_erase1(vns, nvns, fs, fe);
            vec_insert1(vns, nvns, to, seg, n);
            if (fs < T) vec_erase1(tg, T, fs, min(fe, T - 1));
            if (to < T) vec_insert1(tg, T, to, seg, n);
         }
      } else if (kind == 2 && sizeV >= 3) {
         int r[4];
         draw_distinct_sorted(rng, sizeV, 4, r);
         int fs = r[0], fe = r[1], ts = r[2], te = r[3], up = 1;
         if (rng_int(rng, 0, 1) == 1) { fs = r[2]; fe = r[3]; ts = r[0]; te = r[1]; up = 0; }  // :1022-1028
         const int nf = fe - fs + 1, nt = te - ts + 1;
         int *vf = seg, *vt = seg + nf;
         for (int x = 0; x < nf; ++x) vf[x] = vns[fs + x];
         for (int x = 0; x < nt; ++x) vt[x] = vns[ts + x];
         first = min(fs, ts);
         if (up) {                                                                      // :1030-1051
            vec_erase1(vns, nvns, ts, te); vec_insert1(vns, nvns, ts, vf, nf);
            vec_erase1(vns, nvns, fs, fe); vec_insert1(vns, nvns, fs, vt, nt);
            if (ts < T) { vec_erase1(tg, T, ts, min(te, T - 1)); vec_insert1(tg, T, ts, vf, nf); }
            if (fs < T) { vec_erase1(tg, T, fs, min(fe, T - 1)); vec_insert1(tg, T, fs, vt, nt); }
         } else {                                                                       // :1053-1080
            vec_erase1(vns, nvns, fs, fe); vec_insert1(vns, nvns, fs, vt, nt);
            vec_erase1(vns, nvns, ts, te); vec_insert1(vns, nvns, ts, vf, nf);
            if (fs < T) { vec_erase1(tg, T, fs, min(fe, T - 1)); vec_insert1(tg, T, fs, vt, nt); }
            if (ts < T) { vec_erase1(tg, T, ts, min(te, T - 1)); vec_insert1(tg, T, ts, vf, nf); }
         }
      }
      s_T2 = T; s_first = first;
   }
   __syncthreads();
   int T = s_T2;
   // forward DP of the spliced path (layers below the first changed target are still valid)
   if ((threadIdx.x >> 5) == 0) warp_fwd_scan(I, P, T, min(s_first + 1, T + 1), threadIdx.x & 31);
   __syncthreads();
   cta_fit(I, P, T, budget, vbuf);                                                      // :992,:1082
   if (threadIdx.x == 0) P.hd[HD_REWARD] = path_reward(I, P.tg, T);
   __syncthreads();
}

// ---- greedy_insertion vns_sop_solver.cpp:1091-1155 --------------------------------------------------
// Each round evaluates the full (candidate x position) grid - (|VNS|-T)*(T+1) independent tryToAdd -
// and inserts the strict minimum of (d - L)/reward among d <= budget; ties keep the first grid index
// (candidate-major, position-minor: the reference's loop order with strict '<' at :1119).
static __device__ void cta_greedy(const InstDev &I, const PathDev &P, double budget, double *vbuf_all)
{
   __shared__ double s_best[VNS_CHAIN_WARPS];
   __shared__ double s_d[VNS_CHAIN_WARPS];
   __shared__ int s_idx[VNS_CHAIN_WARPS];
   __shared__ int s_pick;
   const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
   double *vbuf = vbuf_all + (size_t)warp * 2 * I.mp;
   int T = P.hdr[HDR_T];
   const int nvns = P.hdr[HDR_NVNS];
   __syncthreads();
   for (;;) {
      const double L = P.hd[HD_LEN];
      const int npos = T + 1, ngrid = (nvns - T) * npos;
      double best = VNS_M, bestd = 0;   // minimalAddDistPerReward = DBL_MAX (:1101)
      int bidx = -1;
      for (int g = warp; g < ngrid; g += nw) {
         const int av = T + g / npos, pos = g % npos;
         const int c = P.vns[av];
         const CandPlan pl = make_plan(T, 0, c, pos, I.C);
         const double d = warp_eval(I, P, T, pl, vbuf, lane);
         const double per = (d - L) / __ldg(I.reward + c);      // :1116-1117
         if (per < best && d <= budget) { best = per; bestd = d; bidx = g; }   // :1119
      }
      if (lane == 0) { s_best[warp] = best; s_idx[warp] = bidx; s_d[warp] = bestd; }
      __syncthreads();
      if (threadIdx.x == 0) {
         double b = VNS_M; int bi = -1;
         for (int w = 0; w < nw; ++w)
            if (s_idx[w] >= 0 && (s_best[w] < b || (s_best[w] == b && bi >= 0 && s_idx[w] < bi))) { b = s_best[w]; bi = s_idx[w]; }
         s_pick = bi;
      }
      __syncthreads();
      const int pick = s_pick;
      if (pick < 0) break;                                       // :1129
      const int av = T + pick / npos, pos = pick % npos;
      T = cta_apply_insert(I, P, T, nvns, av, pos);              // :1133-1146 (same final state)
      if (P.hdr[HDR_ERR]) break;
   }
   __syncthreads();
   if (threadIdx.x == 0) { P.hdr[HDR_T] = T; P.hd[HD_REWARD] = path_reward(I, P.tg, T); }
   __syncthreads();
}

// ---- path copy (VnsSopPath value semantics: vns_sop_solver.cpp:195-197,227-228) ------------------------
static __device__ void cta_copy_path(const InstDev &I, const PathDev &dst, const PathDev &src)
{
   __syncthreads();
   const int T = src.hdr[HDR_T], nvns = src.hdr[HDR_NVNS];
   const int rows = T + 2;
   for (int k = threadIdx.x; k < HDR_INTS; k += blockDim.x) dst.hdr[k] = src.hdr[k];
   for (int k = threadIdx.x; k < HD_DBLS; k += blockDim.x) dst.hd[k] = src.hd[k];
   for (int k = threadIdx.x; k < T; k += blockDim.x) dst.tg[k] = src.tg[k];
   for (int k = threadIdx.x; k < nvns; k += blockDim.x) dst.vns[k] = src.vns[k];
   for (int k = threadIdx.x; k < rows * I.mp; k += blockDim.x) { dst.F[k] = src.F[k]; dst.B[k] = src.B[k]; }
   for (int k = threadIdx.x; k < rows; k += blockDim.x) { dst.minF[k] = src.minF[k]; dst.minB[k] = src.minB[k]; }
   __syncthreads();
}

// getPathLengthCalculate vns_sop_path.cpp:1700-1741: arg-min at the goal layer (first index on ties),
// walk the forward back-pointers (smallest predecessor index achieving F[l][j], :1356-1371) and
// re-sum the raw distances goal -> start.  One thread.  node[] (T+2 ints) may be null.
static __device__ double path_walk_back(const InstDev &I, const PathDev &P, int T, int *cluster, int *node)
{
   const int S = T + 1;
   int cur = 0;
   {
      const double *Fs = P.F + (size_t)S * I.mp;
      double mn = VNS_M;
      const int mS = I.m[I.goal];
      for (int i = 0; i < mS; ++i) if (Fs[i] < mn) { mn = Fs[i]; cur = i; }
   }
   double sum = 0;
   for (int l = S; l >= 0; --l) {
      const int cb = layer_cluster(I, P.tg, T, l);
      if (cluster) cluster[l] = cb;
      if (node) node[l] = cur;
      if (l == 0) break;
      const int ca = layer_cluster(I, P.tg, T, l - 1);
      const int ma = I.m[ca], mb = I.m[cb];
      const double *d = blk_ptr(I, ca, cb);
      const double *Fp = P.F + (size_t)(l - 1) * I.mp;
      double acc = VNS_M; int pred = -1;
      for (int i = 0; i < ma; ++i) {
         const double n = Fp[i] + d[i * mb + cur];
         if (n < acc) { acc = n; pred = i; }
      }
      sum += d[pred * mb + cur];
      cur = pred;
   }
   return sum;
}

// ---- VnsSopSolver::iterate vns_sop_solver.cpp:82-260 as one persistent chain ------------------------
__device__ __forceinline__ double now_ms()
{
   unsigned long long t;
   asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
   return (double)t * 1e-6;
}

static __device__ void cta_vns_run(const InstDev &I, const PathDev &tour, const PathDev &act, const VnsParamsDev &prm, VnsResultDev *res,
                            double *impr_reward, double *impr_length, double *impr_ms, LsShared &sh, int *seg, double *vbuf_all)
{
   __shared__ int s_stop, s_improved;
   const double t0 = now_ms();
   int T = tour.hdr[HDR_T];
   cta_update(I, tour, T);
   if (prm.greedy) cta_greedy(I, tour, prm.budget, vbuf_all);            // :127-129
   const int num_avail = tour.hdr[HDR_NVNS];                            // :115
   int last_impr = 1, it = 0, n_impr = 0;                               // :145
   double t_last = now_ms() - t0;
   if (threadIdx.x == 0) {                                              // :150-154
      if (impr_reward && n_impr < prm.max_improvements) {
         impr_reward[n_impr] = tour.hd[HD_REWARD]; impr_length[n_impr] = tour.hd[HD_LEN]; impr_ms[n_impr] = t_last;
      }
      s_stop = 0;
   }
   n_impr = 1;
   __syncthreads();
   bool stop = false;
   while (!stop) {
      ++it;
      if (it >= prm.num_iterations) stop = true;                         // :178
      if (it - last_impr >= prm.num_iterations_unimproved) stop = true;  // :182
      int k = 1;
      while (k <= 2) {
         if (threadIdx.x == 0 && prm.max_time_ms > 0 && now_ms() - t0 >= prm.max_time_ms) s_stop = 1;  // :189-193
         cta_copy_path(I, act, tour);                                    // :195-197
         const double r0 = act.hd[HD_REWARD], l0 = act.hd[HD_LEN];
         cta_shake(I, act, k, prm.budget, sh.rng, seg, vbuf_all);        // :202
         cta_local_search(I, act, k, num_avail * num_avail, prm.budget, prm.skip, sh, vbuf_all);  // :207,:346
         if (threadIdx.x == 0) {
            const double r1 = act.hd[HD_REWARD], l1 = act.hd[HD_LEN];
            s_improved = (r1 > r0 || (r1 == r0 && l1 < l0));             // :213
         }
         __syncthreads();
         if (act.hdr[HDR_ERR]) { stop = true; break; }
         if (s_improved) {
            last_impr = it;
            t_last = now_ms() - t0;
            if (threadIdx.x == 0) {
               if (impr_reward && n_impr < prm.max_improvements) {
                  impr_reward[n_impr] = act.hd[HD_REWARD]; impr_length[n_impr] = act.hd[HD_LEN]; impr_ms[n_impr] = t_last;
               }
               // :229-237 sanity check of the improved tour
               const double calc = path_walk_back(I, act, act.hdr[HDR_T], nullptr, nullptr);
               const double nl = act.hd[HD_LEN];
               if (calc > prm.budget || nl > prm.budget || calc - nl > 0.01) act.hdr[HDR_ERR] = 4;
            }
            ++n_impr;
            cta_copy_path(I, tour, act);                                 // :227-228
            if (tour.hdr[HDR_ERR]) { stop = true; break; }
            k = 1;                                                       // :241
         } else {
            ++k;                                                         // :243
         }
      }
      __syncthreads();
      if (s_stop) stop = true;
   }
   __syncthreads();
   if (threadIdx.x == 0) {
      res->num_iters = it; res->num_iters_last_impr = last_impr; res->n_improvements = n_impr;
      res->status = tour.hdr[HDR_ERR] ? tour.hdr[HDR_ERR] : act.hdr[HDR_ERR];
      res->reward = tour.hd[HD_REWARD]; res->length = tour.hd[HD_LEN];
      res->time_ms = now_ms() - t0; res->time_last_impr_ms = t_last;
      res->proposals = sh.st.proposals; res->evals = sh.st.evals; res->accepts = sh.st.accepts;
   }
   __syncthreads();
}
