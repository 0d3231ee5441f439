// Trajectory-exact local search with the WHOLE search state in shared memory (SURVEY.md 8a rows
// a15/a16): insertRandom (vns_sop_solver.cpp:469-684) and exchangeRandom (:686-864) as one CTA.
//
// Per round
//   1. warp 0 extends a ring of raw glibc rand() outputs (31 new values per step: with
//      x_i = x_{i-3} + x_{i-31} the next 31 values are three interleaved prefix sums of the last 31 -
//      four shuffle-adds) and turns the next 32 (from, to) draw pairs into up to W proposals in draw
//      order (ballot + popc; a pair with |from-to| <= gap is the reference's "redraw both");
//   2. every sub-warp group evaluates one proposal against the current state (groups_eval of
//      eval.cuh: lower bound + tryToMove / Add / Exchange / Replace, DP rows read from shared memory);
//   3. the first accepted proposal in draw order wins (shared-memory atomicMin); the RNG position moves
//      to just after ITS draws - everything drawn behind it is discarded, exactly as if the
//      reference's sequential loop had run - and the move is applied: targets / VNS vector shifted in
//      shared memory, forward DP re-scanned by warp 0 and backward DP by warp 1 with the running row
//      in registers.
// The backward table is stored by distance from the goal (slot = S - layer), so inserting a target
// never moves a row: rows before the edit point are recomputed anyway, rows behind it keep their slot.
#pragma once
#include "common.cuh"
#include "eval.cuh"
#include "search.cuh"

#define VNS_CH_THREADS 512
#define VNS_CH_WARPS (VNS_CH_THREADS / 32)
#define VNS_CH_RING 512
#define VNS_CH_MAXPROP 64   // (VNS_CH_WARPS * 4): proposals per round with 8-lane groups

struct ChainCtl {
   double L;            // current path length (actualLength)
   int T, nvns;
   int first, W, err;
   unsigned pos, gen;   // ring positions: consumed / generated raw draws
   int pfrom[VNS_CH_MAXPROP], pto[VNS_CH_MAXPROP], pdraws[VNS_CH_MAXPROP], pskip[VNS_CH_MAXPROP];
   unsigned ring[VNS_CH_RING];
   LsStats st;
   long long prof[4];   // cycles of thread 0: proposals, evaluation, selection, apply (vnsgpu_chain_cycles)
};

template <typename V> struct ChainSm {
   double *minF, *minBrev;   // R
   V *F, *Brev;              // R x mp
   int2 *cinfo;              // C
   int *lcl, *vns;           // R
};

template <typename V> __host__ __device__ inline size_t chain_smem_bytes(int C, int mp)
{
   const size_t R = C + 2;
   return 2 * R * sizeof(double) + 2 * R * mp * sizeof(V) + (size_t)C * sizeof(int2) + 2 * R * sizeof(int) + 16;
}

template <typename V> __device__ __forceinline__ ChainSm<V> chain_carve(unsigned char *base, int C, int mp)
{
   const size_t R = C + 2;
   ChainSm<V> s;
   s.minF = reinterpret_cast<double *>(base);
   s.minBrev = s.minF + R;
   s.F = reinterpret_cast<V *>(s.minBrev + R);
   s.Brev = s.F + R * mp;
   size_t o = (2 * R * sizeof(double) + 2 * R * mp * sizeof(V) + 7) & ~(size_t)7;
   s.cinfo = reinterpret_cast<int2 *>(base + o);
   s.lcl = reinterpret_cast<int *>(s.cinfo + C);
   s.vns = s.lcl + R;
   return s;
}

// ---- DP re-scan in shared memory: forward layers [lf, S] on warp 0, backward layers [lb, 0] on warp 1,
//      running row in registers (lane = sample; needs mp <= 32) ----------------------------------------
template <typename V>
__device__ void chain_rescan(const InstDev &I, const V *__restrict__ Dv, const ChainSm<V> &s, ChainCtl &ctl, int T, int lf, int lb)
{
   typedef ValOps<V> Ops;
   constexpr unsigned FULL = 0xffffffffu;
   const V INF = Ops::inf();
   const int S = T + 1, mp = I.mp;
   const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
   __syncthreads();
   if (warp == 0) {            // F[l] = F[l-1] (x) D(c_{l-1}, c_l)   vns_sop_path.cpp:1426-1448 / 1488-1511
      int2 cia = s.cinfo[s.lcl[lf - 1]];
      V v = lane < cia.y ? s.F[(lf - 1) * mp + lane] : INF;
      for (int l = lf; l <= S; ++l) {
         const int2 cib = s.cinfo[s.lcl[l]];
         const unsigned base = (unsigned)cia.x * (unsigned)I.N + (unsigned)cia.y * (unsigned)cib.x + (unsigned)min(lane, cib.y - 1);
         V acc = INF;
         for (int i0 = 0; i0 < cia.y; i0 += 8) {          // all 8 loads in flight before the first min
            V dd[8];
#pragma unroll
            for (int t = 0; t < 8; ++t) dd[t] = __ldg(Dv + (base + (unsigned)(min(i0 + t, cia.y - 1) * cib.y)));
#pragma unroll
            for (int t = 0; t < 8; ++t) acc = Ops::add_min(acc, __shfl_sync(FULL, v, min(i0 + t, cia.y - 1)), dd[t]);
         }
         v = lane < cib.y ? acc : INF;
         if (lane < cib.y) s.F[l * mp + lane] = acc;
         cia = cib;
      }
   } else if (warp == 1) {     // B[l][i] = min_j (B[l+1][j] + D(c_l, c_{l+1})[i][j])   :1453-1472 / 1515-1539
      int2 cib = s.cinfo[s.lcl[lb + 1]];
      V v = lane < cib.y ? s.Brev[(S - (lb + 1)) * mp + lane] : INF;
      for (int l = lb; l >= 0; --l) {
         const int2 cia = s.cinfo[s.lcl[l]];
         const unsigned base = (unsigned)cia.x * (unsigned)I.N + (unsigned)cia.y * (unsigned)cib.x + (unsigned)(min(lane, cia.y - 1) * cib.y);
         V acc = INF;
         for (int j0 = 0; j0 < cib.y; j0 += 8) {
            V dd[8];
#pragma unroll
            for (int t = 0; t < 8; ++t) dd[t] = __ldg(Dv + (base + (unsigned)min(j0 + t, cib.y - 1)));
#pragma unroll
            for (int t = 0; t < 8; ++t) acc = Ops::add_min(acc, __shfl_sync(FULL, v, min(j0 + t, cib.y - 1)), dd[t]);
         }
         v = lane < cia.y ? acc : INF;
         if (lane < cia.y) s.Brev[(S - l) * mp + lane] = acc;
         cib = cia;
      }
   }
   __syncthreads();
   // per-layer minima (the lower bounds read them); one thread per layer
   for (int l = threadIdx.x; l <= S; l += blockDim.x) {
      const int m = s.cinfo[s.lcl[l]].y;
      if (l >= lf) {
         V mn = INF;
         for (int j = 0; j < m; ++j) mn = Ops::vmin(mn, s.F[l * mp + j]);
         s.minF[l] = (double)mn;
      }
      if (l <= lb) {
         V mn = INF;
         for (int j = 0; j < m; ++j) mn = Ops::vmin(mn, s.Brev[(S - l) * mp + j]);
         s.minBrev[S - l] = (double)mn;
      }
   }
   __syncthreads();
   if (threadIdx.x == 0) {
      ctl.L = s.minF[S];                                                  // getPathLength :1660-1678
      if (fabs(s.minF[S] - s.minBrev[S]) > 0.01 && !ctl.err) ctl.err = 4;   // addPoint's check :80-88
   }
   __syncthreads();
}

// arr[lo .. hi] moves by one position (delta = +1 / -1); needs hi - lo < blockDim.x; all threads call
__device__ __forceinline__ void chain_shift1(int *arr, int lo, int hi, int delta)
{
   const int k = lo + (int)threadIdx.x;
   int v = 0;
   if (k <= hi) v = arr[k];
   __syncthreads();
   if (k <= hi) arr[k + delta] = v;
   __syncthreads();
}

// insertRandom accept (vns_sop_solver.cpp:621-646): VNS[from] is re-inserted immediately before the
// element that was at index `to`; only the final state matters.  tg = lcl + 1.
template <typename V>
__device__ void chain_apply_insert(const InstDev &I, const V *__restrict__ Dv, const ChainSm<V> &s, ChainCtl &ctl, int from, int to)
{
   const int T = ctl.T;
   int *tg = s.lcl + 1;
   const int c = s.vns[from];
   __syncthreads();
   if (from < T) {
      if (to > from) {        // travels up to index to-1; layers from+1 .. to change
         chain_shift1(s.vns, from + 1, to - 1, -1);
         chain_shift1(tg, from + 1, to - 1, -1);
         if (threadIdx.x == 0) { s.vns[to - 1] = c; tg[to - 1] = c; }
         chain_rescan(I, Dv, s, ctl, T, from + 1, to);
      } else {                // travels down to index `to`; layers to+1 .. from+1 change
         chain_shift1(s.vns, to, from - 1, +1);
         chain_shift1(tg, to, from - 1, +1);
         if (threadIdx.x == 0) { s.vns[to] = c; tg[to] = c; }
         chain_rescan(I, Dv, s, ctl, T, to + 1, from + 1);
      }
   } else {                   // from outside the path (from >= T >= to): one more target at index `to`
      chain_shift1(s.vns, to, from - 1, +1);
      chain_shift1(tg, to, T, +1);            // includes the goal's entry tg[T] = lcl[T+1]
      if (threadIdx.x == 0) { s.vns[to] = c; tg[to] = c; ctl.T = T + 1; }
      // B rows of the old layers to+1 .. S keep their slots (distance from the goal is unchanged)
      chain_rescan(I, Dv, s, ctl, T + 1, to + 1, to + 1);
   }
}

// exchangeRandom accept (:807-832): VNS[from] <-> VNS[to]; `to` is always in the path
template <typename V>
__device__ void chain_apply_exchange(const InstDev &I, const V *__restrict__ Dv, const ChainSm<V> &s, ChainCtl &ctl, int from, int to)
{
   const int T = ctl.T;
   int *tg = s.lcl + 1;
   __syncthreads();
   if (threadIdx.x == 0) {
      const int cf = s.vns[from], ct = s.vns[to];
      s.vns[to] = cf; s.vns[from] = ct;
      tg[to] = cf;
      if (from < T) tg[from] = ct;
   }
   const int lo = (from < T && from < to) ? from : to;
   const int hi = (from < T && from > to) ? from : to;
   chain_rescan(I, Dv, s, ctl, T, lo + 1, hi + 1);
}

// ---- glibc rand() ring (math_common.cpp:21-28; random_r TYPE_3: x_i = x_{i-3} + x_{i-31}, output >> 1) --
// ring[(p) & (RING-1)] = raw x value number p of this kernel's stream; numbers 0..30 are the state at
// entry.  Warp 0 only.
__device__ __forceinline__ void chain_ring_init(ChainCtl &ctl, const RngState &g, int lane)
{
   if (lane < 31) ctl.ring[lane] = (unsigned)g.r[(g.f + lane) % 31];   // oldest first: r[f] is x_{i-31}
   if (lane == 0) { ctl.gen = 31; ctl.pos = 31; }
   __syncwarp();
}

__device__ __forceinline__ void chain_ring_fill(ChainCtl &ctl, unsigned need_until, int lane)
{
   constexpr unsigned FULL = 0xffffffffu, M = VNS_CH_RING - 1;
   unsigned gen = ctl.gen;
   while (gen < need_until) {
      unsigned a = 0;
      if (lane < 31) a = ctl.ring[(gen - 31 + lane) & M];
      if (lane < 3) a += ctl.ring[(gen - 3 + lane) & M];
      // x_{gen+k} = a_k + x_{gen+k-3}: inclusive scan with stride 3
#pragma unroll
      for (int d = 3; d < 32; d <<= 1) {
         const unsigned t = __shfl_up_sync(FULL, a, d);
         if (lane >= d) a += t;
      }
      if (lane < 31) ctl.ring[(gen + lane) & M] = a;
      gen += 31;
      __syncwarp();
   }
   if (lane == 0) ctl.gen = gen;
   __syncwarp();
}

__device__ __forceinline__ void chain_ring_export(const ChainCtl &ctl, RngState &g, int lane)
{
   constexpr unsigned M = VNS_CH_RING - 1;
   if (lane < 31) g.r[lane] = (int32_t)ctl.ring[(ctl.pos - 31 + lane) & M];
   if (lane == 0) { g.f = 0; g.b = 28; g.pad = 0; }
   __syncwarp();
}

// ---- the local search --------------------------------------------------------------------------------
template <typename V, int G, int NR>
__device__ void chain_local_search(const InstDev &I, const V *__restrict__ Dv, const ChainSm<V> &s, ChainCtl &ctl, int kind,
                                   int num_changes, double budget, int skipflag)
{
   constexpr int NG = 32 / G, WCAP = VNS_CH_WARPS * NG;
   constexpr unsigned M = VNS_CH_RING - 1;
   const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, gl = lane % G, g = lane / G;
   const int nvns = ctl.nvns, mp = I.mp;
   const int gap = kind == 1 ? 1 : 2;
   __syncthreads();
   // guards :498 and :711
   const bool enabled = kind == 1 ? (nvns >= 3) : (nvns >= 4 && ctl.T > 0);
   int var = 0, empty_rounds = 0;
   while (enabled && var < num_changes) {
      int T = ctl.T;
      const long long t0 = clock64();
      // ---- 1. proposals in draw order -------------------------------------------------------------
      if (warp == 0) {
         chain_ring_fill(ctl, ctl.pos + 64, lane);
         const unsigned pos = ctl.pos;
         const unsigned r1 = ctl.ring[(pos + 2 * lane) & M] >> 1, r2 = ctl.ring[(pos + 2 * lane + 1) & M] >> 1;
         const int tohi = kind == 1 ? T : T - 1;
         const int from = (int)(((unsigned long long)r1 * (unsigned)nvns) >> 31);           // :507 / :723
         const int to = (int)(((unsigned long long)r2 * (unsigned)(tohi + 1)) >> 31);       // :508 / :724
         const bool valid = abs(from - to) > gap;                                          // :510-519 / :726-730
         const unsigned bal = __ballot_sync(0xffffffffu, valid);
         const int w = __popc(bal & ((1u << lane) - 1));
         const int W = min(min(__popc(bal), WCAP), num_changes - var);
         if (valid && w < W) { ctl.pfrom[w] = from; ctl.pto[w] = to; ctl.pdraws[w] = 2 * (lane + 1); }
         if (lane == 0) { ctl.W = W; ctl.first = W; }
      }
      __syncthreads();
      const int W = ctl.W;
      const long long t1 = clock64();
      // ---- 2. one group per proposal ---------------------------------------------------------------
      if (warp * NG < W) {
         const int gidx = warp * NG + g;
         const bool act = gidx < W;
         const int from = act ? ctl.pfrom[gidx] : 0, to = act ? ctl.pto[gidx] : 0;
         const bool in = from < T;
         const int cf = s.vns[from];
         int op = -1, a = 0;
         if (act) {
            if (kind == 1) { op = in ? VNSGPU_OP_MOVE : VNSGPU_OP_ADD; a = in ? from : cf; }
            else { op = in ? VNSGPU_OP_EXCHANGE : VNSGPU_OP_REPLACE; a = in ? from : cf; }
         }
         const CandCode cc = decode_cand(T, I.C, op, a, to);
         const int S = T + 1, K = cc.K;
         const int m0 = K ? s.cinfo[s.lcl[cc.s0]].y : 0, mK = K ? s.cinfo[s.lcl[cc.sK]].y : 0;
         V fin;
         double between;
         groups_eval<V, G, NR, true>(Dv, K, cc.tr, cc.ps, cc.pe, m0, mK, gl,
            [&](int k) {
               const int la = cand_layer(cc, k - 1), lb_ = cand_layer(cc, k);
               const int ca = la < 0 ? cc.foreign : s.lcl[la], cb = lb_ < 0 ? cc.foreign : s.lcl[lb_];
               const int2 cia = s.cinfo[ca], cib = s.cinfo[cb];
               EdgeDesc e;
               e.rb = (long long)cia.x * I.N + (long long)cia.y * cib.x;
               e.ma = cia.y; e.mb = cib.y;
               e.dm = __ldg(I.dmin + (size_t)ca * I.C + cb);
               return e;
            },
            [&](int j) { return s.F[cc.s0 * mp + j]; }, [&](int j) { return s.Brev[(S - cc.sK) * mp + j]; }, fin, between);
         if (act && gl == 0) {
            const double L = ctl.L;
            double dR = 0.0;
            if (!in) dR = kind == 1 ? __ldg(I.reward + cf)                                          // :581
                                    : 0.0 + (__ldg(I.reward + cf) - __ldg(I.reward + s.vns[to]));   // :784
            const double lbv = s.minF[cc.s0] + between + s.minBrev[S - cc.sK];   // :536,:570-571 (- 0.0),:745,:776
            const int skip = K == 0 || (skipflag && lbv > budget);               // :539,:574,:749,:779
            double len = (double)fin;
            if (kind == 1 && !in) len = (L - 0.0) + (len - L);                   // :580,:583 with lengthRemoved = L - L
            const bool acc = !skip && len <= budget && (dR > 0 || (L - len) > VNS_MIN_CHANGE_EPS);   // :619,:805-806
            ctl.pskip[gidx] = skip;
            if (acc) atomicMin(&ctl.first, gidx);
            if (K == 0 && !ctl.err) ctl.err = 3;
         }
      }
      __syncthreads();
      // ---- 3. first accepted in draw order ---------------------------------------------------------
      const long long t2 = clock64();
      const int first = ctl.first;
      const int used = first < W ? first + 1 : W;                    // proposals the reference consumed
      if (threadIdx.x < 32) {
         const int l2 = threadIdx.x + 32;
         const int sk = __popc(__ballot_sync(0xffffffffu, (int)threadIdx.x < used && ctl.pskip[threadIdx.x])) +
                        __popc(__ballot_sync(0xffffffffu, l2 < used && ctl.pskip[l2]));
         if (threadIdx.x == 0) {
         ctl.st.proposals += used; ctl.st.lb_skips += sk; ctl.st.evals += used - sk; ctl.st.rounds += 1;
         if (first < W) ctl.st.accepts += 1;
         ctl.pos += used ? (unsigned)ctl.pdraws[used - 1] : 64u;     // no valid pair among 32: all 64 draws consumed
         }
      }
      var += used;
      const long long t3 = clock64();
      empty_rounds = used ? 0 : empty_rounds + 1;
      if (empty_rounds > VNS_MAX_DRAWS / 64) { if (threadIdx.x == 0) ctl.err = 3; }
      if (first < W) {
         const int from = ctl.pfrom[first], to = ctl.pto[first];
         if (kind == 1) chain_apply_insert(I, Dv, s, ctl, from, to);
         else chain_apply_exchange(I, Dv, s, ctl, from, to);
      }
      __syncthreads();
      if (threadIdx.x == 0) {
         ctl.prof[0] += t1 - t0; ctl.prof[1] += t2 - t1; ctl.prof[2] += t3 - t2; ctl.prof[3] += clock64() - t3;
      }
      if (ctl.err) break;
   }
   __syncthreads();
}

// ---- global <-> shared path state -------------------------------------------------------------------------
template <typename V>
__device__ void chain_load(const InstDev &I, const PathDev &P, const ChainSm<V> &s, ChainCtl &ctl)
{
   const int T = P.hdr[HDR_T], nvns = P.hdr[HDR_NVNS], S = T + 1, mp = I.mp;
   for (int c = threadIdx.x; c < I.C; c += blockDim.x) s.cinfo[c] = make_int2(I.off[c], I.m[c]);
   for (int l = threadIdx.x; l <= S; l += blockDim.x) {
      s.lcl[l] = layer_cluster(I, P.tg, T, l);
      s.minF[l] = P.minF[l];
      s.minBrev[S - l] = P.minB[l];
   }
   for (int k = threadIdx.x; k < nvns; k += blockDim.x) s.vns[k] = P.vns[k];
   for (int e = threadIdx.x; e < (S + 1) * mp; e += blockDim.x) {
      const int l = e / mp, j = e - l * mp;
      s.F[e] = (V)P.F[e];
      s.Brev[(S - l) * mp + j] = (V)P.B[e];
   }
   if (threadIdx.x == 0) { ctl.T = T; ctl.nvns = nvns; ctl.L = P.hd[HD_LEN]; ctl.err = 0; }
   __syncthreads();
}

// entries beyond a layer's cluster size are never read by anyone and are not written back
template <typename V>
__device__ void chain_store(const InstDev &I, const PathDev &P, const ChainSm<V> &s, ChainCtl &ctl)
{
   __syncthreads();
   const int T = ctl.T, nvns = ctl.nvns, S = T + 1, mp = I.mp;
   for (int l = threadIdx.x; l <= S; l += blockDim.x) {
      if (l >= 1 && l <= T) P.tg[l - 1] = s.lcl[l];
      P.minF[l] = s.minF[l];
      P.minB[l] = s.minBrev[S - l];
   }
   for (int k = threadIdx.x; k < nvns; k += blockDim.x) P.vns[k] = s.vns[k];
   for (int e = threadIdx.x; e < (S + 1) * mp; e += blockDim.x) {
      const int l = e / mp, j = e - l * mp;
      if (j < s.cinfo[s.lcl[l]].y) {
         P.F[e] = (double)s.F[e];
         P.B[e] = (double)s.Brev[(S - l) * mp + j];
      }
   }
   __syncthreads();
   if (threadIdx.x == 0) {
      P.hdr[HDR_T] = T;
      P.hd[HD_LEN] = s.minF[S];
      P.hd[HD_LEN_REAR] = s.minBrev[S];
      P.hd[HD_REWARD] = path_reward(I, P.tg, T);                // :650,:835
      if (ctl.err && !P.hdr[HDR_ERR]) P.hdr[HDR_ERR] = ctl.err;
   }
   __syncthreads();
}

template <typename V, int G, int NR>
__global__ void __launch_bounds__(VNS_CH_THREADS) k_chain_ls(InstDev I, const V *__restrict__ Dv, PathDev P, int kind, int num_changes,
                                                             double budget, int skipflag, RngState *rng, LsStats *stats, long long *prof)
{
   extern __shared__ __align__(16) unsigned char smem_c[];
   __shared__ ChainCtl ctl;
   const ChainSm<V> s = chain_carve<V>(smem_c, I.C, I.mp);
   if (threadIdx.x < 32) {
      RngState g = *rng;
      chain_ring_init(ctl, g, threadIdx.x);
      if (threadIdx.x == 0) { ctl.st = LsStats{0, 0, 0, 0, 0}; ctl.prof[0] = ctl.prof[1] = ctl.prof[2] = ctl.prof[3] = 0; }
   }
   chain_load<V>(I, P, s, ctl);
   chain_local_search<V, G, NR>(I, Dv, s, ctl, kind, num_changes, budget, skipflag);
   chain_store<V>(I, P, s, ctl);
   if (threadIdx.x < 32) {
      __shared__ RngState gout;
      chain_ring_export(ctl, gout, threadIdx.x);
      if (threadIdx.x == 0) {
         *rng = gout; *stats = ctl.st;
         for (int k = 0; k < 4; ++k) prof[k] += ctl.prof[k];
      }
   }
}

// ---- VnsSopSolver::iterate (vns_sop_solver.cpp:82-260) as one persistent chain; the two local searches
//      run out of shared memory, shake / fit / greedy / path copies are the global-memory routines of
//      search.cuh (once per VNS step) --------------------------------------------------------------------
template <typename V, int G, int NR>
__device__ void chain_vns_body(const InstDev &I, const V *__restrict__ Dv, const PathDev &tour, const PathDev &act, const VnsParamsDev &prm,
                               RngState *rng, VnsResultDev *res, double *impr_reward, double *impr_length, double *impr_ms, int *seg,
                               long long *prof)
{
   extern __shared__ __align__(16) unsigned char smem_c[];
   __shared__ ChainCtl ctl;
   __shared__ RngState g;
   __shared__ int s_stop, s_improved;
   double *vbuf_all = reinterpret_cast<double *>(smem_c);                       // shake / fit / greedy scratch
   const ChainSm<V> s = chain_carve<V>(smem_c + sizeof(double) * 2 * I.mp * VNS_CH_WARPS, I.C, I.mp);
   if (threadIdx.x == 0) {
      g = *rng; s_stop = 0;
      ctl.st = LsStats{0, 0, 0, 0, 0}; ctl.prof[0] = ctl.prof[1] = ctl.prof[2] = ctl.prof[3] = 0; ctl.err = 0;
   }
   __syncthreads();
   const double t0 = now_ms();
   cta_update(I, tour, tour.hdr[HDR_T]);
   if (prm.greedy) cta_greedy(I, tour, prm.budget, vbuf_all);            // :127-129
   const int num_avail = tour.hdr[HDR_NVNS];                            // :115
   int last_impr = 1, it = 0, n_impr = 0;                               // :145
   double t_last = now_ms() - t0;
   if (threadIdx.x == 0 && impr_reward && n_impr < prm.max_improvements) {   // :150-154
      impr_reward[n_impr] = tour.hd[HD_REWARD]; impr_length[n_impr] = tour.hd[HD_LEN]; impr_ms[n_impr] = t_last;
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
         if (prm.stop_flag) {      // multi-start race: not part of the reference's loop
            if (threadIdx.x == 0) {
               if (tour.hd[HD_REWARD] >= prm.target_reward) atomicExch(prm.stop_flag, 1);
               if (*(volatile int *)prm.stop_flag) s_stop = 2;
            }
            __syncthreads();
            if (s_stop == 2) break;
         }
         cta_copy_path(I, act, tour);                                    // :195-197
         const double r0 = act.hd[HD_REWARD], l0 = act.hd[HD_LEN];
         cta_shake(I, act, k, prm.budget, g, seg, vbuf_all);             // :202
         if (threadIdx.x < 32) chain_ring_init(ctl, g, threadIdx.x);
         chain_load<V>(I, act, s, ctl);
         chain_local_search<V, G, NR>(I, Dv, s, ctl, k, num_avail * num_avail, prm.budget, prm.skip);   // :207,:346
         chain_store<V>(I, act, s, ctl);
         if (threadIdx.x < 32) chain_ring_export(ctl, g, threadIdx.x);
         __syncthreads();
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
      res->proposals = ctl.st.proposals; res->evals = ctl.st.evals; res->accepts = ctl.st.accepts;
      *rng = g;
      if (prof)
         for (int q = 0; q < 4; ++q) prof[q] += ctl.prof[q];
   }
   __syncthreads();
}

template <typename V, int G, int NR>
__global__ void __launch_bounds__(VNS_CH_THREADS) k_chain_vns_run(InstDev I, const V *__restrict__ Dv, PathDev tour, PathDev act,
                                                                  VnsParamsDev prm, RngState *rng, VnsResultDev *res, double *impr_reward,
                                                                  double *impr_length, double *impr_ms, int *seg, long long *prof)
{
   chain_vns_body<V, G, NR>(I, Dv, tour, act, prm, rng, res, impr_reward, impr_length, impr_ms, seg, prof);
}

// Multi-start (the reference's --batch=N: N independent solves, algorithm.cc:120-161): one CTA per run,
// each with its own seed, paths and scratch; no communication between runs.
struct VnsRunSlot {
   PathDev tour, act;
   RngState rng;
   VnsResultDev res;
   int *seg;
};

template <typename V, int G, int NR>
__global__ void __launch_bounds__(VNS_CH_THREADS) k_chain_vns_multi(InstDev I, const V *__restrict__ Dv, VnsRunSlot *slots, VnsParamsDev prm)
{
   VnsRunSlot &sl = slots[blockIdx.x];
   chain_vns_body<V, G, NR>(I, Dv, sl.tour, sl.act, prm, &sl.rng, &sl.res, nullptr, nullptr, nullptr, sl.seg, nullptr);
}
