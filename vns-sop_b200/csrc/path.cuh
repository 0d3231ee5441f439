// Layered min-plus DP of one path, warp-level building blocks.
// Notation: layers l = 0..S (S = T+1), c_l = cluster of layer l, F = shortest, B = shortest_back,
// (v (x) D)[j] = min_i (v[i] + D[i][j]).  All arithmetic is IEEE double with the reference's
// association; min is exact so its reduction order is free.
#pragma once
#include "common.cuh"

// ---- one DP layer, executed by one warp (lane = destination sample) ------------------------------
// F[l] = F[l-1] (x) D(c_{l-1}, c_l)   vns_sop_path.cpp:1353-1373 / 1426-1448 / 1488-1511
__device__ __forceinline__ void warp_fwd_layer(const InstDev &I, const PathDev &P, int T, int l, int lane)
{
   const int ca = layer_cluster(I, P.tg, T, l - 1), cb = layer_cluster(I, P.tg, T, l);
   const int ma = __ldg(I.m + ca), mb = __ldg(I.m + cb);
   const double *d = blk_ptr(I, ca, cb);
   const double *Fp = P.F + (size_t)(l - 1) * I.mp;
   double *Fl = P.F + (size_t)l * I.mp;
   double lmin = VNS_M;
   for (int j = lane; j < mb; j += 32) {
      double acc = VNS_M;
      for (int i = 0; i < ma; ++i) {
         double n = Fp[i] + __ldg(d + i * mb + j);
         acc = n < acc ? n : acc;
      }
      Fl[j] = acc;
      lmin = acc < lmin ? acc : lmin;
   }
   lmin = warp_min(lmin);
   if (lane == 0) P.minF[l] = lmin;
   __syncwarp();
}

// B[l][i] = min_j (B[l+1][j] + D(c_l, c_{l+1})[i][j])   vns_sop_path.cpp:1380-1399 / 1453-1472 / 1515-1539
__device__ __forceinline__ void warp_bwd_layer(const InstDev &I, const PathDev &P, int T, int l, int lane)
{
   const int ca = layer_cluster(I, P.tg, T, l), cb = layer_cluster(I, P.tg, T, l + 1);
   const int ma = __ldg(I.m + ca), mb = __ldg(I.m + cb);
   const double *d = blk_ptr(I, ca, cb);
   const double *Bn = P.B + (size_t)(l + 1) * I.mp;
   double *Bl = P.B + (size_t)l * I.mp;
   double lmin = VNS_M;
   for (int i = lane; i < ma; i += 32) {
      double acc = VNS_M;
      for (int j = 0; j < mb; ++j) {
         double n = Bn[j] + __ldg(d + i * mb + j);
         acc = n < acc ? n : acc;
      }
      Bl[i] = acc;
      lmin = acc < lmin ? acc : lmin;
   }
   lmin = warp_min(lmin);
   if (lane == 0) P.minB[l] = lmin;
   __syncwarp();
}

// forward layers [lfrom, S] by the calling warp (lfrom >= 1; lfrom == 0 also initialises layer 0)
static __device__ void warp_fwd_scan(const InstDev &I, const PathDev &P, int T, int lfrom, int lane)
{
   const int S = T + 1;
   if (lfrom <= 0) {  // F[0][*] = 0: default ClusterNodeDist, heuristic_types.h:184-187
      const int m0 = __ldg(I.m + I.start);
      for (int j = lane; j < m0; j += 32) P.F[j] = 0.0;
      if (lane == 0) P.minF[0] = 0.0;
      __syncwarp();
      lfrom = 1;
   }
   for (int l = lfrom; l <= S; ++l) warp_fwd_layer(I, P, T, l, lane);
   if (lane == 0) P.hd[HD_LEN] = P.minF[S];  // getPathLength vns_sop_path.cpp:1660-1678
}

// backward layers [lfrom, 0] by the calling warp (lfrom >= S initialises layer S)
static __device__ void warp_bwd_scan(const InstDev &I, const PathDev &P, int T, int lfrom, int lane)
{
   const int S = T + 1;
   if (lfrom >= S) {
      const int mS = __ldg(I.m + I.goal);
      for (int j = lane; j < mS; j += 32) P.B[(size_t)S * I.mp + j] = 0.0;
      if (lane == 0) P.minB[S] = 0.0;
      __syncwarp();
      lfrom = S - 1;
   }
   for (int l = lfrom; l >= 0; --l) warp_bwd_layer(I, P, T, l, lane);
   if (lane == 0) P.hd[HD_LEN_REAR] = P.minB[0];  // getPathLengthRear :1680-1698
}

// getReward vns_sop_path.cpp:2016-2031 (left-to-right sum in targets order), one thread
static __device__ double path_reward(const InstDev &I, const int *tg, int T)
{
   double r = 0;
   for (int t = 0; t < T; ++t) r += __ldg(I.reward + tg[t]);
   const double rs = __ldg(I.reward + I.start);
   if (rs > 0) {
      r += rs;
      if (I.start != I.goal) r += __ldg(I.reward + I.goal);
   }
   return r;
}

// ---- candidate evaluation ---------------------------------------------------------------------
// Every tryTo* is a chain  F[q_0] (x) D(q_0,q_1) (x) ... closed against B[q_K]; the node sequence q
// per operation (i, j = layers):
//   ADD(c,idx)      idx, c, idx+1                      tryToAdd        vns_sop_path.cpp:245-304
//   REPLACE(c,idx)  idx, c, idx+2                      tryToReplace    :456-513
//   REMOVE(idx)     idx, idx+2                         tryToRemove     :515-535
//   MOVE up         i-1, i+1..j-1, i, j                tryToMove       :988-1071
//   MOVE down       i-1, j, i..j-1, j+1                tryToMove       :1074-1151
//   EXCHANGE        i, j, i+2..j-1, i+1, j+1           tryToExchange   :635-797
struct CandPlan {
   int K;        // number of distance blocks (edges)
   int i, j;     // layer anchors
   int kind;     // 0 ADD 1 REPLACE 2 REMOVE 3 MOVE-up 4 MOVE-down 5 EXCHANGE
   int foreign;  // cluster id for ADD/REPLACE
   int rowF, rowB;
   int dBv;      // closing association (d + B) + v  (ADD/REPLACE/REMOVE) else (v + d) + B
   int bad;
};

__device__ __forceinline__ CandPlan make_plan(int T, int op, int a, int b, int C)
{
   CandPlan p;
   p.bad = 0; p.foreign = -1; p.i = 0; p.j = 0; p.dBv = 0; p.K = 0; p.kind = 0; p.rowF = 0; p.rowB = 0;
   if (op == 0) {         // idx in [0, T]  (:247)
      p.bad = (b < 0 || b > T || a < 0 || a >= C);
      p.kind = 0; p.K = 2; p.foreign = a; p.rowF = b; p.rowB = b + 1; p.dBv = 1; p.i = b;
   } else if (op == 1) {  // idx in [0, T-1]
      p.bad = (b < 0 || b >= T || a < 0 || a >= C);
      p.kind = 1; p.K = 2; p.foreign = a; p.rowF = b; p.rowB = b + 2; p.dBv = 1; p.i = b;
   } else if (op == 2) {
      p.bad = (a < 0 || a >= T);
      p.kind = 2; p.K = 1; p.rowF = a; p.rowB = a + 2; p.dBv = 1; p.i = a;
   } else if (op == 3) {  // :938-945
      p.bad = (a >= T || a < 0 || b > T || b < 0 || abs(b - a) < 2);
      const int lo = a < b ? a : b, hi = a < b ? b : a;
      p.i = lo + 1; p.j = hi + 1;
      if (b > a) { p.kind = 3; p.K = p.j - p.i + 1; p.rowF = p.i - 1; p.rowB = p.j; }
      else       { p.kind = 4; p.K = p.j - p.i + 2; p.rowF = p.i - 1; p.rowB = p.j + 1; }
   } else if (op == 4) {  // :637-657 (idxTo < T)
      p.bad = (a >= T || a < 0 || b >= T || b < 0 || abs(a - b) < 2);
      p.i = a < b ? a : b; p.j = (a < b ? b : a) + 1;
      p.kind = 5; p.K = p.j - p.i + 1; p.rowF = p.i; p.rowB = p.j + 1;
   } else {
      p.bad = 1;
   }
   return p;
}

// layer index of node k (0..K) of the chain; -1 = the foreign cluster
__device__ __forceinline__ int plan_layer(const CandPlan &p, int k)
{
   switch (p.kind) {
   case 0: return k == 0 ? p.i : (k == 1 ? -1 : p.i + 1);
   case 1: return k == 0 ? p.i : (k == 1 ? -1 : p.i + 2);
   case 2: return k == 0 ? p.i : p.i + 2;
   case 3: return k == 0 ? p.i - 1 : (k <= p.j - p.i - 1 ? p.i + k : (k == p.j - p.i ? p.i : p.j));
   case 4: return k == 0 ? p.i - 1 : (k == 1 ? p.j : (k <= p.j - p.i + 1 ? p.i + k - 2 : p.j + 1));
   default: return k == 0 ? p.i : (k == 1 ? p.j : (k <= p.j - p.i - 1 ? p.i + k : (k == p.j - p.i ? p.i + 1 : p.j + 1)));
   }
}

__device__ __forceinline__ int plan_cluster(const InstDev &I, const PathDev &P, int T, const CandPlan &p, int k)
{
   const int l = plan_layer(p, k);
   return l < 0 ? p.foreign : layer_cluster(I, P.tg, T, l);
}

// Lower bound = minF(q_0) + dist_between + minB(q_K) with dist_between accumulated in the exact
// source order of tryTo*LowerBound (vns_sop_path.cpp:207-243, 423-454, 573-633, 799-930):
// MOVE-up adds its last two edges as one pre-summed term (:873-874), MOVE-down its first two
// (:912-913); everything else is left-to-right from 0.0.  Uniform across the warp.
static __device__ double plan_lower_bound(const InstDev &I, const PathDev &P, int T, const CandPlan &p)
{
   const int K = p.K;
   if (p.kind == 2) return 0.0;  // tryToRemove has no lower bound in the reference
   double between = 0.0;
   int prev = plan_cluster(I, P, T, p, 0);
   if (p.kind <= 1) {            // dmin[from][c] + dmin[c][to]  (:221-222, :450-451)
      const int c1 = plan_cluster(I, P, T, p, 1), c2 = plan_cluster(I, P, T, p, 2);
      between = __ldg(I.dmin + (size_t)prev * I.C + c1) + __ldg(I.dmin + (size_t)c1 * I.C + c2);
   } else {
      int k = 1;
      if (p.kind == 4) {         // first two edges pre-summed
         const int c1 = plan_cluster(I, P, T, p, 1), c2 = plan_cluster(I, P, T, p, 2);
         between += __ldg(I.dmin + (size_t)prev * I.C + c1) + __ldg(I.dmin + (size_t)c1 * I.C + c2);
         prev = c2; k = 3;
      }
      const int kend = p.kind == 3 ? K - 2 : K;   // MOVE-up: last two edges pre-summed
      for (; k <= kend; ++k) {
         const int c = plan_cluster(I, P, T, p, k);
         between += __ldg(I.dmin + (size_t)prev * I.C + c);
         prev = c;
      }
      if (p.kind == 3) {
         const int c1 = plan_cluster(I, P, T, p, K - 1), c2 = plan_cluster(I, P, T, p, K);
         between += __ldg(I.dmin + (size_t)prev * I.C + c1) + __ldg(I.dmin + (size_t)c1 * I.C + c2);
      }
   }
   return P.minF[p.rowF] + between + P.minB[p.rowB];
}

// Full evaluation by one warp.  vbuf: 2*I.mp doubles of shared memory private to the warp.
static __device__ double warp_eval(const InstDev &I, const PathDev &P, int T, const CandPlan &p, double *vbuf, int lane)
{
   double *v = vbuf, *w = vbuf + I.mp;
   int ca = plan_cluster(I, P, T, p, 0);
   int ma = __ldg(I.m + ca);
   const double *F0 = P.F + (size_t)p.rowF * I.mp;
   for (int j = lane; j < ma; j += 32) v[j] = F0[j];
   __syncwarp();
   for (int k = 1; k < p.K; ++k) {     // intermediate blocks: w = v (x) D(ca, cb)
      const int cb = plan_cluster(I, P, T, p, k);
      const int mb = __ldg(I.m + cb);
      const double *d = blk_ptr(I, ca, cb);
      for (int j = lane; j < mb; j += 32) {
         double acc = VNS_M;
         for (int i = 0; i < ma; ++i) {
            double n = v[i] + __ldg(d + i * mb + j);
            acc = n < acc ? n : acc;
         }
         w[j] = acc;
      }
      __syncwarp();
      double *t = v; v = w; w = t;
      ca = cb; ma = mb;
   }
   // closing block against B[rowB]
   const int cb = plan_cluster(I, P, T, p, p.K);
   const int mb = __ldg(I.m + cb);
   const double *d = blk_ptr(I, ca, cb);
   const double *Bt = P.B + (size_t)p.rowB * I.mp;
   double best = VNS_M;
   for (int j = lane; j < mb; j += 32) {
      const double bj = Bt[j];
      for (int i = 0; i < ma; ++i) {
         const double dij = __ldg(d + i * mb + j);
         const double n = p.dBv ? (dij + bj) + v[i] : (v[i] + dij) + bj;
         best = n < best ? n : best;
      }
   }
   best = warp_min(best);
   __syncwarp();
   return best;
}

// algorithmic bytes of one candidate: every block entry once + first F row + last B
// row + the result
static __device__ long long plan_bytes(const InstDev &I, const PathDev &P, int T, const CandPlan &p, int w)
{
   long long e = 0;
   int ca = plan_cluster(I, P, T, p, 0);
   const int m0 = __ldg(I.m + ca);
   int ma = m0;
   for (int k = 1; k <= p.K; ++k) {
      const int cb = plan_cluster(I, P, T, p, k);
      const int mb = __ldg(I.m + cb);
      e += (long long)ma * mb;
      ma = mb;
   }
   return w * e + 8 * (m0 + ma) + 8;   // w = bytes per table entry read (8 double, 4 int32 path)
}
