// __global__ entry points over the device functions of path.cuh / search.cuh: path maintenance + the generic
// batched evaluation (included by api_path.cu only).
#pragma once
#include <climits>

#include "common.cuh"
#include "path.cuh"
#include "search.cuh"

// ---- path maintenance (single CTA per path) -------------------------------------------------------
// VnsSopPath::update vns_sop_path.cpp:1321-1404
static __global__ void __launch_bounds__(64) k_path_update(InstDev I, PathDev P)
{
   cta_update(I, P, P.hdr[HDR_T]);
}

// addPoint(c, idx) (kind 0) / removePoint(idx) (kind 1) with the incremental re-scan ranges of
// updateAfterInsert / updateAfterRemove (vns_sop_path.cpp:1406-1540).  Targets only.
static __global__ void __launch_bounds__(256) k_path_edit(InstDev I, PathDev P, int kind, int c, int idx)
{
   int T = P.hdr[HDR_T];
   __syncthreads();
   if (kind == 0) {
      cta_shift(P.tg, idx, T - idx, +1);
      cta_shift_B(I, P, idx + 1, T + 2 - (idx + 1), +1);
      if (threadIdx.x == 0) P.tg[idx] = c;
      T += 1;
      cta_rescan(I, P, T, idx + 1, idx + 1);          // :1426, :1453
   } else {
      cta_shift(P.tg, idx + 1, T - idx - 1, -1);
      cta_shift_B(I, P, idx + 2, T + 2 - (idx + 2), -1);
      T -= 1;
      cta_rescan(I, P, T, idx + 1, idx);              // :1488, :1515
   }
   if (threadIdx.x == 0) P.hd[HD_REWARD] = path_reward(I, P.tg, T);
}

static __global__ void __launch_bounds__(256) k_path_copy(InstDev I, PathDev dst, PathDev src)
{
   cta_copy_path(I, dst, src);
}

// getPath + getPathLengthCalculate (vns_sop_path.cpp:1542-1599, 1700-1741)
static __global__ void k_path_solution(InstDev I, PathDev P, int *cluster, int *node, double *calc)
{
   if (threadIdx.x == 0 && blockIdx.x == 0) *calc = path_walk_back(I, P, P.hdr[HDR_T], cluster, node);
}

// ---- batched candidate evaluation (a8-a13), wide-cluster fallback: one warp per candidate -----------
// len = tryTo*(...), lb = tryTo*LowerBound(...), flags = feasibility / accept predicate / LB skip;
// d_first receives the smallest accepted (and not LB-skipped when skip != 0) candidate index.
// generic fallback for clusters with more than 32 samples (the group kernel of eval.cuh covers the rest)
static __global__ void __launch_bounds__(VNS_EVAL_THREADS) k_eval_batch(InstDev I, PathDev P, long long n, long long idx0, const vnsgpu_cand *__restrict__ cands,
                                                                 double budget, int skipflag, double *__restrict__ len,
                                                                 double *__restrict__ lb, uint8_t *__restrict__ flags, unsigned long long *d_first)
{
   extern __shared__ double smem_v[];
   const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
   double *vbuf = smem_v + (size_t)warp * 2 * I.mp;
   const int T = P.hdr[HDR_T];
   const double L = P.hd[HD_LEN];
   const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
   unsigned long long best = ULLONG_MAX;   // d_first is preset to all-ones (= -1 as int64: none)
   for (long long k = (((long long)blockIdx.x * blockDim.x) >> 5) + warp; k < n; k += nwarps) {
      int4 cd = __ldg(reinterpret_cast<const int4 *>(cands) + k);   // op, a, b, gain_cluster
      if (cd.w >= I.C) cd.x = -1;                 // gain cluster out of range: a bad descriptor like any other bad index
      const CandPlan pl = make_plan(T, cd.x, cd.y, cd.z, I.C);
      double l, bound;
      unsigned f = 0;
      if (pl.bad) {
         l = CUDART_NAN; bound = CUDART_NAN; f = VNSGPU_F_BAD;
      } else {
         l = warp_eval(I, P, T, pl, vbuf, lane);
         bound = pl.kind == 2 ? l : plan_lower_bound(I, P, T, pl);
         double dR = 0.0, lm = l;                               // lm = lengthAfterMove of the drivers
         if (cd.x == 0) {                                        // insertRandom, outside -> path
            dR = cd.w >= 0 ? __ldg(I.reward + cd.w) : 0.0;       // vns_sop_solver.cpp:581
            lm = (L - 0.0) + (l - L);                            // :580,:583 (lengthRemoved = L - L)
         } else if (cd.x == 1) {                                 // exchangeRandom, outside replaces b
            const int ct = P.tg[cd.z];
            dR = cd.w >= 0 ? 0.0 + (__ldg(I.reward + cd.w) - __ldg(I.reward + ct)) : 0.0;  // :784
         }
         if (lm <= budget) f |= VNSGPU_F_FEASIBLE;
         if (lm <= budget && (dR > 0 || (L - lm) > VNS_MIN_CHANGE_EPS)) f |= VNSGPU_F_ACCEPT; // :619,:805
         if (bound > budget && pl.kind != 2) f |= VNSGPU_F_LB_SKIP;                          // :539,...
      }
      if (lane == 0) {
         if (len) len[k] = l;
         if (lb) lb[k] = bound;
         if (flags) flags[k] = (uint8_t)f;
         if ((f & VNSGPU_F_ACCEPT) && !(skipflag && (f & VNSGPU_F_LB_SKIP)) && (unsigned long long)(k + idx0) < best) best = (unsigned long long)(k + idx0);
      }
   }
   if (d_first && lane == 0 && best != ULLONG_MAX) atomicMin(d_first, best);
}

static __global__ void __launch_bounds__(256) k_cand_bytes(InstDev I, PathDev P, long long n, const vnsgpu_cand *__restrict__ cands,
                                                    unsigned long long *total, int w)
{
   const int T = P.hdr[HDR_T];
   long long acc = 0;
   for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x) {
      const int4 cd = __ldg(reinterpret_cast<const int4 *>(cands) + k);
      const CandPlan pl = make_plan(T, cd.x, cd.y, cd.z, I.C);
      if (!pl.bad) acc += plan_bytes(I, P, T, pl, w);
   }
   for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
   if ((threadIdx.x & 31) == 0) atomicAdd(total, (unsigned long long)acc);
}

