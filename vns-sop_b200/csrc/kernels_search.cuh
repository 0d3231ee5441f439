// __global__ entry points of the global-memory search chain (search.cuh); included by api_search.cu only.
#pragma once
#include "common.cuh"
#include "path.cuh"
#include "search.cuh"

// ---- chain kernels: exactly one CTA of VNS_CHAIN_THREADS; dynamic smem = 2*mp*warps doubles -----------
static __global__ void __launch_bounds__(VNS_CHAIN_THREADS) k_local_search(InstDev I, PathDev P, int kind, int num_changes, double budget,
                                                                    int skipflag, RngState *rng, LsStats *stats)
{
   extern __shared__ double smem_v[];
   __shared__ LsShared sh;
   if (threadIdx.x == 0) { sh.rng = *rng; sh.err = 0; sh.st = LsStats{0, 0, 0, 0, 0}; }
   __syncthreads();
   cta_local_search(I, P, kind, num_changes, budget, skipflag, sh, smem_v);
   if (threadIdx.x == 0) { *rng = sh.rng; *stats = sh.st; }
}

static __global__ void __launch_bounds__(VNS_CHAIN_THREADS) k_shake(InstDev I, PathDev P, int kind, double budget, RngState *rng, int *seg)
{
   extern __shared__ double smem_v[];
   __shared__ RngState g;
   if (threadIdx.x == 0) g = *rng;
   __syncthreads();
   cta_shake(I, P, kind, budget, g, seg, smem_v);
   if (threadIdx.x == 0) *rng = g;
}

static __global__ void __launch_bounds__(VNS_CHAIN_THREADS) k_fit(InstDev I, PathDev P, double budget)
{
   extern __shared__ double smem_v[];
   int T = P.hdr[HDR_T];
   cta_fit(I, P, T, budget, smem_v);
   if (threadIdx.x == 0) P.hd[HD_REWARD] = path_reward(I, P.tg, T);
}

static __global__ void __launch_bounds__(VNS_CHAIN_THREADS) k_greedy(InstDev I, PathDev P, double budget)
{
   extern __shared__ double smem_v[];
   cta_greedy(I, P, budget, smem_v);
}

static __global__ void __launch_bounds__(VNS_CHAIN_THREADS) k_vns_run(InstDev I, PathDev tour, PathDev act, VnsParamsDev prm, RngState *rng,
                                                               VnsResultDev *res, double *impr_reward, double *impr_length,
                                                               double *impr_ms, int *seg)
{
   extern __shared__ double smem_v[];
   __shared__ LsShared sh;
   if (threadIdx.x == 0) { sh.rng = *rng; sh.err = 0; sh.st = LsStats{0, 0, 0, 0, 0}; }
   __syncthreads();
   cta_vns_run(I, tour, act, prm, res, impr_reward, impr_length, impr_ms, sh, seg, smem_v);
   if (threadIdx.x == 0) *rng = sh.rng;
}
