// Micro-benchmark of the third chain form's building blocks on a synthetic krob200-like state (42 clusters, mixed sizes,
// T = 30): cycles of one forward / backward re-scan per layer, of one proposal evaluation per block step, of one window
// derivation per chunk of 32 draw pairs.  One thread block of C3_THREADS; warp 0 is timed.
// nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I ../../include -o c3_parts c3_parts.cu
#include <cstdio>
#include <vector>
#include <math_constants.h>
#include "../../include/vnsgpu.h"
#include "../../vns-sop_b200/csrc/chain3.cuh"

__global__ void __launch_bounds__(C3_THREADS) k_parts(C3Env E, const int *lcl_g, int T, long long *cyc, int *sink, int nrep)
{
   extern __shared__ __align__(16) unsigned char smem_c[];
   typedef int TT;
   const C3Sm<TT> s = c3_carve<TT>(smem_c, E.I, E.tab_g);
   c3_load_tables(E, s);
   const int S = T + 1, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
   for (int l = threadIdx.x; l <= S; l += blockDim.x) s.lcl[l] = lcl_g[l];
   for (int k = threadIdx.x; k < E.I.C; k += blockDim.x) s.vns[k] = k < T ? lcl_g[k + 1] : 0;
   for (int k = threadIdx.x; k < C3_RING; k += blockDim.x) s.ring[k] = 1234567u * (k + 1) + 89u * k * k;
   __syncthreads();
   if (warp == 0) {
      const int m0 = s.cinfo[s.lcl[0]].y, mS = s.cinfo[s.lcl[S]].y;
      if (lane < m0) s.F[lane] = 0;
      if (lane < mS) s.Brev[lane] = 0;
      if (lane == 0) { s.minF[0] = 0; s.minBrev[0] = 0; }
      __syncwarp();
      c3_fwd_scan(s, S, 1, lane);
      c3_bwd_scan(s, S, S - 1, lane);
      c3_pd_scan(s, S, lane);
   }
   __syncthreads();
   long long t_f = 0, t_b = 0, t_e = 0, t_w = 0, t_p = 0, t_e5 = 0, t_e2 = 0, t_d = 0;
   int acc = 0;
   for (int rep = 0; rep < nrep; ++rep) {
      __syncthreads();
      if (warp == 0) {
         long long t0 = clock64();
         c3_fwd_scan(s, S, 1, lane);
         long long t1 = clock64();
         c3_bwd_scan(s, S, S - 1, lane);
         long long t2 = clock64();
         t_f += t1 - t0; t_b += t2 - t1;
         // one evaluation: tryToExchange(3, 3 + span), span = 10 -> K = 12 blocks
         const CandCode cc = decode_cand(T, s.C, VNSGPU_OP_EXCHANGE, 3, 13);
         t0 = clock64();
         acc += c3_eval<TT, 16>(s, T, cc, lane % 16, s.gw + (lane / 16) * s.gwcap);
         t1 = clock64();
         t_e += t1 - t0;
         {  // shorter chains: tryToExchange(3, 6) -> K = 5, tryToAdd(cluster 5, 4) -> K = 2
            const CandCode c5 = decode_cand(T, s.C, VNSGPU_OP_EXCHANGE, 3, 6);
            t0 = clock64();
            acc += c3_eval<TT, 16>(s, T, c5, lane % 16, s.gw + (lane / 16) * s.gwcap);
            t1 = clock64();
            t_e5 += t1 - t0;
            const CandCode c2 = decode_cand(T, s.C, VNSGPU_OP_ADD, 5, 4);
            t0 = clock64();
            acc += c3_eval<TT, 16>(s, T, c2, lane % 16, s.gw + (lane / 16) * s.gwcap);
            t1 = clock64();
            t_e2 += t1 - t0;
            t0 = clock64();
            const CandCode c9 = decode_cand(T, s.C, VNSGPU_OP_MOVE, 3 + (lane & 1), 9);
            acc += c9.K;
            t_d += clock64() - t0;
         }
         // lower bounds of 32 proposals
         const int from = (lane * 7 + rep) % T, to = (from + 3 + lane % 5) % T;
         t0 = clock64();
         int lbv = 0;
         if (abs(from - to) > 2) lbv = c3_lower_bound(s, T, 2, from, to);
         const unsigned bal = __ballot_sync(0xffffffffu, lbv > 1000);
         t1 = clock64();
         t_w += t1 - t0;
         acc += lbv + (int)bal;
         t0 = clock64();
         c3_pd_scan(s, S, lane);
         t_p += clock64() - t0;
      }
      __syncthreads();
   }
   if (threadIdx.x == 0) { cyc[0] = t_f; cyc[1] = t_b; cyc[2] = t_e; cyc[3] = t_w; cyc[4] = t_p; cyc[5] = t_e5; cyc[6] = t_e2; cyc[7] = t_d; sink[0] = acc; }
}

int main()
{
   const int C = 42, T = 30;
   std::vector<int> m(C), off(C);
   int N = 0;
   for (int c = 0; c < C; ++c) { m[c] = 1 + (c * 5) % 9; if (c == 7) m[c] = 13; off[c] = N; N += m[c]; }
   std::vector<int> tab((size_t)N * N + 64), dmin(C * C, 3), lcl(T + 2);
   for (size_t k = 0; k < tab.size(); ++k) tab[k] = 10 + (int)((k * 7919) % 500);
   std::vector<int2> ci(C);
   std::vector<double> rw(C, 1.0);
   for (int c = 0; c < C; ++c) ci[c] = make_int2(off[c], m[c]);
   lcl[0] = 0; lcl[T + 1] = 0;
   for (int l = 1; l <= T; ++l) lcl[l] = 1 + (l * 11) % (C - 1);
   int *d_tab, *d_dmin, *d_lcl, *d_sink; int2 *d_ci; double *d_rw; long long *d_cyc;
   cudaMalloc(&d_tab, tab.size() * 4); cudaMalloc(&d_dmin, dmin.size() * 4); cudaMalloc(&d_lcl, lcl.size() * 4); cudaMalloc(&d_ci, C * 8);
   cudaMalloc(&d_rw, C * 8); cudaMalloc(&d_cyc, 64); cudaMalloc(&d_sink, 64);
   cudaMemcpy(d_tab, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice); cudaMemcpy(d_dmin, dmin.data(), dmin.size() * 4, cudaMemcpyHostToDevice);
   cudaMemcpy(d_lcl, lcl.data(), lcl.size() * 4, cudaMemcpyHostToDevice); cudaMemcpy(d_ci, ci.data(), C * 8, cudaMemcpyHostToDevice);
   cudaMemcpy(d_rw, rw.data(), C * 8, cudaMemcpyHostToDevice);
   C3Env E{};
   E.I.C = C; E.I.N = N; E.I.mp = 13; E.I.start = 0; E.I.goal = 0; E.I.reward = d_rw;
   E.tab_g = d_tab; E.dmin_g = d_dmin; E.cinfo_g = d_ci;
   const size_t smem = c3_smem_bytes(C, N, 13, 4);
   cudaFuncSetAttribute(k_parts, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
   const int nrep = 200;
   for (int it = 0; it < 2; ++it) { k_parts<<<1, C3_THREADS, smem>>>(E, d_lcl, T, d_cyc, d_sink, nrep); cudaDeviceSynchronize(); }
   long long h[8];
   cudaMemcpy(h, d_cyc, sizeof(h), cudaMemcpyDeviceToHost);
   printf("N = %d, T = %d (%s)\n", N, T, cudaGetErrorString(cudaGetLastError()));
   printf("forward re-scan   %8.1f cycles / layer  (%d layers)\n", (double)h[0] / nrep / (T + 1), T + 1);
   printf("backward re-scan  %8.1f cycles / layer\n", (double)h[1] / nrep / (T + 1));
   printf("evaluation        %8.1f cycles / call, K = 12 block steps\n", (double)h[2] / nrep);
   printf("evaluation        %8.1f cycles / call, K = 5;  %8.1f cycles / call, K = 2;  decode_cand %6.1f cycles\n", (double)h[5] / nrep, (double)h[6] / nrep, (double)h[7] / nrep);
   printf("lower bound + ballot %5.1f cycles / 32 proposals\n", (double)h[3] / nrep);
   printf("prefix sums       %8.1f cycles / call\n", (double)h[4] / nrep);
   return 0;
}
