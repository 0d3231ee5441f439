// Micro-benchmark: cycles per DP layer of a single-warp min-plus scan over a shared-memory table, in the variants
// that chain3.cuh had to choose between.  nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o scan_latency scan_latency.cu
#include <cstdio>
#include <cuda_runtime.h>
#define INF 0x3f000000
#define NL 64
#define FULL 0xffffffffu

template <int VAR>
__global__ void __launch_bounds__(512) k_scan(const int *gtab, const unsigned *gw, int *out, long long *cyc, int nrep)
{
   extern __shared__ int sm[];
   int *tab = sm;                 // 40000 ints
   unsigned *wl = (unsigned *)(sm + 40000 + 2048);
   int *F = sm + 40000 + 2048 + 128;
   int *minF = F + NL * 32;
   for (int k = threadIdx.x; k < 40000 + 2048; k += blockDim.x) tab[k] = k < 40000 ? gtab[k] : 0;
   for (int k = threadIdx.x; k < NL; k += blockDim.x) wl[k] = gw[k];
   __syncthreads();
   const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
   long long total = 0;
   for (int rep = 0; rep < nrep; ++rep) {
      __syncthreads();
      if (warp == 0) {
         const long long t0 = clock64();
         int v = lane < 5 ? lane * 3 : INF;
         unsigned w = wl[0];
         int mn_prev = 0;
#pragma unroll 1
         for (int l = 0; l < NL; ++l) {
            const unsigned wn = wl[(l + 1) & (NL - 1)];
            const int ma = (w >> 20) & 63, mb = w >> 26;
            const int *p = tab + (w & 0xfffffu) + min(lane, mb - 1);
            int a0 = INF, a1 = INF;
            if (VAR == 3) {   // fixed 8 rows, unrolled, no branch
#pragma unroll
               for (int t = 0; t < 8; t += 2) {
                  a0 = __viaddmin_s32(__shfl_sync(FULL, v, t), p[t * mb], a0);
                  a1 = __viaddmin_s32(__shfl_sync(FULL, v, t + 1), p[(t + 1) * mb], a1);
               }
            } else if (ma <= 4) {
               const int d0 = p[0], d1 = p[mb], d2 = p[2 * mb], d3 = p[3 * mb];
               const int x0 = __shfl_sync(FULL, v, 0), x1 = __shfl_sync(FULL, v, 1), x2 = __shfl_sync(FULL, v, 2), x3 = __shfl_sync(FULL, v, 3);
               a0 = __viaddmin_s32(x0, d0, a0); a1 = __viaddmin_s32(x1, d1, a1);
               a0 = __viaddmin_s32(x2, d2, a0); a1 = __viaddmin_s32(x3, d3, a1);
            } else {
#pragma unroll 1
               for (int t0 = 0; t0 < ma; t0 += 8) {
                  int d[8], x[8];
#pragma unroll
                  for (int t = 0; t < 8; ++t) d[t] = p[t * mb];
#pragma unroll
                  for (int t = 0; t < 8; ++t) x[t] = __shfl_sync(FULL, v, t0 + t);
#pragma unroll
                  for (int t = 0; t < 8; t += 2) { a0 = __viaddmin_s32(x[t], d[t], a0); a1 = __viaddmin_s32(x[t + 1], d[t + 1], a1); }
                  p += 8 * mb;
               }
            }
            int acc = min(a0, a1);
            acc = lane < mb ? acc : INF;
            if (VAR != 2 && lane < mb) F[l * 32 + lane] = acc;
            if (VAR == 0) {
               if (lane == 0 && l > 0) minF[l - 1] = mn_prev;
               mn_prev = __reduce_min_sync(FULL, acc);
            }
            if (VAR == 4) {   // minimum by shuffles instead of REDUX
               int m = acc;
#pragma unroll
               for (int sft = 16; sft > 0; sft >>= 1) m = min(m, __shfl_xor_sync(FULL, m, sft));
               if (lane == 0) minF[l] = m;
            }
            v = acc;
            w = wn;
         }
         if (lane == 0) minF[NL - 1] = mn_prev;
         total += clock64() - t0;
         if (lane == 0) out[0] = v;
      }
      __syncthreads();
   }
   if (threadIdx.x == 0) cyc[0] = total;
}

template <int VAR> void run(const char *name, int threads, const int *gtab, const unsigned *gw, int *out, long long *cyc)
{
   const size_t smem = sizeof(int) * (40000 + 2048 + 128 + NL * 32 + NL + 64);
   cudaFuncSetAttribute(k_scan<VAR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
   const int nrep = 200;
   for (int it = 0; it < 2; ++it) {
      k_scan<VAR><<<1, threads, smem>>>(gtab, gw, out, cyc, nrep);
      cudaDeviceSynchronize();
   }
   long long h = 0;
   cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
   printf("%-44s threads %3d: %.1f cycles / layer (%s)\n", name, threads, (double)h / nrep / NL, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
   int *gtab, *out;
   unsigned *gw;
   long long *cyc;
   cudaMalloc(&gtab, sizeof(int) * 40000); cudaMalloc(&out, 64); cudaMalloc(&cyc, 64); cudaMalloc(&gw, sizeof(unsigned) * NL);
   int *h = new int[40000];
   for (int k = 0; k < 40000; ++k) h[k] = (k * 7919) % 1000;
   cudaMemcpy(gtab, h, sizeof(int) * 40000, cudaMemcpyHostToDevice);
   unsigned hw[NL];
   for (int mode = 0; mode < 2; ++mode) {
      // mode 0: blocks of 5 x 5 (the krob200 mean), mode 1: 12 x 12 (DOP with 12 headings)
      const int m = mode == 0 ? 5 : 12;
      for (int l = 0; l < NL; ++l) hw[l] = (unsigned)((l * 577) % 30000) | ((unsigned)m << 20) | ((unsigned)m << 26);
      cudaMemcpy(gw, hw, sizeof(hw), cudaMemcpyHostToDevice);
      printf("---- blocks %d x %d ----\n", m, m);
      run<0>("loop rows + REDUX (stored a layer late)", 32, gtab, gw, out, cyc);
      run<0>("loop rows + REDUX (stored a layer late)", 512, gtab, gw, out, cyc);
      run<1>("loop rows, no minimum", 512, gtab, gw, out, cyc);
      run<2>("loop rows, no minimum, no F store", 512, gtab, gw, out, cyc);
      run<3>("8 rows unrolled, no branch, no minimum", 512, gtab, gw, out, cyc);
      run<4>("loop rows + minimum by 5 shuffles", 512, gtab, gw, out, cyc);
   }
   return 0;
}
