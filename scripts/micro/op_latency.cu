// Micro-benchmark: dependent-issue latency of the instructions a single-warp min-plus chain is made of
// (one warp, everything in registers / shared memory).  nvcc -O3 -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cuda_runtime.h>
#define FULL 0xffffffffu
#define N 256

template <int VAR> __global__ void k_lat(int *out, long long *cyc, const int *in)
{
   __shared__ int sm[1024];
   for (int k = threadIdx.x; k < 1024; k += blockDim.x) sm[k] = (k * 37 + 11) & 1023;
   __syncthreads();
   int a = in[threadIdx.x & 31], b = in[32 + (threadIdx.x & 31)], c = in[64];
   const int lane = threadIdx.x & 31;
   const long long t0 = clock64();
   if (VAR == 0) {          // VIADDMNMX chain
#pragma unroll
      for (int k = 0; k < N; ++k) a = __viaddmin_s32(a, b, c + k);
   } else if (VAR == 1) {   // IADD + VIMNMX chain
#pragma unroll
      for (int k = 0; k < N; ++k) a = min(a + b, c - k);
   } else if (VAR == 2) {   // SHFL chain
#pragma unroll
      for (int k = 0; k < N; ++k) a = __shfl_sync(FULL, a, (lane + 1) & 31);
   } else if (VAR == 3) {   // LDS chain (pointer chase)
#pragma unroll
      for (int k = 0; k < N; ++k) a = sm[a & 1023];
   } else if (VAR == 4) {   // 8 independent SHFL + 8 VIADDMNMX in two chains, repeated (one DP layer without loads)
#pragma unroll
      for (int k = 0; k < N / 8; ++k) {
         int a0 = c, a1 = c;
#pragma unroll
         for (int t = 0; t < 8; t += 2) {
            a0 = __viaddmin_s32(__shfl_sync(FULL, a, t), b + t, a0);
            a1 = __viaddmin_s32(__shfl_sync(FULL, a, t + 1), b - t, a1);
         }
         a = min(a0, a1);
      }
   } else if (VAR == 5) {   // the same with add + min tree (no DPX)
#pragma unroll
      for (int k = 0; k < N / 8; ++k) {
         int s[8];
#pragma unroll
         for (int t = 0; t < 8; ++t) s[t] = __shfl_sync(FULL, a, t) + (b ^ t);
         a = min(min(min(s[0], s[1]), min(s[2], s[3])), min(min(s[4], s[5]), min(s[6], s[7])));
      }
   } else if (VAR == 6) {   // v through shared memory instead of shuffles: STS, then 2 x LDS.128 broadcast, add + min tree
#pragma unroll
      for (int k = 0; k < N / 8; ++k) {
         sm[lane] = a;
         __syncwarp();
         const int4 q0 = *reinterpret_cast<const int4 *>(sm), q1 = *reinterpret_cast<const int4 *>(sm + 4);
         __syncwarp();
         const int s0 = q0.x + b, s1 = q0.y + (b ^ 1), s2 = q0.z + (b ^ 2), s3 = q0.w + (b ^ 3), s4 = q1.x + (b ^ 4), s5 = q1.y + (b ^ 5), s6 = q1.z + (b ^ 6),
                   s7 = q1.w + (b ^ 7);
         a = min(min(min(s0, s1), min(s2, s3)), min(min(s4, s5), min(s6, s7)));
      }
   } else if (VAR == 7) {   // REDUX chain
#pragma unroll
      for (int k = 0; k < N; ++k) a = __reduce_min_sync(FULL, a + lane);
   } else if (VAR == 8) {   // ballot chain
#pragma unroll
      for (int k = 0; k < N; ++k) a = (int)__ballot_sync(FULL, a > lane + k) + b;
   }
   const long long t1 = clock64();
   out[threadIdx.x] = a;
   if (threadIdx.x == 0) cyc[0] = t1 - t0;
}

template <int VAR> void run(const char *name, int per, int *out, long long *cyc, const int *in)
{
   for (int it = 0; it < 3; ++it) { k_lat<VAR><<<1, 32>>>(out, cyc, in); cudaDeviceSynchronize(); }
   long long h = 0;
   cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
   printf("%-70s %.1f cycles per %s\n", name, (double)h / (N / per), per == 1 ? "op" : "layer of 8 rows");
}

int main()
{
   int *out, *in; long long *cyc;
   cudaMalloc(&out, 4096); cudaMalloc(&cyc, 64); cudaMalloc(&in, 4096);
   int h[128]; for (int k = 0; k < 128; ++k) h[k] = k * 3 + 1;
   cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice);
   run<0>("VIADDMNMX dependent chain", 1, out, cyc, in);
   run<1>("IADD + VIMNMX dependent chain", 1, out, cyc, in);
   run<2>("SHFL dependent chain", 1, out, cyc, in);
   run<3>("LDS dependent chain", 1, out, cyc, in);
   run<7>("REDUX.MIN dependent chain", 1, out, cyc, in);
   run<8>("VOTE (ballot) dependent chain", 1, out, cyc, in);
   run<4>("layer: 8 SHFL + 8 VIADDMNMX (two chains of 4)", 8, out, cyc, in);
   run<5>("layer: 8 SHFL + 8 IADD + VIMNMX tree", 8, out, cyc, in);
   run<6>("layer: STS + 2 LDS.128 broadcast + 8 IADD + VIMNMX tree", 8, out, cyc, in);
   return 0;
}
