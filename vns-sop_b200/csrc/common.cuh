// Device-side data model of the engine (see DESIGN.md "Data layout in HBM").
//
//  * samples are cluster-contiguous: cluster c owns samples off[c] .. off[c+1]-1 (m[c] of them)
//  * the distance table is BLOCK-MAJOR: the m[c1] x m[c2] block of ordered cluster pair (c1,c2) is
//    one contiguous row-major run starting at off[c1]*N + m[c1]*off[c2]  (closed form, no offset
//    table, no padding: exactly 8*N^2 bytes).  Replaces allDistances[c1][c2][i][j]
//    (sources/vns_sop_path.h:157, four dependent pointer loads per entry in the reference).
//  * a path keeps targets[], the solver's VNS vector, and the two layered DP tables
//    shortest / shortest_back (vns_sop_path.h:184,190) as rows of stride mp (= max m[c]) plus the
//    per-layer minima the lower bounds need.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define VNS_M 1.7976931348623157e308  // numeric_limits<double>::max(), vns_sop_path.cpp:10
#define VNS_MIN_CHANGE_EPS 0.00001    // vns_sop_solver.cpp:15
#define VNS_WARP 32
#define VNS_CHAIN_THREADS 512   // CTA size of the global-memory chain kernels (search.cuh)
#define VNS_CHAIN_WARPS (VNS_CHAIN_THREADS / 32)

struct InstDev {
   int C, N, mp, start, goal;
   int rows;        // row capacity of a path (2C+2: the shakes add a range before removing one)
   int exact_int;   // every table entry is an integer and all path sums stay < 2^53
   int symmetric;   // D(c1,c2)[i][j] == D(c2,c1)[j][i] bit for bit (Euclidean builds): a column strip is a row band
   const int *m, *off, *clof;  // clof[s] = cluster of sample s
   const double *reward;
   const double *D;
   const double *dmin;
};

// hdr: [0]=T [1]=nvns [2]=error flag   hd: [0]=length (min F[S]) [1]=length_rear (min B[0]) [2]=reward
struct PathDev {
   int *hdr;
   double *hd;
   int *tg;       // targets, capacity rows
   int *vns;      // VNS vector, capacity rows
   double *F, *B; // rows x mp
   double *minF, *minB;  // rows
};

enum { HDR_T = 0, HDR_NVNS = 1, HDR_ERR = 2, HDR_INTS = 8 };
enum { HD_LEN = 0, HD_LEN_REAR = 1, HD_REWARD = 2, HD_DBLS = 4 };

__host__ __device__ __forceinline__ long long blk_off(const InstDev &I, int c1, int c2, const int *m, const int *off)
{
   return (long long)off[c1] * I.N + (long long)m[c1] * off[c2];
}

__device__ __forceinline__ const double *blk_ptr(const InstDev &I, int c1, int c2)
{
   return I.D + ((long long)__ldg(I.off + c1) * I.N + (long long)__ldg(I.m + c1) * __ldg(I.off + c2));
}

// layer -> cluster (VnsSopPath::update vns_sop_path.cpp:1327-1336)
__device__ __forceinline__ int layer_cluster(const InstDev &I, const int *tg, int T, int l)
{
   return l == 0 ? I.start : (l == T + 1 ? I.goal : tg[l - 1]);
}

__device__ __forceinline__ double warp_min(double v)
{
#pragma unroll
   for (int s = 16; s > 0; s >>= 1) {
      double o = __shfl_xor_sync(0xffffffffu, v, s);
      v = o < v ? o : v;
   }
   return v;
}

// ---- glibc rand() (random_r TYPE_3: r[i] = r[i-3] + r[i-31], output >> 1) -----------------------
struct RngState {
   int32_t r[31];
   int32_t f, b;
   int32_t pad;
};

__host__ __device__ __forceinline__ int rng_next(RngState &g)
{
   uint32_t v = (uint32_t)g.r[g.f] + (uint32_t)g.r[g.b];
   g.r[g.f] = (int32_t)v;
   g.f = g.f + 1 == 31 ? 0 : g.f + 1;
   g.b = g.b + 1 == 31 ? 0 : g.b + 1;
   return (int)(v >> 1);
}

// randIntMinMax math_common.cpp:21-28: lo + (int)((rand() / (RAND_MAX + 1.0)) * (hi - lo + 1));
// with RAND_MAX = 2^31-1 the double expression equals ((uint64)r * n) >> 31 exactly (r*n < 2^53)
__host__ __device__ __forceinline__ int rng_int(RngState &g, int lo, int hi)
{
   uint64_t n = (uint64_t)(hi - lo + 1);
   return lo + (int)(((uint64_t)rng_next(g) * n) >> 31);
}

inline void rng_seed_host(RngState &g, uint32_t seed)
{
   int32_t word = seed == 0 ? 1 : (int32_t)seed;
   g.r[0] = word;
   for (int i = 1; i < 31; ++i) {
      long hi = word / 127773, lo = word % 127773;
      long w = 16807 * lo - 2836 * hi;
      if (w < 0) w += 2147483647;
      word = (int32_t)w;
      g.r[i] = word;
   }
   g.f = 3;
   g.b = 0;
   g.pad = 0;
   for (int i = 0; i < 310; ++i) rng_next(g);
}

// ---- plain structs shared by the host API units and the search kernels ---------------------------
struct LsStats {
   long long proposals, lb_skips, evals, accepts, rounds;
};
struct VnsParamsDev {
   int num_iterations, num_iterations_unimproved, skip, greedy;
   double budget, max_time_ms;
   int max_improvements;
   // multi-start race (vnsgpu_set_target_reward): the first run whose tour reaches target_reward raises *stop_flag and
   // every run leaves at its next shake; null = off (the reference's stopping rules only)
   double target_reward;
   int *stop_flag;
};
struct VnsResultDev {
   int num_iters, num_iters_last_impr, n_improvements, status;
   double reward, length, time_ms, time_last_impr_ms;
   long long proposals, evals, accepts;
};

