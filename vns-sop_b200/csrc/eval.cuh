// Batched candidate evaluation, the kernel bench.py times.
//
// One sub-warp GROUP of G lanes (G = 8 / 16 / 32, the smallest that holds the largest cluster) owns one
// candidate; a warp advances its 32/G candidates in lock step, one distance block per iteration:
//   forward step     w[j] = min_i (v[i] + D(ca,cb)[i][j])       lane = j, v[i] arrives by shuffle
//   transposed step  u[i] = min_j (D(ca,cb)[i][j] + B[j])       lane = i, B[j] arrives by shuffle
// The running vector v never leaves registers.  Every tryTo* is a chain of forward steps closed by
//   (v + d) + B  (MOVE / EXCHANGE, vns_sop_path.cpp:789,1066,1145): forward step, then + B[j], min over j
//   (d + B) + v  (ADD / REPLACE / REMOVE, :295,504,526):            transposed step, then + v[i], min over i
// Both are bit-identical to the reference's min over all (i,j) of the three-term sum because rounding
// is monotone: min_i fl(a_i + c) == fl((min_i a_i) + c).
//
// V = double: IEEE arithmetic in the reference's association (any table).
// V = int:    tables whose entries are integers with every path sum < 2^29 (all shipped datasets):
//             exact int32 arithmetic on a 4-byte copy of the table, one DPX instruction
//             (VIADDMNMX: min(a + b, c)) per table entry; results convert to double exactly.
#pragma once
#include <climits>

#include "common.cuh"
#include "path.cuh"

template <typename V> struct ValOps;
template <> struct ValOps<double> {
   static __device__ __forceinline__ double inf() { return VNS_M; }
   static __device__ __forceinline__ double add_min(double acc, double x, double d)
   {
      const double n = x + d;
      return n < acc ? n : acc;
   }
   static __device__ __forceinline__ double vmin(double a, double b) { return b < a ? b : a; }
};
#define VNS_INT_INF 0x3f000000   // INF + INF < 2^31
template <> struct ValOps<int> {
   static __device__ __forceinline__ int inf() { return VNS_INT_INF; }
   static __device__ __forceinline__ int add_min(int acc, int x, int d) { return __viaddmin_s32(x, d, acc); }
   static __device__ __forceinline__ int vmin(int a, int b) { return min(a, b); }
};

// decoded candidate: node k of the chain sits at layer
//   k == 0 ? s0 : k == K ? sK : k == k1 ? s1 : k == kq ? sq : base + k      (layer -1 = the foreign cluster)
struct CandCode {
   int K, s0, sK, k1, s1, kq, sq, base, foreign;
   int tr;       // closing association (d + B) + v: transposed last step
   int ps, pe;   // lower bound: edges ps and pe enter dist_between as one pre-summed term
   int nolb;     // tryToRemove has no lower bound
};

__device__ __forceinline__ CandCode decode_cand(int T, int C, int op, int a, int b)
{
   CandCode c;
   c.K = 0; c.s0 = 0; c.sK = 0; c.k1 = -1; c.s1 = 0; c.kq = -1; c.sq = 0; c.base = 0; c.foreign = 0;
   c.tr = 0; c.ps = -1; c.pe = -1; c.nolb = 0;
   if (op == VNSGPU_OP_ADD) {                 // tryToAdd(c = a, idx = b), idx in [0, T]  (:247)
      if (b < 0 || b > T || a < 0 || a >= C) return c;
      c.K = 2; c.s0 = b; c.k1 = 1; c.s1 = -1; c.sK = b + 1; c.foreign = a; c.tr = 1; c.ps = 1; c.pe = 2;
   } else if (op == VNSGPU_OP_REPLACE) {      // tryToReplace(c = a, idx = b, 1)
      if (b < 0 || b >= T || a < 0 || a >= C) return c;
      c.K = 2; c.s0 = b; c.k1 = 1; c.s1 = -1; c.sK = b + 2; c.foreign = a; c.tr = 1; c.ps = 1; c.pe = 2;
   } else if (op == VNSGPU_OP_REMOVE) {       // tryToRemove(idx = a)
      if (a < 0 || a >= T) return c;
      c.K = 1; c.s0 = a; c.sK = a + 2; c.tr = 1; c.nolb = 1;
   } else if (op == VNSGPU_OP_MOVE) {         // tryToMove(what = a, where = b)  (:938-945)
      if (a >= T || a < 0 || b > T || b < 0 || abs(b - a) < 2) return c;
      const int i = min(a, b) + 1, j = max(a, b) + 1;
      if (b > a) {   // up:   i-1, i+1 .. j-1, i, j
         c.K = j - i + 1; c.s0 = i - 1; c.base = i; c.kq = c.K - 1; c.sq = i; c.sK = j; c.ps = c.K - 1; c.pe = c.K;
      } else {       // down: i-1, j, i .. j-1, j+1
         c.K = j - i + 2; c.s0 = i - 1; c.k1 = 1; c.s1 = j; c.base = i - 2; c.sK = j + 1; c.ps = 1; c.pe = 2;
      }
   } else if (op == VNSGPU_OP_EXCHANGE) {     // tryToExchange(a, b)  (:637-657): i, j, i+2 .. j-1, i+1, j+1
      if (a >= T || a < 0 || b >= T || b < 0 || abs(a - b) < 2) return c;
      const int i = min(a, b), j = max(a, b) + 1;
      c.K = j - i + 1; c.s0 = i; c.k1 = 1; c.s1 = j; c.base = i; c.kq = c.K - 1; c.sq = i + 1; c.sK = j + 1;
   }
   return c;
}

#define VNS_EVAL_THREADS 256

// shared memory: layer -> cluster map of the path (T+2 ints, padded to even) + per-cluster {offset, size}
__host__ __device__ inline size_t eval_smem_bytes(int T, int C) { return sizeof(int) * (size_t)(((T + 2 + 1) & ~1) + 2 * C); }

// layer of chain node k (see CandCode)
__device__ __forceinline__ int cand_layer(const CandCode &cc, int k)
{
   int l = cc.base + k;
   l = k == cc.kq ? cc.sq : l;
   l = k == cc.k1 ? cc.s1 : l;
   l = k == cc.K ? cc.sK : l;
   l = k == 0 ? cc.s0 : l;
   return l;
}

// G lanes x NR registers hold one vector (sample j lives in lane j % G, register j / G); G * NR >= mp.
// perm (may be null): evaluation order = candidates sorted by chain length, so the 32/G candidates a
// warp advances in lock step have (almost always) the same number of blocks.
template <typename V, int G, int NR>
__global__ void __launch_bounds__(VNS_EVAL_THREADS)
k_eval_groups(InstDev I, const V *__restrict__ Dv, PathDev P, long long n, long long idx0, const vnsgpu_cand *__restrict__ cands,
              const unsigned *__restrict__ perm, double budget, int skipflag, double *__restrict__ len, double *__restrict__ lb,
              uint8_t *__restrict__ flags, unsigned long long *d_first)
{
   typedef ValOps<V> Ops;
   extern __shared__ int smem_i[];
   const int T = P.hdr[HDR_T];
   int *lcl = smem_i;
   int2 *cinfo = reinterpret_cast<int2 *>(smem_i + ((T + 2 + 1) & ~1));
   for (int l = threadIdx.x; l < T + 2; l += blockDim.x) lcl[l] = layer_cluster(I, P.tg, T, l);
   for (int c = threadIdx.x; c < I.C; c += blockDim.x) cinfo[c] = make_int2(I.off[c], I.m[c]);
   __syncthreads();

   constexpr int NG = 32 / G;
   constexpr unsigned FULL = 0xffffffffu;
   const int lane = threadIdx.x & 31, gl = lane % G, g = lane / G;
   const double L = P.hd[HD_LEN];
   const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
   const long long warp0 = (((long long)blockIdx.x * blockDim.x) >> 5) + (threadIdx.x >> 5);
   const V INF = Ops::inf();
   unsigned long long best = ULLONG_MAX;

   for (long long base = warp0 * NG; base < n; base += nwarps * NG) {
      const long long slot = base + g;
      long long kidx = slot;
      if (perm && slot < n) kidx = __ldg(perm + slot);
      int4 cd = make_int4(-1, 0, 0, -1);
      if (slot < n) cd = __ldg(reinterpret_cast<const int4 *>(cands) + kidx);
      const CandCode cc = decode_cand(T, I.C, cd.x, cd.y, cd.z);
      const int K = cc.K;                         // 0: out of range -> flags BAD
      // chain start: v = F[s0], closing row B[sK]
      V v[NR], breg[NR];
      {
         const int m0 = K ? cinfo[lcl[cc.s0]].y : 0, mK = K ? cinfo[lcl[cc.sK]].y : 0;
#pragma unroll
         for (int r = 0; r < NR; ++r) {
            const int j = gl + r * G;
            v[r] = j < m0 ? (V)P.F[(size_t)cc.s0 * I.mp + j] : INF;
            breg[r] = j < mK ? (V)P.B[(size_t)cc.sK * I.mp + j] : INF;
         }
      }
      double between = 0.0, pend = 0.0;
      V fin = INF;
      const int Kmax = __reduce_max_sync(FULL, K);
      for (int e0 = 0; e0 < Kmax; e0 += G) {
         // lane gl describes edge k = e0 + gl + 1 of its group's chain: block offset, sizes, dmin
         unsigned e_rb = 0, e_mm = 1 | (1 << 8);
         double e_dm = 0.0;
         {
            const int k = e0 + gl + 1;
            if (k <= K) {
               const int la = cand_layer(cc, k - 1), lb_ = cand_layer(cc, k);
               const int ca = la < 0 ? cc.foreign : lcl[la], cb = lb_ < 0 ? cc.foreign : lcl[lb_];
               const int2 cia = cinfo[ca], cib = cinfo[cb];
               e_rb = (unsigned)cia.x * (unsigned)I.N + (unsigned)cia.y * (unsigned)cib.x;
               e_mm = (unsigned)cia.y | ((unsigned)cib.y << 8);
               e_dm = __ldg(I.dmin + (size_t)ca * I.C + cb);
            }
         }
         const int kend = min(G, Kmax - e0);
         for (int kk = 0; kk < kend; ++kk) {
            const int k = e0 + kk + 1;
            const bool act = k <= K;
            const unsigned rb = __shfl_sync(FULL, e_rb, kk, G);
            const unsigned mm = __shfl_sync(FULL, e_mm, kk, G);
            const double dm = __shfl_sync(FULL, e_dm, kk, G);
            const int ma = mm & 0xff, mb = mm >> 8;
            // lower bound: dist_between in the source order of tryTo*LowerBound (path.cuh plan_lower_bound)
            {
               const bool is_ps = k == cc.ps, is_pe = k == cc.pe;
               const double add = is_pe ? pend + dm : dm;
               pend = is_ps ? dm : pend;
               between = (act && !is_ps) ? between + add : between;
            }
            const bool trn = k == K && cc.tr;
            // Loads are never predicated: idle groups, exhausted groups and lanes beyond the block width
            // re-read a valid entry of their block and pair it with a value that cannot win (min is
            // idempotent), so the loop body is branch-free and the loads pipeline.
            const int trip = trn ? mb : ma, nvalid = trn ? ma : mb;
            const int sl = trn ? 1 : mb, sn = trn ? mb : 1;
            const int nvmax = NR > 1 ? __reduce_max_sync(FULL, nvalid) : 0;
            unsigned cbase[NR];
            V acc[NR];
#pragma unroll
            for (int r = 0; r < NR; ++r) {
               cbase[r] = rb + (unsigned)(min(gl + r * G, nvalid - 1) * sn);
               acc[r] = INF;
            }
#pragma unroll
            for (int rs = 0; rs < NR; ++rs) {
               const int trip_r = min(max(trip - rs * G, 0), G);
               const int tmax = __reduce_max_sync(FULL, trip_r);
               const V opnd = trn ? breg[rs] : v[rs];
#pragma unroll 2
               for (int t = 0; t < tmax; ++t) {
                  const int tt = max(min(t, trip_r - 1), 0);
                  const V x = __shfl_sync(FULL, opnd, tt, G);
                  const unsigned so = (unsigned)(min(rs * G + tt, trip - 1) * sl);
#pragma unroll
                  for (int rd = 0; rd < NR; ++rd) {
                     if (rd == 0 || nvmax > rd * G) {           // warp-uniform
                        const V dd = __ldg(Dv + (cbase[rd] + so));
                        acc[rd] = Ops::add_min(acc[rd], x, dd);
                     }
                  }
               }
            }
            if (act) {
               if (k == K) {
#pragma unroll
                  for (int r = 0; r < NR; ++r) {
                     const V f = trn ? acc[r] + v[r] : acc[r] + breg[r];
                     if (gl + r * G < nvalid) fin = Ops::vmin(fin, f);
                  }
               } else {
#pragma unroll
                  for (int r = 0; r < NR; ++r) v[r] = gl + r * G < nvalid ? acc[r] : INF;
               }
            }
         }
      }
#pragma unroll
      for (int s = G / 2; s > 0; s >>= 1) fin = Ops::vmin(fin, __shfl_xor_sync(FULL, fin, s, G));
      if (gl == 0 && slot < n) {
         double l, bound;
         unsigned f = 0;
         if (K == 0) {
            l = CUDART_NAN; bound = CUDART_NAN; f = VNSGPU_F_BAD;
         } else {
            l = (double)fin;
            bound = cc.nolb ? l : P.minF[cc.s0] + between + P.minB[cc.sK];
            double dR = 0.0, lm = l;                               // lm = lengthAfterMove of the drivers
            if (cd.x == VNSGPU_OP_ADD) {                           // insertRandom, outside -> path
               dR = cd.w >= 0 ? __ldg(I.reward + cd.w) : 0.0;      // vns_sop_solver.cpp:581
               lm = (L - 0.0) + (l - L);                           // :580,:583 (lengthRemoved = L - L)
            } else if (cd.x == VNSGPU_OP_REPLACE) {                // exchangeRandom, outside replaces b
               dR = cd.w >= 0 ? 0.0 + (__ldg(I.reward + cd.w) - __ldg(I.reward + lcl[cd.z + 1])) : 0.0;   // :784
            }
            if (lm <= budget) f |= VNSGPU_F_FEASIBLE;
            if (lm <= budget && (dR > 0 || (L - lm) > VNS_MIN_CHANGE_EPS)) f |= VNSGPU_F_ACCEPT;          // :619,:805
            if (bound > budget && !cc.nolb) f |= VNSGPU_F_LB_SKIP;                                      // :539,...
         }
         if (len) len[kidx] = l;
         if (lb) lb[kidx] = bound;
         if (flags) flags[kidx] = (uint8_t)f;
         if ((f & VNSGPU_F_ACCEPT) && !(skipflag && (f & VNSGPU_F_LB_SKIP)) && (unsigned long long)(kidx + idx0) < best)
            best = (unsigned long long)(kidx + idx0);
      }
   }
   if (d_first && best != ULLONG_MAX) atomicMin(d_first, best);
}

// double -> int32 copy of an all-integer table (same block-major layout)
__global__ void __launch_bounds__(256) k_table_to_int(const double *__restrict__ D, int *__restrict__ Di, long long n)
{
   for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (long long)gridDim.x * blockDim.x) Di[q] = (int)D[q];
}

// ---- counting sort of a batch by chain length (descending): three light passes over the descriptors ----
// hist / cursor: nb = rows + 1 counters each (chain lengths are < rows); perm: n indices.
__global__ void __launch_bounds__(256) k_sort_hist(InstDev I, PathDev P, long long n, const vnsgpu_cand *__restrict__ cands, int nb,
                                                   unsigned *__restrict__ hist)
{
   extern __shared__ unsigned smem_u[];
   for (int k = threadIdx.x; k < nb; k += blockDim.x) smem_u[k] = 0;
   __syncthreads();
   const int T = P.hdr[HDR_T];
   for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (long long)gridDim.x * blockDim.x) {
      const int4 cd = __ldg(reinterpret_cast<const int4 *>(cands) + q);
      atomicAdd(&smem_u[decode_cand(T, I.C, cd.x, cd.y, cd.z).K], 1u);
   }
   __syncthreads();
   for (int k = threadIdx.x; k < nb; k += blockDim.x)
      if (smem_u[k]) atomicAdd(&hist[k], smem_u[k]);
}

__global__ void k_sort_scan(int nb, const unsigned *__restrict__ hist, unsigned *__restrict__ cursor)
{
   if (threadIdx.x == 0 && blockIdx.x == 0) {
      unsigned pos = 0;
      for (int k = nb - 1; k >= 0; --k) { cursor[k] = pos; pos += hist[k]; }
   }
}

__global__ void __launch_bounds__(256) k_sort_scatter(InstDev I, PathDev P, long long n, const vnsgpu_cand *__restrict__ cands, int nb,
                                                      unsigned *__restrict__ cursor, unsigned *__restrict__ perm)
{
   extern __shared__ unsigned smem_u[];
   unsigned *cnt = smem_u, *base = smem_u + nb;
   for (int k = threadIdx.x; k < nb; k += blockDim.x) cnt[k] = 0;
   __syncthreads();
   const int T = P.hdr[HDR_T];
   const long long per = (n + gridDim.x - 1) / gridDim.x;
   const long long lo = per * blockIdx.x, hi = min(n, lo + per);
   for (long long q = lo + threadIdx.x; q < hi; q += blockDim.x) {
      const int4 cd = __ldg(reinterpret_cast<const int4 *>(cands) + q);
      atomicAdd(&cnt[decode_cand(T, I.C, cd.x, cd.y, cd.z).K], 1u);
   }
   __syncthreads();
   for (int k = threadIdx.x; k < nb; k += blockDim.x) {
      if (cnt[k]) base[k] = atomicAdd(&cursor[k], cnt[k]);
      cnt[k] = 0;
   }
   __syncthreads();
   for (long long q = lo + threadIdx.x; q < hi; q += blockDim.x) {
      const int4 cd = __ldg(reinterpret_cast<const int4 *>(cands) + q);
      const int K = decode_cand(T, I.C, cd.x, cd.y, cd.z).K;
      perm[base[K] + atomicAdd(&cnt[K], 1u)] = (unsigned)q;
   }
}
