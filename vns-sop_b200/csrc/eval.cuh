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
__host__ __device__ inline size_t eval_smem_bytes(int T, int C)
{
   return sizeof(int) * (size_t)(((T + 2 + 1) & ~1) + 2 * C + 4 * (T + 3));   // + rowoff (64-bit), cum, pd of the segment cache
}

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

// One lock-step evaluation of the warp's 32/G candidates (cc = this lane's group's candidate, K = 0:
// idle).  rowF(l, j) / rowB(l, j) return F[l][j] / B[l][j] as V.  Must be called by all 32 lanes.
// Out: fin = tryTo*(...) of the group's candidate (valid in every lane of the group), between =
// dist_between of its lower bound.
//
// Per round of G edges every lane first DESCRIBES one edge of its group's chain (block offset, block
// shape, dmin, and - by xor-shuffles across the groups - the warp-wide loop bounds); the edge loop
// then gets everything it needs from two broadcasts per edge.
struct EdgeDesc {
   long long rb;  // element offset of the block (row-major ma x mb) from Dv (64-bit: tables beyond 2^32 entries)
   int ma, mb;
   double dm;     // its contribution to dist_between of the lower bound
};

// chain = K blocks; edge(k), 1 <= k <= K, describes block k; the chain starts from the vector
// row0(j), j < m0, and closes against rowK(j), j < mK; tr = closing association (d + B) + v;
// ps, pe = the pair of edges the reference pre-sums in its lower bound (-1: none).
// LAT: latency mode for the single-CTA chains (nothing else hides the L1 latency there): every block
// step runs G fully unrolled iterations so that all its loads are in flight before the first min.
template <typename V, int G, int NR, bool LAT = false, typename Edge, typename Row0, typename RowK>
__device__ __forceinline__ void groups_eval(const V *__restrict__ Dv, int K, int tr, int ps, int pe, int m0, int mK, int gl, Edge edge,
                                            Row0 row0, RowK rowK, V &fin_out, double &between_out)
{
   typedef ValOps<V> Ops;
   constexpr bool EXACT = sizeof(V) == 4;   // integer tables: the lower bound's summation order is free
   constexpr unsigned FULL = 0xffffffffu;
   const V INF = Ops::inf();
   // chain start: v = F[s0], closing row B[sK]
   V v[NR], breg[NR];
#pragma unroll
   for (int r = 0; r < NR; ++r) {
      const int j = gl + r * G;
      v[r] = j < m0 ? row0(j) : INF;
      breg[r] = j < mK ? rowK(j) : INF;
   }
   double between = 0.0, pend = 0.0;   // ordered accumulation (double tables)
   double lane_dm = 0.0;               // per-lane partial sum (integer tables)
   V fin = INF;
   const int Kmax = __reduce_max_sync(FULL, K);
   for (int e0 = 0; e0 < Kmax; e0 += G) {
      // ---- describe edge k = e0 + gl + 1 ------------------------------------------------------
      long long e_rb = 0;
      unsigned e_w;
      double e_dm = 0.0;
      {
         const int k = e0 + gl + 1;
         int ma = 1, mb = 1, trn = 0;
         if (k <= K) {
            const EdgeDesc ed = edge(k);
            e_rb = ed.rb; ma = ed.ma; mb = ed.mb; e_dm = ed.dm;
            trn = (k == K) & tr;
         }
         int tmax = trn ? mb : ma, nvmax = trn ? ma : mb;       // warp-wide loop bounds of this edge slot
#pragma unroll
         for (int s = G; s < 32; s <<= 1) {
            tmax = max(tmax, __shfl_xor_sync(FULL, tmax, s));
            nvmax = max(nvmax, __shfl_xor_sync(FULL, nvmax, s));
         }
         e_w = (unsigned)ma | ((unsigned)mb << 8) | ((unsigned)(tmax | (trn << 6)) << 16) | ((unsigned)nvmax << 24);
         if (EXACT) lane_dm += e_dm;
      }
      const int kend = min(G, Kmax - e0);
      for (int kk = 0; kk < kend; ++kk) {
         const int k = e0 + kk + 1;
         const V *blk = Dv + __shfl_sync(FULL, e_rb, kk, G);
         const unsigned w = __shfl_sync(FULL, e_w, kk, G);
         const int ma = w & 0xff, mb = (w >> 8) & 0xff, tmaxAll = (w >> 16) & 0x3f, nvmax = w >> 24;
         const bool trn = (w >> 22) & 1;
         if (!EXACT) {   // lower bound in the source order of tryTo*LowerBound (path.cuh plan_lower_bound)
            const double dm = __shfl_sync(FULL, e_dm, kk, G);
            const bool is_ps = k == ps, is_pe = k == pe;
            const double add = is_pe ? pend + dm : dm;
            pend = is_ps ? dm : pend;
            between = (k <= K && !is_ps) ? between + add : between;
         }
         // Loads are never predicated: idle groups, exhausted groups and lanes beyond the block width
         // re-read a valid entry of their block and pair it with a value that cannot win (min is
         // idempotent), so the loop body is branch-free and the loads pipeline.
         const int trip = trn ? mb : ma, nvalid = trn ? ma : mb;
         const int sl = trn ? 1 : mb, sn = trn ? mb : 1;
         V acc[NR];
#pragma unroll
         for (int rd = 0; rd < NR; ++rd) {
            acc[rd] = INF;
            if (rd == 0 || nvmax > rd * G) {                 // warp-uniform
               const unsigned cbase = (unsigned)(min(gl + rd * G, nvalid - 1) * sn);
#pragma unroll
               for (int rs = 0; rs < NR; ++rs) {
                  if (rs == 0 || tmaxAll > rs * G) {         // warp-uniform
                     const int tlim = max(min(trip - rs * G, G) - 1, 0);
                     const unsigned sb = cbase + (unsigned)(min(rs * G, trip - 1) * sl);
                     const int tm = min(tmaxAll - rs * G, G);
                     const V opnd = trn ? breg[rs] : v[rs];
                     if (LAT) {
                        V dd[G];
#pragma unroll
                        for (int t = 0; t < G; ++t) dd[t] = __ldg(blk + (sb + (unsigned)(min(t, tlim) * sl)));
#pragma unroll
                        for (int t = 0; t < G; ++t) acc[rd] = Ops::add_min(acc[rd], __shfl_sync(FULL, opnd, min(t, tlim), G), dd[t]);
                     } else {
#pragma unroll 2
                        for (int t = 0; t < tm; ++t) {
                           const int tt = min(t, tlim);
                           const V x = __shfl_sync(FULL, opnd, tt, G);
                           const V dd = __ldg(blk + (sb + (unsigned)(tt * sl)));
                           acc[rd] = Ops::add_min(acc[rd], x, dd);
                        }
                     }
                  }
               }
            }
         }
         const bool isK = k == K, mid = k < K;
#pragma unroll
         for (int r = 0; r < NR; ++r) {
            const bool lv = gl + r * G < nvalid;
            const V f = trn ? acc[r] + v[r] : acc[r] + breg[r];
            fin = (isK && lv) ? Ops::vmin(fin, f) : fin;
            v[r] = mid ? (lv ? acc[r] : INF) : v[r];
         }
      }
   }
   if (EXACT) {
#pragma unroll
      for (int s = G / 2; s > 0; s >>= 1) lane_dm += __shfl_xor_sync(FULL, lane_dm, s, G);
      between = lane_dm;
   }
#pragma unroll
   for (int s = G / 2; s > 0; s >>= 1) fin = Ops::vmin(fin, __shfl_xor_sync(FULL, fin, s, G));
   fin_out = fin;
   between_out = between;
}

// ---- exact segment products (integer tables only) -------------------------------------------------
// MOVE and EXCHANGE chains contain a run of consecutive path layers a -> a+1 -> ... -> b.  Its min-plus
// product M(a,b) = D(a,a+1) (x) ... (x) D(b-1,b) does not depend on the candidate, and with integer
// entries min-plus products are exactly associative, so one pre-pass per path state builds every
// M(a,b), b >= a+2 (k_seg_build) and a chain of `span` blocks shrinks to at most 5 blocks with
// bit-identical results.  (The double path never uses this: floating-point sums do not reassociate.)
// M(a,b) is row-major m_a x m_b at element offset rowoff[a] + m_a * (cum[b] - cum[a+2]) from the table.
struct SegCache {
   const int *cum;          // cum[l] = samples in layers < l, l = 0 .. S+1
   const long long *rowoff; // per layer a
   const int *pd;           // pd[l] = sum_{l' < l} dmin(layer l', layer l'+1): dist_between of a run
};

struct SegCode {
   int K, n0, n1, n2, n3, n4, n5, kseg, foreign, tr, nolb;
};

__device__ __forceinline__ SegCode decode_cand_seg(int T, int C, int op, int a, int b)
{
   SegCode c;
   c.K = 0; c.n0 = c.n1 = c.n2 = c.n3 = c.n4 = c.n5 = 0; c.kseg = -1; c.foreign = 0; c.tr = 0; c.nolb = 0;
   if (op == VNSGPU_OP_ADD) {
      if (b < 0 || b > T || a < 0 || a >= C) return c;
      c.K = 2; c.n0 = b; c.n1 = -1; c.n2 = b + 1; c.foreign = a; c.tr = 1;
   } else if (op == VNSGPU_OP_REPLACE) {
      if (b < 0 || b >= T || a < 0 || a >= C) return c;
      c.K = 2; c.n0 = b; c.n1 = -1; c.n2 = b + 2; c.foreign = a; c.tr = 1;
   } else if (op == VNSGPU_OP_REMOVE) {
      if (a < 0 || a >= T) return c;
      c.K = 1; c.n0 = a; c.n1 = a + 2; c.tr = 1; c.nolb = 1;
   } else if (op == VNSGPU_OP_MOVE) {
      if (a >= T || a < 0 || b > T || b < 0 || abs(b - a) < 2) return c;
      const int i = min(a, b) + 1, j = max(a, b) + 1;
      if (b > a) {   // up: i-1, [i+1 .. j-1], i, j
         if (j - i == 2) { c.K = 3; c.n0 = i - 1; c.n1 = i + 1; c.n2 = i; c.n3 = j; }
         else { c.K = 4; c.n0 = i - 1; c.n1 = i + 1; c.n2 = j - 1; c.n3 = i; c.n4 = j; c.kseg = 2; }
      } else {       // down: i-1, j, [i .. j-1], j+1
         c.K = 4; c.n0 = i - 1; c.n1 = j; c.n2 = i; c.n3 = j - 1; c.n4 = j + 1; c.kseg = 3;
      }
   } else if (op == VNSGPU_OP_EXCHANGE) {   // i, j, [i+2 .. j-1], i+1, j+1
      if (a >= T || a < 0 || b >= T || b < 0 || abs(a - b) < 2) return c;
      const int i = min(a, b), j = max(a, b) + 1;
      if (j - i == 3) { c.K = 4; c.n0 = i; c.n1 = j; c.n2 = i + 2; c.n3 = i + 1; c.n4 = j + 1; }
      else { c.K = 5; c.n0 = i; c.n1 = j; c.n2 = i + 2; c.n3 = j - 1; c.n4 = i + 1; c.n5 = j + 1; c.kseg = 3; }
   }
   return c;
}

__device__ __forceinline__ int seg_node(const SegCode &c, int k)
{
   int l = c.n0;
   l = k == 1 ? c.n1 : l;
   l = k == 2 ? c.n2 : l;
   l = k == 3 ? c.n3 : l;
   l = k == 4 ? c.n4 : l;
   l = k == 5 ? c.n5 : l;
   return l;
}

// G lanes x NR registers hold one vector (sample j lives in lane j % G, register j / G); G * NR >= mp.
// perm (may be null): evaluation order = candidates sorted by chain length, so the 32/G candidates a
// warp advances in lock step have (almost always) the same number of blocks.
// SEG: chains compressed with the segment products (seg valid; V = int only).
template <typename V, int G, int NR, bool SEG>
__global__ void __launch_bounds__(VNS_EVAL_THREADS)
k_eval_groups(InstDev I, const V *__restrict__ Dv, PathDev P, SegCache seg, long long n, long long idx0,
              const vnsgpu_cand *__restrict__ cands, const unsigned *__restrict__ perm, double budget, int skipflag,
              double *__restrict__ len, double *__restrict__ lb, uint8_t *__restrict__ flags, unsigned long long *d_first)
{
   typedef ValOps<V> Ops;
   extern __shared__ int smem_i[];
   const int T = P.hdr[HDR_T];
   int *lcl = smem_i;
   int2 *cinfo = reinterpret_cast<int2 *>(smem_i + ((T + 2 + 1) & ~1));
   for (int l = threadIdx.x; l < T + 2; l += blockDim.x) lcl[l] = layer_cluster(I, P.tg, T, l);
   for (int c = threadIdx.x; c < I.C; c += blockDim.x) cinfo[c] = make_int2(I.off[c], I.m[c]);
   long long *s_rowoff = reinterpret_cast<long long *>(cinfo + I.C);
   int *s_cum = reinterpret_cast<int *>(s_rowoff + (T + 2)), *s_pd = s_cum + (T + 3);
   if (SEG) {
      for (int l = threadIdx.x; l < T + 3; l += blockDim.x) s_cum[l] = seg.cum[l];
      for (int l = threadIdx.x; l < T + 2; l += blockDim.x) { s_pd[l] = seg.pd[l]; s_rowoff[l] = seg.rowoff[l]; }
   }
   __syncthreads();

   constexpr int NG = 32 / G;
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
      if (cd.w >= I.C) cd.x = -1;                 // gain cluster out of range: a bad descriptor like any other bad index
      int K, s0, sK, nolb;                        // K = 0: out of range -> flags BAD
      V fin;
      double between;
      auto table_edge = [&](int ca, int cb) {
         const int2 cia = cinfo[ca], cib = cinfo[cb];
         EdgeDesc e;
         e.rb = (long long)cia.x * I.N + (long long)cia.y * cib.x;
         e.ma = cia.y; e.mb = cib.y;
         e.dm = __ldg(I.dmin + (size_t)ca * I.C + cb);
         return e;
      };
      if (SEG) {
         const SegCode sc = decode_cand_seg(T, I.C, cd.x, cd.y, cd.z);
         K = sc.K; s0 = sc.n0; sK = seg_node(sc, sc.K); nolb = sc.nolb;
         const int m0 = K ? cinfo[lcl[s0]].y : 0, mK = K ? cinfo[lcl[sK]].y : 0;
         groups_eval<V, G, NR>(Dv, K, sc.tr, -1, -1, m0, mK, gl,
            [&](int k) {
               const int la = seg_node(sc, k - 1), lb_ = seg_node(sc, k);
               if (k == sc.kseg && lb_ > la + 1) {          // run of path layers la .. lb_: one product
                  EdgeDesc e;
                  e.ma = s_cum[la + 1] - s_cum[la]; e.mb = s_cum[lb_ + 1] - s_cum[lb_];
                  e.rb = s_rowoff[la] + (long long)e.ma * (s_cum[lb_] - s_cum[la + 2]);
                  e.dm = (double)(s_pd[lb_] - s_pd[la]);
                  return e;
               }
               return table_edge(la < 0 ? sc.foreign : lcl[la], lb_ < 0 ? sc.foreign : lcl[lb_]);
            },
            [&](int j) { return (V)P.F[(size_t)s0 * I.mp + j]; }, [&](int j) { return (V)P.B[(size_t)sK * I.mp + j]; }, fin, between);
      } else {
         const CandCode cc = decode_cand(T, I.C, cd.x, cd.y, cd.z);
         K = cc.K; s0 = cc.s0; sK = cc.sK; nolb = cc.nolb;
         const int m0 = K ? cinfo[lcl[s0]].y : 0, mK = K ? cinfo[lcl[sK]].y : 0;
         groups_eval<V, G, NR>(Dv, K, cc.tr, cc.ps, cc.pe, m0, mK, gl,
            [&](int k) {
               const int la = cand_layer(cc, k - 1), lb_ = cand_layer(cc, k);
               return table_edge(la < 0 ? cc.foreign : lcl[la], lb_ < 0 ? cc.foreign : lcl[lb_]);
            },
            [&](int j) { return (V)P.F[(size_t)s0 * I.mp + j]; }, [&](int j) { return (V)P.B[(size_t)sK * I.mp + j]; }, fin, between);
      }
      if (gl == 0 && slot < n) {
         double l, bound;
         unsigned f = 0;
         if (K == 0) {
            l = CUDART_NAN; bound = CUDART_NAN; f = VNSGPU_F_BAD;
         } else {
            l = (double)fin;
            bound = nolb ? l : P.minF[s0] + between + P.minB[sK];
            double dR = 0.0, lm = l;                               // lm = lengthAfterMove of the drivers
            if (cd.x == VNSGPU_OP_ADD) {                           // insertRandom, outside -> path
               dR = cd.w >= 0 ? __ldg(I.reward + cd.w) : 0.0;      // vns_sop_solver.cpp:581
               lm = (L - 0.0) + (l - L);                           // :580,:583 (lengthRemoved = L - L)
            } else if (cd.x == VNSGPU_OP_REPLACE) {                // exchangeRandom, outside replaces b
               dR = cd.w >= 0 ? 0.0 + (__ldg(I.reward + cd.w) - __ldg(I.reward + lcl[cd.z + 1])) : 0.0;   // :784
            }
            if (lm <= budget) f |= VNSGPU_F_FEASIBLE;
            if (lm <= budget && (dR > 0 || (L - lm) > VNS_MIN_CHANGE_EPS)) f |= VNSGPU_F_ACCEPT;          // :619,:805
            if (bound > budget && !nolb) f |= VNSGPU_F_LB_SKIP;                                      // :539,...
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

// sort key of the endpoint-vector kernel: candidates with a block step first.  byrow (tables beyond the caches):
// block candidates are ordered by the first layer la of their untouched run - the products M(la, .) are one contiguous
// row of at most (S - la) blocks, so the candidates a wave of CTAs works on share a few L2-resident rows instead of
// gathering 4 KB blocks from all over HBM (ncu, 33 GB table: 5.8 GB of DRAM reads per 4 Mi candidates before)
__device__ __forceinline__ int vec_sort_key(int T, int C, const int4 &cd, bool byrow)
{
   const int op = cd.x, a = cd.y, b = cd.z;
   if (op == VNSGPU_OP_MOVE) {
      if (a >= T || a < 0 || b > T || b < 0 || abs(b - a) < 2) return 0;
      const int blk = b > a ? (b - a > 2 ? 1 : 0) : 1;     // up: layers i+1 .. j-1 = a+2 .. b;  down: always >= one block
      return byrow && blk ? 1 + min(a, b) + (b > a ? 2 : 1) : blk;
   }
   if (op == VNSGPU_OP_EXCHANGE) {
      if (a >= T || a < 0 || b >= T || b < 0 || abs(a - b) < 2) return 0;
      const int blk = abs(a - b) > 2 ? 1 : 0;              // layers p+1 .. q-1
      return byrow && blk ? 1 + min(a, b) + 2 : blk;
   }
   return 0;
}

// useseg: 0 block by block, 1 segment products, 2 endpoint vectors (vec.cuh), 3 endpoint vectors ordered by product row
__device__ __forceinline__ int sort_key(int T, int C, const int4 &cd, int useseg)
{
   if (useseg >= 2) return vec_sort_key(T, C, cd, useseg == 3);
   return useseg ? decode_cand_seg(T, C, cd.x, cd.y, cd.z).K : decode_cand(T, C, cd.x, cd.y, cd.z).K;
}

// ---- counting sort of a batch by chain length (descending): three light passes over the descriptors ----
// hist / cursor: nb = rows + 1 counters each (chain lengths are < rows); perm: n indices.
static __global__ void __launch_bounds__(256) k_sort_hist(InstDev I, PathDev P, long long n, const vnsgpu_cand *__restrict__ cands, int nb,
                                                   unsigned *__restrict__ hist, int useseg)
{
   extern __shared__ unsigned smem_u[];
   for (int k = threadIdx.x; k < nb; k += blockDim.x) smem_u[k] = 0;
   __syncthreads();
   const int T = P.hdr[HDR_T];
   for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (long long)gridDim.x * blockDim.x) {
      const int4 cd = __ldg(reinterpret_cast<const int4 *>(cands) + q);
      atomicAdd(&smem_u[sort_key(T, I.C, cd, useseg)], 1u);
   }
   __syncthreads();
   for (int k = threadIdx.x; k < nb; k += blockDim.x)
      if (smem_u[k]) atomicAdd(&hist[k], smem_u[k]);
}

// exclusive scan of the histogram in DESCENDING key order (longest chains first); one block of 1024 threads,
// thread t owns the t-th chunk of keys counted from the top
static __global__ void __launch_bounds__(1024) k_sort_scan(int nb, const unsigned *__restrict__ hist, unsigned *__restrict__ cursor)
{
   __shared__ unsigned part[1024];
   const int chunk = (nb + 1023) / 1024, t = threadIdx.x;
   const int hi = nb - 1 - t * chunk, lo = max(hi - chunk + 1, 0);   // keys hi, hi-1, .., lo
   unsigned sum = 0;
   for (int k = hi; k >= lo; --k) sum += hist[k];
   part[t] = sum;
   __syncthreads();
   for (int d = 1; d < 1024; d <<= 1) {
      const unsigned add = t >= d ? part[t - d] : 0;
      __syncthreads();
      part[t] += add;
      __syncthreads();
   }
   unsigned pos = part[t] - sum;   // exclusive prefix of this chunk
   for (int k = hi; k >= lo; --k) { cursor[k] = pos; pos += hist[k]; }
}

static __global__ void __launch_bounds__(256) k_sort_scatter(InstDev I, PathDev P, long long n, const vnsgpu_cand *__restrict__ cands, int nb,
                                                      unsigned *__restrict__ cursor, unsigned *__restrict__ perm, int useseg)
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
      atomicAdd(&cnt[sort_key(T, I.C, cd, useseg)], 1u);
   }
   __syncthreads();
   for (int k = threadIdx.x; k < nb; k += blockDim.x) {
      if (cnt[k]) base[k] = atomicAdd(&cursor[k], cnt[k]);
      cnt[k] = 0;
   }
   __syncthreads();
   for (long long q = lo + threadIdx.x; q < hi; q += blockDim.x) {
      const int4 cd = __ldg(reinterpret_cast<const int4 *>(cands) + q);
      const int K = sort_key(T, I.C, cd, useseg);
      perm[base[K] + atomicAdd(&cnt[K], 1u)] = (unsigned)q;
   }
}

// ---- segment-product cache (see SegCache) ------------------------------------------------------------
// meta (ints): cum[R], pd[R], rowoff[R] as 64-bit (2R ints), lcl[R] (layer -> cluster), loff[R] (first sample of the
// layer's cluster) with R = I.rows + 3 (rowoff starts 2R ints in: 8-byte aligned).  One block.
// exclusive prefix sum over the 256 threads of the block (s_w: 8 cells); *total = the sum of all
static __device__ __forceinline__ long long block_scan_256(long long mine, long long *s_w, long long *total)
{
   const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
   long long x = mine;
#pragma unroll
   for (int d = 1; d < 32; d <<= 1) {
      const long long y = __shfl_up_sync(0xffffffffu, x, d);
      if (lane >= d) x += y;
   }
   __syncthreads();                                     // s_w may still be read from the scan before
   if (lane == 31) s_w[w] = x;
   __syncthreads();
   long long base = 0, all = 0;
#pragma unroll
   for (int i = 0; i < 8; ++i) { const long long v = s_w[i]; all += v; if (i < w) base += v; }
   *total = all;
   return base + x - mine;
}

// every thread owns a run of consecutive layers; two block-wide scans (sizes + lower-bound distances packed into one word,
// then the rows' places behind the table) - the one-thread form walked the layers twice: 59 us on a 794-layer tour
static __global__ void __launch_bounds__(256) k_seg_offsets(InstDev I, PathDev P, int *__restrict__ meta)
{
   extern __shared__ int smem_i[];
   __shared__ long long s_w[8];
   const int T = P.hdr[HDR_T], S = T + 1, R = I.rows + 3;
   int *m_l = smem_i, *dm_l = smem_i + R;
   for (int l = threadIdx.x; l <= S; l += blockDim.x) {
      const int c = layer_cluster(I, P.tg, T, l);
      m_l[l] = I.m[c];
      dm_l[l] = l < S ? (int)I.dmin[(size_t)c * I.C + layer_cluster(I, P.tg, T, l + 1)] : 0;
      meta[4 * R + l] = c;
      meta[5 * R + l] = I.off[c];
   }
   __syncthreads();
   int *cum = meta, *pd = meta + R;
   long long *rowoff = reinterpret_cast<long long *>(meta + 2 * R);
   const int chunk = (S + 1 + 255) / 256, l0 = min(S + 1, (int)threadIdx.x * chunk), l1 = min(S + 1, l0 + chunk);
   long long mine = 0, all = 0;
   for (int l = l0; l < l1; ++l) mine += ((long long)dm_l[l] << 32) + m_l[l];      // both sums stay below 2^31
   long long run = block_scan_256(mine, s_w, &all);
   const int ctot = (int)(all & 0xffffffffll);
   for (int l = l0; l < l1; ++l) {
      const int c = (int)(run & 0xffffffffll);
      cum[l] = c; pd[l] = (int)(run >> 32);
      run += ((long long)dm_l[l] << 32) + m_l[l];
      dm_l[l] = c;                                      // from here on: cum, for the second scan
   }
   if (threadIdx.x == 0) { cum[S + 1] = ctot; dm_l[S + 1] = ctot; }
   __syncthreads();
   mine = 0;
   for (int a = l0; a < l1; ++a) mine += (long long)m_l[a] * (ctot - dm_l[min(a + 2, S + 1)]);
   run = (long long)I.N * I.N + block_scan_256(mine, s_w, &all);               // the products follow the table
   for (int a = l0; a < l1; ++a) {
      rowoff[a] = run;
      run += (long long)m_l[a] * (ctot - dm_l[min(a + 2, S + 1)]);
   }
}

// block a builds M(a,b) for b = a+2 .. S by M(a,b) = M(a,b-1) (x) D(b-1,b), M(a,a+1) = D(a,a+1)
// (global-memory form: any cluster size; every step is a global round trip)
static __global__ void __launch_bounds__(128) k_seg_build_g(InstDev I, PathDev P, int *Dv, const int *meta)
{
   const int T = P.hdr[HDR_T], S = T + 1, R = I.rows + 3;
   const int a = blockIdx.x;
   if (a + 2 > S) return;
   const int *cum = meta;
   const long long *rowoff = reinterpret_cast<const long long *>(meta + 2 * R);
   const int ca = layer_cluster(I, P.tg, T, a);
   const int ma = I.m[ca];
   const int *prev = Dv + ((long long)I.off[ca] * I.N + (long long)ma * I.off[layer_cluster(I, P.tg, T, a + 1)]);
   int mk = I.m[layer_cluster(I, P.tg, T, a + 1)];
   for (int b = a + 2; b <= S; ++b) {
      const int ck = layer_cluster(I, P.tg, T, b - 1), cb = layer_cluster(I, P.tg, T, b);
      const int mb = I.m[cb];
      const int *d = Dv + ((long long)I.off[ck] * I.N + (long long)mk * I.off[cb]);
      int *out = Dv + (rowoff[a] + (long long)ma * (cum[b] - cum[a + 2]));
      for (int e = threadIdx.x; e < ma * mb; e += blockDim.x) {
         const int i = e / mb, j = e - i * mb;
         int acc = VNS_INT_INF;
         for (int k = 0; k < mk; ++k) acc = __viaddmin_s32(prev[i * mk + k], d[k * mb + j], acc);
         out[e] = acc;
      }
      __syncthreads();
      prev = out;
      mk = mb;
   }
}

// The same recurrence with the running product in shared memory and the table blocks of the next NST - 1 steps in
// flight (cp.async ring of NST).  Row i of M(a,b) depends on row i of M(a,b-1) only, so the rows are independent
// chains: warp w of the 8 owns rows w, w+8, w+16, w+24 (lane = column), keeps them in its own slice of shared memory
// (read back as broadcast 16-byte loads: four k per load) and needs no barrier for them; the one barrier per step
// publishes the staged table block.  The chain of block a has S - a - 1 dependent steps; block 0's chain is the
// kernel's duration, and a step cannot be shorter than (memory latency) / (NST - 1): on HBM-resident tables the
// ring of four of the first version (three steps ahead) held every step at ~1-2 us, hence 8 / 16 stages.
// Needs mp <= 32 and N < 65536 (group_kernel_ok).  dynamic shared memory: seg_build_smem(rows, mp, NST) bytes
#define VNS_SEG_THREADS 256
__host__ __device__ inline size_t seg_build_smem(int rows, int mp, int nst)
{
   return sizeof(int) * ((size_t)(((rows + 3) + 3) & ~3) + (size_t)nst * ((mp * mp + 3) & ~3) + (size_t)32 * 32);
}

// U steps per barrier (small clusters: a step is a few hundred cycles of work, the barrier + wait + fetch round costs as
// much): the U table blocks of a round have landed before its barrier, the rows stay warp-private in between.
// tile > 0 (two-level build, see k_seg_cross): block a stops at the first layer of the next tile, b <= (a / tile + 1) * tile
template <int NST, int U>
__global__ void __launch_bounds__(VNS_SEG_THREADS) k_seg_build(InstDev I, PathDev P, int *Dv, const int *meta, int tile)
{
   extern __shared__ __align__(16) int smem_i[];
   constexpr int INF = VNS_INT_INF;
   const int T = P.hdr[HDR_T], Sfull = T + 1, R = I.rows + 3, mp2 = (I.mp * I.mp + 3) & ~3;
   const int a = blockIdx.x;
   if (a + 2 > Sfull) return;
   const int S = tile > 0 ? min(Sfull, (a / tile + 1) * tile) : Sfull;     // last layer of this block's chain
   if (a + 2 > S) return;
   unsigned short *l_m = reinterpret_cast<unsigned short *>(smem_i), *l_off = l_m + R;   // per layer l >= a: cluster size, first sample
   int *dblk = smem_i + ((R + 3) & ~3);                      // ring of NST table blocks
   int *prow = dblk + NST * mp2;                             // 32 rows x 32 of the running product (INF beyond its columns)
   for (int l = a + threadIdx.x; l <= S; l += blockDim.x) {
      const int c = layer_cluster(I, P.tg, T, l);
      l_m[l] = (unsigned short)I.m[c]; l_off[l] = (unsigned short)I.off[c];
   }
   // ring and rows start as zeros: the product loop reads up to three rows past a staged block (below) and must find
   // numbers in [0, INF] there - zeros, older blocks, or the rows of the running product that follow the ring
   for (int e = threadIdx.x; e < NST * mp2 + 32 * 32; e += blockDim.x) dblk[e] = 0;
   __syncthreads();
   const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
   const int ma = l_m[a];
   long long ooff = reinterpret_cast<const long long *>(meta + 2 * R)[a];   // M(a,a+2) starts the row; M(a,b+1) follows M(a,b)
   // D(l, l+1) -> dst (nothing when l >= S); always one commit group so that the group count per step is fixed
   auto fetch = [&](int *dst, int l) {
      if (l < S) {
         const int *src = Dv + ((long long)l_off[l] * I.N + (long long)l_m[l] * l_off[l + 1]);
         const int n = l_m[l] * l_m[l + 1];
         const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
         if ((((unsigned)(size_t)src) & 15u) == 0 && (n & 3) == 0) {     // block-uniform: 16-byte copies
            for (int e = threadIdx.x * 4; e < n; e += VNS_SEG_THREADS * 4)
               asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d + 4u * e), "l"(src + e));
         } else {
            for (int e = threadIdx.x; e < n; e += VNS_SEG_THREADS)
               asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d + 4u * e), "l"(src + e));
         }
      }
      asm volatile("cp.async.commit_group;");
   };
   static_assert(NST >= 2 * U && NST % U == 0, "ring of NST blocks, U per round");
   for (int b = a + 2; b < a + 2 + NST - U; ++b) fetch(dblk + (b % NST) * mp2, b - 1);   // step b multiplies by D(b-1,b)
   {   // M(a,a+1) = D(a,a+1): this warp's rows
      const int m1 = l_m[a + 1];
      const int *src = Dv + ((long long)l_off[a] * I.N + (long long)ma * l_off[a + 1]);
#pragma unroll
      for (int r = 0; r < 4; ++r) {
         const int i = w + 8 * r;
         if (i < ma) prow[i * 32 + lane] = lane < m1 ? __ldg(src + i * m1 + lane) : INF;
      }
      __syncwarp();
   }
   for (int b0 = a + 2; b0 <= S; b0 += U) {
      // invariant: fetched through step b0 + NST - U - 1; the blocks of steps b0 .. b0 + U - 1 are the oldest U groups
      asm volatile("cp.async.wait_group %0;" ::"n"(NST - 2 * U));
      __syncthreads();
#pragma unroll
      for (int u = 0; u < U; ++u) fetch(dblk + ((b0 + NST - U + u) % NST) * mp2, b0 + NST - U + u - 1);   // slots of the previous round
#pragma unroll 1
      for (int b = b0; b < b0 + U && b <= S; ++b) {
      const int mk = l_m[b - 1], mb = l_m[b];
      const int *d = dblk + (b % NST) * mp2 + min(lane, mb - 1);
      int acc[4] = {INF, INF, INF, INF};
      for (int k4 = 0; k4 < mk; k4 += 4, d += 4 * mb) {
         // rows k >= mk (at most three, when mk is not a multiple of 4) are read UNCLAMPED: they pair whatever follows the
         // block in shared memory (see above) with the INF the product's row holds beyond column mk - never the minimum
         const int d0 = d[0], d1 = d[mb], d2 = d[2 * mb], d3 = d[3 * mb];
#pragma unroll
         for (int r = 0; r < 4; ++r) {
            if (w + 8 * r < ma) {                            // warp-uniform
               const int4 p = *reinterpret_cast<const int4 *>(prow + (w + 8 * r) * 32 + k4);   // INF beyond mk
               acc[r] = __viaddmin_s32(p.x, d0, acc[r]);
               acc[r] = __viaddmin_s32(p.y, d1, acc[r]);
               acc[r] = __viaddmin_s32(p.z, d2, acc[r]);
               acc[r] = __viaddmin_s32(p.w, d3, acc[r]);
            }
         }
      }
      __syncwarp();
      int *out = Dv + ooff;
#pragma unroll
      for (int r = 0; r < 4; ++r) {
         const int i = w + 8 * r;
         if (i < ma) {
            if (lane < mb) out[i * mb + lane] = acc[r];
            prow[i * 32 + lane] = lane < mb ? acc[r] : INF;
         }
      }
      __syncwarp();
      ooff += ma * mb;
      }
   }
   asm volatile("cp.async.wait_group 0;");
}


// ---- two-level build of the segment products (paths of a hundred layers and more) --------------------------------------
// k_seg_build's block a runs S - a DEPENDENT products (793 on the 20 k-node state: 722 us for 65 MB, ncu r2i).  With tiles
// of SEG_TILE layers (s_t = t SEG_TILE) every M(a,b) is at most one product of two products that are themselves wanted:
//   1. k_seg_build(tile)   M(a,b), b <= s_{t(a)+1}: chains of <= SEG_TILE steps, all a in parallel
//   2. k_seg_tiles         M(s_t1, s_t2) = M(s_t1, s_{t2-1}) (x) M(s_{t2-1}, s_t2): one chain of <= S / SEG_TILE steps per t1
//   3. k_seg_cross<0>      M(a, s_t2) = M(a, s_{t(a)+1}) (x) M(s_{t(a)+1}, s_t2), t2 >= t(a) + 2, a not a tile start
//   4. k_seg_cross<1>      M(a, b)   = M(a, s_{t(b)}) (x) M(s_{t(b)}, b), t(b) > t(a), b not a tile start, b > s_{t(a)+1}
// Critical path SEG_TILE + S / SEG_TILE + 2 products instead of S; steps 3 and 4 are independent products (one warp per
// (a, tile), the left operand staged in shared memory).  Exact: integer min-plus products are associative.
#define SEG_TILE 32

struct SegMeta {
   const int *cum, *loff;
   const long long *rowoff;
   int N;
   __device__ __forceinline__ int m(int l) const { return __ldg(cum + l + 1) - __ldg(cum + l); }
   // M(a,b), b > a: the table block when b == a + 1, else the product stored behind the table
   __device__ __forceinline__ long long off(int a, int b) const
   {
      if (b == a + 1) return (long long)__ldg(loff + a) * N + (long long)m(a) * __ldg(loff + b);
      return __ldg(rowoff + a) + (long long)m(a) * (__ldg(cum + b) - __ldg(cum + a + 2));
   }
};

static __device__ __forceinline__ SegMeta seg_meta(const InstDev &I, const int *meta)
{
   const int R = I.rows + 3;
   return SegMeta{meta, meta + 5 * R, reinterpret_cast<const long long *>(meta + 2 * R), I.N};
}

// A warp's passes over b = b0, b0 + bstep .. b1: 32 / G products C_q = A (x) B_q at once (q = lane / G), A (ma x mk) staged in
// shared memory with a pitch of 32 ints and INF beyond column mk, lane j = lane % G owns column j of B_q in KU registers (one
// coalesced load per row of B_q) and walks the rows of A with 16-byte broadcast loads: per row KU DPX add-mins and KU / 4
// shared loads.  The NEXT pass's columns are loaded before the current pass is reduced (the passes were one exposed L2 round
// trip each).
template <int G, int KU>
static __device__ __forceinline__ void seg_warp_passes(int *__restrict__ Dv, const SegMeta &sm, const int *__restrict__ As, int a, int x,
                                                       int b0, int b1, int bstep, int ma, int mk, int q, int j)
{
   constexpr int NB = 32 / G;
   struct Pass { bool on; int mb; long long out; };
   int breg[KU], nreg[KU];
   auto fetch = [&](int bb, int *r) {
      Pass p;
      const int b = bb + q * bstep;
      p.on = b <= b1 && q < NB;                        // G = 10: lanes 30, 31 idle
      p.mb = p.on ? sm.m(b) : 0;
      p.out = p.on ? sm.off(a, b) : 0;
      const int *Bm = Dv + (p.on ? sm.off(x, b) : 0);
#pragma unroll
      for (int t = 0; t < KU; ++t) r[t] = (p.on && t < mk && j < p.mb) ? __ldcg(Bm + t * p.mb + j) : VNS_INT_INF;
      return p;
   };
   Pass cur = fetch(b0, breg);
   for (int bb = b0; bb <= b1; bb += NB * bstep) {
      const int bn = bb + NB * bstep;
      Pass nxt = cur;
      if (bn <= b1) nxt = fetch(bn, nreg);
      // the row loop is the kernel (ncu r5f: 54 instructions per row, 12 of them add-mins): pointers stepped instead of
      // indexed, the store predicated in place instead of a divergent region
      int *op = Dv + cur.out + j;
      const int4 *ap = reinterpret_cast<const int4 *>(As);
      const int st_ok = cur.on && j < cur.mb;
      for (int i = 0; i < ma; ++i, op += cur.mb, ap += 8) {
         int u = VNS_INT_INF, v = VNS_INT_INF;
#pragma unroll
         for (int t = 0; t < KU; t += 4) {
            const int4 p = ap[t / 4];
            u = __viaddmin_s32(p.x, breg[t], u); v = __viaddmin_s32(p.y, breg[t + 1], v);
            u = __viaddmin_s32(p.z, breg[t + 2], u); v = __viaddmin_s32(p.w, breg[t + 3], v);
         }
         asm volatile("{ .reg .pred p; setp.ne.s32 p, %2, 0; @p st.global.b32 [%0], %1; }" ::"l"(op), "r"(min(u, v)), "r"(st_ok) : "memory");
      }
      if (bn <= b1) {
#pragma unroll
         for (int t = 0; t < KU; ++t) breg[t] = nreg[t];
      }
      cur = nxt;
   }
}

// step 2: block t1 (t1 = 0 .. ntiles - 2): M(s_t1, s_t2) for t2 = t1 + 2 ..  (t2 = t1 + 1 came out of step 1).
// The running product stays in shared memory (it is the next step's left operand) and the right operands - products of
// step 1, independent of the chain - arrive through a ring of SEG_TILES_RING cp.async stages, so a step is its add-mins and
// one barrier instead of two global round trips (55 -> 15 us on a 794-layer tour).  Dynamic shared memory:
// (2 + SEG_TILES_RING) mp^2 ints.
#define SEG_TILES_RING 6
static __global__ void __launch_bounds__(256) k_seg_tiles(InstDev I, PathDev P, int *Dv, const int *meta)
{
   extern __shared__ int smem_i[];
   const int T = P.hdr[HDR_T], S = T + 1, mp2 = I.mp * I.mp;
   const SegMeta sm = seg_meta(I, meta);
   const int a = blockIdx.x * SEG_TILE;
   if (a + 2 * SEG_TILE > S) return;
   const int ma = sm.m(a);
   int *cur = smem_i, *nxt = smem_i + mp2, *ring = smem_i + 2 * mp2;
   const int nsteps = (S - a) / SEG_TILE - 1;            // b = a + (k + 2) SEG_TILE, k = 0 .. nsteps - 1
   auto stage = [&](int k) {                              // right operand of step k: M(b - SEG_TILE, b)
      if (k < nsteps) {
         const int b = a + (k + 2) * SEG_TILE, x = b - SEG_TILE;
         const int *Bm = Dv + sm.off(x, b);
         int *dst = ring + (k % SEG_TILES_RING) * mp2;
         const int n = sm.m(x) * sm.m(b);
         for (int e = threadIdx.x; e < n; e += blockDim.x)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(dst + e)), "l"(Bm + e) : "memory");
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
   };
   for (int k = 0; k < SEG_TILES_RING - 1; ++k) stage(k);
   {
      const int x = a + SEG_TILE;
      const int *A = Dv + sm.off(a, x);
      for (int e = threadIdx.x; e < ma * sm.m(x); e += blockDim.x) cur[e] = __ldcg(A + e);
   }
   for (int k = 0; k < nsteps; ++k) {
      const int b = a + (k + 2) * SEG_TILE, x = b - SEG_TILE, mk = sm.m(x), mb = sm.m(b);
      int *out = Dv + sm.off(a, b);
      asm volatile("cp.async.wait_group %0;" ::"n"(SEG_TILES_RING - 2) : "memory");
      __syncthreads();                                   // step k's operand and the running product are in place
      stage(k + SEG_TILES_RING - 1);                     // into the slot step k - 1 read: every thread is past that step now
      const int *Bs = ring + (k % SEG_TILES_RING) * mp2;
      for (int e = threadIdx.x; e < ma * mb; e += blockDim.x) {
         const int i = e / mb, j = e - i * mb;
         int acc = VNS_INT_INF;
         for (int t = 0; t < mk; ++t) acc = __viaddmin_s32(cur[i * mk + t], Bs[t * mb + j], acc);
         nxt[e] = acc;
         out[e] = acc;
      }
      int *t_ = cur; cur = nxt; nxt = t_;
   }
   asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// steps 3 (MODE 0: one warp per a, b over the tile starts beyond the next one) and 4 (MODE 1: one warp per (a, tile tb),
// b over the tile's layers).  G = lanes per product (>= mp; 10 = three products per warp for clusters of 9 or 10 samples, where
// 16 lanes per product and 16 register rows left 61 % of the add-mins on padding), KU = register rows (>= mp, chosen by the host:
// one size per kernel keeps the double-buffered columns at 2 KU registers).  Static shared memory: 8 warps x 32 x 32 ints.
template <int MODE, int G, int KU>
__global__ void __launch_bounds__(256) k_seg_cross(InstDev I, PathDev P, int *Dv, const int *meta)
{
   __shared__ __align__(16) int s_A[8][32 * 32];
   const int T = P.hdr[HDR_T], S = T + 1;
   const SegMeta sm = seg_meta(I, meta);
   const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, q = lane / G, j = lane % G;
   int *As = s_A[wib];
   // the warps stride over the tasks: the host sizes the grid from T as it last read it, the result does not depend on it
   const int nt = (S - 1) / SEG_TILE + 1;                        // tiles with a layer b that is not a tile start
   // MODE 1: only the pairs (a, tb) with tb beyond a's tile are enumerated - tile row ta holds SEG_TILE (nt - 1 - ta) of them -
   // so every warp of a block has work (a rectangular enumeration left half of the resident warps idle)
   const long long ntasks = MODE == 0 ? S + 1 : (long long)(SEG_TILE / 2) * nt * (nt - 1), nwarps = (long long)gridDim.x * (blockDim.x >> 5);
   for (long long task = (long long)blockIdx.x * (blockDim.x >> 5) + wib; task < ntasks; task += nwarps) {
      int a, x, b0, b1, bstep;
      if (MODE == 0) {
         a = (int)task;
         if (a % SEG_TILE == 0) continue;                        // tile starts: step 2
         x = (a / SEG_TILE + 1) * SEG_TILE;
         b0 = x + SEG_TILE; b1 = S; bstep = SEG_TILE;
      } else {
         int ta = 0;
         long long base = 0;
         while (base + (long long)SEG_TILE * (nt - 1 - ta) <= task) { base += (long long)SEG_TILE * (nt - 1 - ta); ++ta; }
         const int r = nt - 1 - ta, idx = (int)(task - base);
         a = ta * SEG_TILE + idx / r;
         const int tb = ta + 1 + idx % r;
         x = tb * SEG_TILE;
         b0 = x + 1; b1 = min(S, x + SEG_TILE - 1); bstep = 1;
      }
      if (b0 > b1 || x > S || x < a + 1) continue;
      const int ma = sm.m(a), mk = sm.m(x);
      __syncwarp();                                              // the task before is done with the staged operand
      {
         const int *A = Dv + sm.off(a, x);
         for (int e = lane; e < ma * 32; e += 32) {
            const int i = e >> 5, t = e & 31;
            As[e] = t < mk ? __ldcg(A + i * mk + t) : VNS_INT_INF;
         }
         __syncwarp();
      }
      seg_warp_passes<G, KU>(Dv, sm, As, a, x, b0, b1, bstep, ma, mk, q, j);
   }
}
