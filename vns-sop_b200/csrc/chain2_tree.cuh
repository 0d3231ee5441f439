// Product tree of the second search chain (exact int32 tables; included by chain2.cuh).
//
// Edge e of a path joins layers e and e+1; its matrix is the table block D(c_e, c_{e+1}).  Node (l, k), l >= 1, holds the
// min-plus product of the 2^l edges [k 2^l, (k+1) 2^l): a matrix of shape m(layer k 2^l) x m(layer (k+1) 2^l).  Level 0 is
// the table itself.  On integers min-plus products are exactly associative, so
//   * a run of consecutive path layers inside a MOVE / EXCHANGE chain (vns_sop_path.cpp:1009-1035, :1111-1136, :725-761)
//     is at most 2 log2(T) nodes instead of `span` table blocks - the dependent chain that bounds a round on long tours;
//   * after an accepted move only the nodes over the changed edges are rebuilt (log2(T) levels, one warp per node),
//     and every layer's F / B row (updateAfterInsert / updateAfterRemove, :1426-1539) is recomputed independently from
//     the row at the edge of the changed range through at most 2 log2(T) nodes - all layers in parallel, where the
//     sequential re-scan is T dependent steps of one warp.
// Only full nodes exist ((k+1) 2^l <= S); a decomposition never asks for another.  Each thread block owns a replica of
// the tree in global memory (L1 / L2 resident; the same SM wrote it), node pitch mp * mp ints.
#pragma once
#include "common.cuh"

#define C2_MAXLEV 14

struct C2Tree {
   int *nodes;               // this block's replica; null: no tree (every run is walked edge by edge)
   int pitch;                // ints per node
   int nlev;                 // levels 1 .. nlev exist (0: none)
   int lvl_off[C2_MAXLEV + 2];
};

// levels for paths of at most `ecap` edges; returns the node count
__host__ __device__ inline int c2_tree_layout(int ecap, int *lvl_off, int *nlev)
{
   int off = 0, l = 1;
   for (; l <= C2_MAXLEV && (ecap >> l) > 0; ++l) { lvl_off[l] = off; off += ecap >> l; }
   *nlev = l - 1;
   return off;
}

struct C2Mat {
   const int *p;
   int ma, mb;
};

// cinfo: per cluster {first sample, size}; lcl: layer -> cluster
static __device__ __forceinline__ C2Mat c2_tree_mat(const C2Tree &tr, const int *Dv, int N, const int2 *cinfo, const int *lcl, int l, int k)
{
   C2Mat m;
   const int2 cia = cinfo[lcl[k << l]], cib = cinfo[lcl[(k + 1) << l]];
   m.ma = cia.y; m.mb = cib.y;
   m.p = l == 0 ? Dv + ((long long)cia.x * N + (long long)cia.y * cib.x) : tr.nodes + (size_t)(tr.lvl_off[l] + k) * tr.pitch;
   return m;
}

// idx-th node (in path order) of the greedy decomposition of the edge range [a, b): at every position the largest
// aligned node that fits.  At most 2 log2(b - a) nodes.  Returns the count when idx < 0.
static __device__ __forceinline__ int c2_decomp(int nlev, int a, int b, int idx, int *lev, int *node)
{
   if (nlev == 0) {
      if (idx >= 0) { *lev = 0; *node = a + idx; }
      return b - a;
   }
   int pos = a, n = 0;
   while (pos < b) {
      int l = pos ? __ffs(pos) - 1 : 30;
      l = min(min(l, 31 - __clz(b - pos)), nlev);
      if (n == idx) { *lev = l; *node = pos >> l; return n; }
      pos += 1 << l;
      ++n;
   }
   return n;
}

// nodes over the edge ranges [lo1, hi1) and [lo2, hi2) (either may be empty), level by level; all threads call
static __device__ void c2_tree_rebuild(const C2Tree &tr, const int *Dv, int N, const int2 *cinfo, const int *lcl, int S, int lo1, int hi1,
                                       int lo2, int hi2)
{
   const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
   __syncthreads();
   for (int l = 1; l <= tr.nlev && (1 << l) <= S; ++l) {
      const int nfull = S >> l;
      // node index ranges (inclusive), clipped to the full nodes; merged when they touch
      int a0 = lo1 >> l, a1 = min((hi1 - 1) >> l, nfull - 1), b0 = lo2 >> l, b1 = min((hi2 - 1) >> l, nfull - 1);
      if (hi1 <= lo1) { a0 = 0; a1 = -1; }
      if (hi2 <= lo2) { b0 = 0; b1 = -1; }
      if (a1 >= a0 && b1 >= b0 && b0 <= a1 + 1 && a0 <= b1 + 1) { a0 = min(a0, b0); a1 = max(a1, b1); b0 = 0; b1 = -1; }
      const int na = max(a1 - a0 + 1, 0), nb = max(b1 - b0 + 1, 0), ncount = na + nb;
      // wpn warps share a node (its elements are dealt out to them) when the level has fewer nodes than warps
      const int wpn = ncount > 0 ? max(1, nw / ncount) : 1, slots = nw / wpn;
      for (int q = warp / wpn; q < ncount; q += slots) {
         if (warp >= slots * wpn) break;
         const int k = q < na ? a0 + q : b0 + (q - na);
         const C2Mat L = c2_tree_mat(tr, Dv, N, cinfo, lcl, l - 1, 2 * k), R = c2_tree_mat(tr, Dv, N, cinfo, lcl, l - 1, 2 * k + 1);
         int *out = tr.nodes + (size_t)(tr.lvl_off[l] + k) * tr.pitch;
         const int ma = L.ma, mk = L.mb, mb = R.mb;
         for (int e = (warp % wpn) * 32 + lane; e < ma * mb; e += wpn * 32) {
            const int i = e / mb, j = e - i * mb;
            const int *lp = L.p + i * mk, *rp = R.p + j;
            int a = VNS_INT_INF, b = VNS_INT_INF;
            int t = 0;
            for (; t + 3 < mk; t += 4) {
               const int l0 = lp[t], l1 = lp[t + 1], l2 = lp[t + 2], l3 = lp[t + 3];
               const int r0 = rp[t * mb], r1 = rp[(t + 1) * mb], r2 = rp[(t + 2) * mb], r3 = rp[(t + 3) * mb];
               a = __viaddmin_s32(l0, r0, a); b = __viaddmin_s32(l1, r1, b);
               a = __viaddmin_s32(l2, r2, a); b = __viaddmin_s32(l3, r3, b);
            }
            for (; t < mk; ++t) a = __viaddmin_s32(lp[t], rp[t * mb], a);
            out[e] = min(a, b);
         }
      }
      __syncthreads();
   }
}

// ---- a chain of min-plus steps by a sub-warp group ------------------------------------------------------------------------
// The group holds a vector v (sample j in lane j % G, register j / G) and applies K blocks, block k described by
// desc(k) -> C2Mat.  TR = false: v <- v (x) M (lane = column of M);  TR = true: v <- M (x) v (lane = row of M).
// Block descriptions are produced one round of G steps ahead (each lane describes one block and prefetches it into L1).
// Returns the final vector in v[]; mlast = its size.  Must be called by all 32 lanes; K = 0: idle group (v unchanged).
template <int G, int NR, bool TR, typename Desc>
static __device__ __forceinline__ void c2_chain(int K, int gl, Desc desc, int (&v)[NR], int &mlast)
{
   constexpr unsigned FULL = 0xffffffffu;
   constexpr int INF = VNS_INT_INF;
   const int Kmax = __reduce_max_sync(FULL, K);
   unsigned long long n_p = 0;
   int n_ma = 1, n_mb = 1;
   auto describe = [&](int k) {
      n_p = 0; n_ma = 1; n_mb = 1;
      if (k <= K) {
         const C2Mat m = desc(k);
         n_p = (unsigned long long)m.p; n_ma = m.ma; n_mb = m.mb;
         const char *p = reinterpret_cast<const char *>(m.p);
         const int bytes = n_ma * n_mb * 4;
         for (int o = 0; o < bytes; o += 128) asm volatile("prefetch.global.L1 [%0];" ::"l"(p + o));
      }
   };
   describe(gl + 1);
   for (int e0 = 0; e0 < Kmax; e0 += G) {
      const unsigned long long e_p = n_p;
      int tmax = TR ? n_mb : n_ma, nvmax = TR ? n_ma : n_mb;     // trip count / output size of this block slot, warp-wide
#pragma unroll
      for (int sft = G; sft < 32; sft <<= 1) {
         tmax = max(tmax, __shfl_xor_sync(FULL, tmax, sft));
         nvmax = max(nvmax, __shfl_xor_sync(FULL, nvmax, sft));
      }
      const unsigned e_w = (unsigned)n_ma | ((unsigned)n_mb << 8) | ((unsigned)tmax << 16) | ((unsigned)nvmax << 24);
      if (e0 + G < Kmax) describe(e0 + G + gl + 1);
      const int kend = min(G, Kmax - e0);
      for (int kk = 0; kk < kend; ++kk) {
         const int k = e0 + kk + 1;
         const int *blk = reinterpret_cast<const int *>(__shfl_sync(FULL, e_p, kk, G));
         const unsigned w = __shfl_sync(FULL, e_w, kk, G);
         const int ma = w & 0xff, mb = (w >> 8) & 0xff, tmaxAll = (w >> 16) & 0xff, nvm = w >> 24;
         const bool on = k <= K;                                 // group-uniform; blk is null beyond the group's chain
         const int trip = TR ? mb : ma, nout = TR ? ma : mb;
         const int sl = TR ? 1 : mb, sn = TR ? mb : 1;           // strides: per trip step / per output lane
         int acc[NR];
#pragma unroll
         for (int rd = 0; rd < NR; ++rd) {
            acc[rd] = INF;
            if (rd == 0 || nvm > rd * G) {                       // warp-uniform
               const unsigned cbase = (unsigned)(min(gl + rd * G, nout - 1) * sn);
#pragma unroll
               for (int rs = 0; rs < NR; ++rs) {
                  if (rs == 0 || tmaxAll > rs * G) {              // warp-uniform
                     const int tlim = max(min(trip - rs * G, G) - 1, 0);
                     const unsigned sb = cbase + (unsigned)(min(rs * G, trip - 1) * sl);
                     const int opnd = v[rs];
                     int dd[G];
#pragma unroll
                     for (int t = 0; t < G; ++t) dd[t] = on ? blk[sb + (unsigned)(min(t, tlim) * sl)] : 0;
                     int a0 = INF, a1 = INF;
#pragma unroll
                     for (int t = 0; t < G; t += 2) {
                        a0 = __viaddmin_s32(__shfl_sync(FULL, opnd, min(t, tlim), G), dd[t], a0);
                        a1 = __viaddmin_s32(__shfl_sync(FULL, opnd, min(t + 1, tlim), G), dd[t + 1], a1);
                     }
                     acc[rd] = min(acc[rd], min(a0, a1));
                  }
               }
            }
         }
#pragma unroll
         for (int r = 0; r < NR; ++r) v[r] = on ? (gl + r * G < nout ? acc[r] : INF) : v[r];
         mlast = on ? nout : mlast;
      }
   }
}
