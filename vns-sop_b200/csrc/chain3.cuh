// Trajectory-exact search chain, third form: small instances (exact int32 tables).  The whole distance table sits in one
// SM's shared memory as int32 or, when every entry is below 65535, as uint16; a table that does not fit (below 96 clusters)
// is read from global memory through the read-only path by the same kernels (C3G<> below).
//
// insertRandom (vns_sop_solver.cpp:469-684) and exchangeRandom (:686-864) accept one proposal in 5 .. 15 on the shipped
// instances, so a round of a trajectory-exact chain is ONE accept: draw -> lower bound -> evaluate a dozen dependent
// block steps -> select -> shift + re-scan 20-40 layers.  Every one of those steps is a dependent memory round trip;
// with the table in L1 / L2 (first and second form) a round costs 13 000 - 20 000 cycles and the seeded krob200 solve
// loses to one CPU core (4.5 s against 3.2 s).  This form removes the round trips instead of adding parallelism:
//   * the distance table, the cluster-pair minima, the cluster table, the rewards and the whole search state live in
//     shared memory (krob200: 162 KB table + 20 KB state): a block step is LDS + SHFL + VIADDMNMX, no L1 / L2 access
//     in a round at all;
//   * every warp derives the round's window by itself - draw pairs from a shared ring of raw rand() values, validity,
//     the O(1) lower bound, the survivor masks (ballots) - so there is no cross-warp compaction, no thread-0 loop and
//     no barrier before the evaluation; every warp writes the same survivor list (pair indices in draw order) for its groups to pick from;
//   * the round's bookkeeping (proposals consumed, generator position, T, L, window size) is replicated in registers
//     of every thread; shared memory only carries the first accepted pair (one atomicMin cell per pass, four cells in
//     rotation: no reset barrier);
//   * the ring of raw generator outputs is topped up by the last warp WHILE the first three re-scan F / B / the prefix
//     sums after an accept, so the draws of the next round are already there;
//   * a round without accept costs one barrier per evaluation pass, a round with accept four.
// Same control flow, same draws, same acceptance tests as the other forms: tests/test_gpu_parity.py runs all of them
// against the reference's states.
#pragma once
#include "common.cuh"
#include "eval.cuh"
#include "search.cuh"

#define C3_THREADS 128                     // four warps: F re-scan, B re-scan, prefix sums, generator; every warp derives the round's window,
                                           // so more warps only add issue pressure (512 threads: krob200 solve 2.40 s, 256: 2.22 s, 128: 2.19 s)
#define C3_WARPS (C3_THREADS / 32)
#define C3_RING 1024                       // raw generator outputs kept (power of two)
#define C3_MAXCH 4                         // chunks of 32 draw pairs per round
#define C3_AHEAD (64 * C3_MAXCH)           // draws that must be in the ring at the start of a round
#define C3_INF VNS_INT_INF
#define C3_NONE 0x7fffffff

// Where the distance table is read from: TT = int / unsigned short -> a copy in shared memory; C3G<int> / C3G<unsigned short> ->
// the table in global memory through the read-only path (ld.global.nc: L1 hit latency is that of shared memory; instances
// whose table does not fit beside the state, e.g. 21 clusters x 16 headings = 226 KB as uint16).  Everything else is the same.
template <typename E> struct C3G {};
template <typename TT> struct C3TT { typedef TT elem; typedef unsigned addr; static constexpr bool global = false; };
template <typename E> struct C3TT<C3G<E>> { typedef E elem; typedef unsigned long long addr; static constexpr bool global = true; };

struct C3Ctl {
   int T, nvns, L, err;
   int bad;                  // a proposal decoded to no chain (impossible by construction)
   int ema;                  // running proposals per accept (window sizing)
   int first[4];             // first accepted pair of a pass (atomicMin), cells in rotation
   unsigned balv[C3_MAXCH], bals[C3_MAXCH];   // valid / surviving pairs of the round's window, per chunk (every warp writes the same)
   LsStats st;
   long long prof[4];
};

template <typename TT> struct C3Sm {
   const typename C3TT<TT>::elem *tab;   // N x N, block-major (shared-memory copy, or the global table)
   const int *dmin;          // C x C cluster-pair minima
   const double *reward;     // C
   const int2 *cinfo;        // C: {first sample, size}
   int *F, *Brev;            // R x mpp
   int *minF, *minBrev, *pd; // R
   int *lcl, *vns, *seg;     // 2R, 2R, alias of Brev (shake scratch)
   unsigned *ew;             // edge words of the path's layers (re-scans), R + 2
   int *surv;                // pair index of survivor k of the round's window (every warp writes the same values)
   unsigned *ring;           // C3_RING
   unsigned *gw;             // per evaluation group: the block words of its chain, gwcap each
   int gwcap;
   int N, C, mp;
};

// dynamic shared memory: table | dmin | reward | cinfo | F Brev minF minBrev pd lcl vns | ring
// the table is followed by 32 * mp + 32 zero entries: the row loops read whole groups of 4 / 8 / 16 / 32 rows (or row
// entries) without clamping; rows beyond a block pair an INF sample with whatever follows the block (table entries or
// this padding: all in [0, 2^29))
__host__ __device__ inline size_t c3_tab_bytes(int N, int mp, int esz) { return (((size_t)N * N + 32 * (size_t)mp + 32) * esz + 15) & ~(size_t)15; }
__host__ __device__ inline size_t c3_smem_bytes(int C, int N, int mp, int esz, bool table_in_global = false)
{
   const size_t R = C + 3, mpp = mp < 2 ? 2 : mp;
   size_t b = table_in_global ? 0 : c3_tab_bytes(N, mp, esz);
   b += ((size_t)C * C * sizeof(int) + 7) & ~(size_t)7;
   b += (size_t)C * sizeof(double) + (size_t)C * sizeof(int2);
   b += sizeof(int) * (2 * R * mpp + 8 * R + 8 + 32 * C3_MAXCH) + sizeof(unsigned) * C3_RING;
   b += sizeof(unsigned) * (size_t)(C3_THREADS / (mp <= 8 ? 8 : (mp <= 16 ? 16 : 32))) * (R + 5);   // block words of every group's chain
   return b + 16;
}

template <typename TT> static __device__ __forceinline__ C3Sm<TT> c3_carve(unsigned char *base, const InstDev &I, const void *tab_g)
{
   typedef typename C3TT<TT>::elem EL;
   const size_t R = I.C + 3, mpp = I.mp < 2 ? 2 : I.mp;
   C3Sm<TT> s;
   s.N = I.N; s.C = I.C; s.mp = I.mp;
   if (C3TT<TT>::global)
      s.tab = reinterpret_cast<const EL *>(tab_g);
   else {
      s.tab = reinterpret_cast<const EL *>(base);
      base += c3_tab_bytes(I.N, I.mp, sizeof(EL));
   }
   s.dmin = reinterpret_cast<const int *>(base);
   base += ((size_t)I.C * I.C * sizeof(int) + 7) & ~(size_t)7;
   s.reward = reinterpret_cast<const double *>(base);
   base += (size_t)I.C * sizeof(double);
   s.cinfo = reinterpret_cast<const int2 *>(base);
   base += (size_t)I.C * sizeof(int2);
   int *p = reinterpret_cast<int *>(base);
   s.F = p; p += R * mpp;
   s.Brev = p; p += R * mpp;
   s.minF = p; p += R;
   s.minBrev = p; p += R;
   s.pd = p; p += R;
   s.lcl = p; p += 2 * R;
   s.vns = p; p += 2 * R;
   s.surv = p; p += 32 * C3_MAXCH;
   s.ew = reinterpret_cast<unsigned *>(p) + 2; p += R + 8;      // indices -2 .. R + 5
   s.ring = reinterpret_cast<unsigned *>(p);
   s.gw = s.ring + C3_RING;
   s.gwcap = (int)R + 5;
   s.seg = s.Brev;
   return s;
}

struct C3Env {
   InstDev I;
   const void *tab_g;        // int32 or uint16 copy of the table in global memory
   const int *dmin_g;        // C x C
   const int2 *cinfo_g;      // C
};

// instance constants -> shared memory (once per kernel); all threads
template <typename TT> static __device__ void c3_load_tables(const C3Env &E, const C3Sm<TT> &s)
{
   typedef typename C3TT<TT>::elem EL;
   const InstDev &I = E.I;
   if (!C3TT<TT>::global) {  // table: 16-byte words (both copies are allocated with slack behind the last entry)
      const size_t n16 = ((size_t)I.N * I.N * sizeof(EL) + 15) / 16, all16 = c3_tab_bytes(I.N, I.mp, sizeof(EL)) / 16;
      const int4 *src = reinterpret_cast<const int4 *>(E.tab_g);
      int4 *dst = reinterpret_cast<int4 *>(const_cast<EL *>(s.tab));
      for (size_t k = threadIdx.x; k < n16; k += blockDim.x) dst[k] = __ldg(src + k);
      for (size_t k = n16 + threadIdx.x; k < all16; k += blockDim.x) dst[k] = make_int4(0, 0, 0, 0);
      __syncthreads();
      // the last 16-byte word of the copy may carry bytes from behind the table: make them zeros as well
      EL *tw = const_cast<EL *>(s.tab);
      for (size_t k = (size_t)I.N * I.N + threadIdx.x; k < n16 * (16 / sizeof(EL)); k += blockDim.x) tw[k] = 0;
   }
   int *dm = const_cast<int *>(s.dmin);
   for (int k = threadIdx.x; k < I.C * I.C; k += blockDim.x) dm[k] = __ldg(E.dmin_g + k);
   double *rw = const_cast<double *>(s.reward);
   int2 *ci = const_cast<int2 *>(s.cinfo);
   for (int c = threadIdx.x; c < I.C; c += blockDim.x) { rw[c] = __ldg(I.reward + c); ci[c] = __ldg(E.cinfo_g + c); }
   __syncthreads();
}

// ---- min-plus rows, unclamped ------------------------------------------------------------------------------------------------
// acc = min_t (v[t] + p[t * stride]) over rows 0 .. tmax-1 (tmax warp-uniform, rounded up to 4 or a multiple of 8); v[t]
// arrives by shuffle (width G); rows beyond the block meet an INF sample.  Per row: LDS, SHFL, VIADDMNMX - no clamp.
// Kept as LOOPS on purpose: one warp executes every instruction of a round once, so straight-line code runs at the
// speed of instruction fetch (L0 instruction cache ~6 KB, L1.5 32 KB); the first, fully unrolled version of these
// kernels took 500 cycles per DP layer.
template <typename TT, int G>
static __device__ __forceinline__ int c3_block(const TT *p, int stride, int v, int tmax)
{
   constexpr unsigned FULL = 0xffffffffu;
   int a0 = C3_INF, a1 = C3_INF;
   if (tmax <= 4) {
      const int d0 = (int)p[0], d1 = (int)p[stride], d2 = (int)p[2 * stride], d3 = (int)p[3 * stride];
      const int x0 = __shfl_sync(FULL, v, 0, G), x1 = __shfl_sync(FULL, v, 1, G), x2 = __shfl_sync(FULL, v, 2, G), x3 = __shfl_sync(FULL, v, 3, G);
      a0 = __viaddmin_s32(x0, d0, a0); a1 = __viaddmin_s32(x1, d1, a1);
      a0 = __viaddmin_s32(x2, d2, a0); a1 = __viaddmin_s32(x3, d3, a1);
   } else {
#pragma unroll 1
      for (int t0 = 0; t0 < tmax; t0 += 8) {
         int d[8], x[8];
#pragma unroll
         for (int t = 0; t < 8; ++t) d[t] = (int)p[t * stride];
#pragma unroll
         for (int t = 0; t < 8; ++t) x[t] = __shfl_sync(FULL, v, t0 + t, G);
#pragma unroll
         for (int t = 0; t < 8; t += 2) { a0 = __viaddmin_s32(x[t], d[t], a0); a1 = __viaddmin_s32(x[t + 1], d[t + 1], a1); }
         p += 8 * stride;
      }
   }
   return min(a0, a1);
}

// edge (l-1 -> l) of the path as one word: block offset | rows << 20 | columns << 26
template <typename TT> static __device__ __forceinline__ unsigned c3_edge_word(const C3Sm<TT> &s, int ca, int cb)
{
   const int2 cia = s.cinfo[ca], cib = s.cinfo[cb];
   return (unsigned)(cia.x * s.N + cia.y * cib.x) | ((unsigned)cia.y << 20) | ((unsigned)cib.y << 26);
}

// ---- DP layers on the shared-memory state, one warp (lane = sample; mp <= 32) -------------------------------------------
// F[l] = F[l-1] (x) D(c_{l-1}, c_l)   vns_sop_path.cpp:1353-1373 / 1426-1448 / 1488-1511
template <typename TT> static __device__ __forceinline__ void c3_fwd_layer(const C3Sm<TT> &s, int l, int lane)
{
   constexpr unsigned FULL = 0xffffffffu;
   const int mp = s.mp;
   const unsigned w = c3_edge_word(s, s.lcl[l - 1], s.lcl[l]);
   const int ma = (w >> 20) & 63, mb = w >> 26;
   const int v = lane < ma ? s.F[(l - 1) * mp + lane] : C3_INF;
   int acc = c3_block<typename C3TT<TT>::elem, 32>(s.tab + (w & 0xfffffu) + min(lane, mb - 1), mb, v, ma);
   acc = lane < mb ? acc : C3_INF;
   if (lane < mb) s.F[l * mp + lane] = acc;
   const int mn = __reduce_min_sync(FULL, acc);
   if (lane == 0) s.minF[l] = mn;
   __syncwarp();
}

// ---- re-scans: one warp, one dependent chain --------------------------------------------------------------------------------
// A layer is SHFL (27 cycles) -> VIADDMNMX x 4 (5 each) -> min -> select -> next layer's SHFL: ~65 cycles when nothing else
// sits on the chain (scripts/micro/op_latency.cu).  One warp issues in order, so every load, shuffle or reduction whose
// result is consumed right behind it stalls the whole chain for its full latency; the first versions of this loop took
// 420-500 cycles per layer (scripts/micro/c3_parts.cu).  Hence:
//   * the edge words (block offset | rows << 20 | columns << 26) of all layers are written to shared memory by the lanes in
//     parallel before the loop; the loop reads them two layers ahead;
//   * the first eight rows of the NEXT layer's block are loaded before the current layer is reduced (rows beyond a block
//     meet INF samples: no clamping); blocks with more than eight rows take a tail loop;
//   * the per-layer minima (the lower bounds read them) come from a lane-parallel pass behind the loop instead of a
//     REDUX per layer.
#define C3_EW_NONE ((1u << 20) | (1u << 26))     // offset 0, 1 x 1: a valid read that nobody uses

//   * the loop works on 32-bit shared-memory addresses held in registers (laundered through an empty asm) and explicit
//     ld.shared / st.shared in program order: left to itself the compiler re-derives every shared-memory pointer from
//     S2R / LDC in each iteration (35 instructions and two S2R round trips per layer) and parks each load right in
//     front of its consumer.
static __device__ __forceinline__ unsigned c3_saddr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
static __device__ __forceinline__ unsigned c3_keep(unsigned x) { asm volatile("" : "+r"(x)); return x; }
static __device__ __forceinline__ int c3_lds(unsigned a)
{
   int v;
   asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(a));
   return v;
}
static __device__ __forceinline__ void c3_sts(unsigned a, int v) { asm volatile("st.shared.s32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
static __device__ __forceinline__ unsigned long long c3_keep(unsigned long long x) { asm volatile("" : "+l"(x)); return x; }
template <typename TT> static __device__ __forceinline__ int c3_ldt(typename C3TT<TT>::addr a);
template <> __device__ __forceinline__ int c3_ldt<int>(unsigned a) { return c3_lds(a); }
template <> __device__ __forceinline__ int c3_ldt<unsigned short>(unsigned a)
{
   int v;
   asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(a));
   return v;
}
template <> __device__ __forceinline__ int c3_ldt<C3G<int>>(unsigned long long a)
{
   int v;
   asm volatile("ld.global.nc.s32 %0, [%1];" : "=r"(v) : "l"(a));
   return v;
}
template <> __device__ __forceinline__ int c3_ldt<C3G<unsigned short>>(unsigned long long a)
{
   int v;
   asm volatile("ld.global.nc.u16 %0, [%1];" : "=r"(v) : "l"(a));
   return v;
}
// address of the table as the loops use it (shared window offset, or the 64-bit global address), held in a register
template <typename TT> static __device__ __forceinline__ typename C3TT<TT>::addr c3_tab_addr(const void *tab)
{
   if (C3TT<TT>::global) return (typename C3TT<TT>::addr)c3_keep((unsigned long long)tab);
   return (typename C3TT<TT>::addr)c3_keep(c3_saddr(tab));
}

// rows t0 .. t0 + 7 of a block (unclamped) against the shuffled samples: the tail of blocks with more than eight rows
template <typename TT, int W>
static __device__ __forceinline__ void c3_rows8(typename C3TT<TT>::addr pa, unsigned strideB, int v, int t0, int &a0, int &a1)
{
   constexpr unsigned FULL = 0xffffffffu;
   int d[8], x[8];
#pragma unroll
   for (int t = 0; t < 8; ++t) d[t] = c3_ldt<TT>(pa + (unsigned)(t0 + t) * strideB);
#pragma unroll
   for (int t = 0; t < 8; ++t) x[t] = __shfl_sync(FULL, v, t0 + t, W);
#pragma unroll
   for (int t = 0; t < 8; t += 2) { a0 = __viaddmin_s32(x[t], d[t], a0); a1 = __viaddmin_s32(x[t + 1], d[t + 1], a1); }
}

template <typename TT> static __device__ void c3_fwd_scan(const C3Sm<TT> &s, int S, int lf, int lane)
{
   constexpr unsigned FULL = 0xffffffffu;
   typedef typename C3TT<TT>::addr AT;
   constexpr int ES = (int)sizeof(typename C3TT<TT>::elem);
   if (lf > S) return;
   const int mp = s.mp;
   for (int l = lf + lane; l <= S; l += 32) s.ew[l] = c3_edge_word(s, s.lcl[l - 1], s.lcl[l]);
   if (lane < 2) s.ew[S + 1 + lane] = C3_EW_NONE;
   __syncwarp();
   lane = (int)c3_keep((unsigned)lane);
   const AT tabA = c3_tab_addr<TT>(s.tab);
   const unsigned mp4 = c3_keep(4u * (unsigned)mp);
   unsigned ewp = c3_keep(c3_saddr(s.ew + lf)), frow = c3_keep(c3_saddr(s.F + lf * mp + lane));
   unsigned w = (unsigned)c3_lds(ewp), wn = (unsigned)c3_lds(ewp + 4);
   ewp += 8;
   int v = lane < (int)((w >> 20) & 63) ? c3_lds(frow - mp4) : C3_INF;
   AT pa = tabA + ES * ((w & 0xfffffu) + min(lane, (int)(w >> 26) - 1));
   int dd[8];
#pragma unroll
   for (int t = 0; t < 8; ++t) dd[t] = c3_ldt<TT>(pa + t * ES * (w >> 26));
#pragma unroll 2
   for (int l = lf; l <= S; ++l) {
      int x[8];
#pragma unroll
      for (int t = 0; t < 8; ++t) x[t] = __shfl_sync(FULL, v, t);
      const unsigned wn2 = (unsigned)c3_lds(ewp);
      ewp += 4;
      const unsigned sbn = ES * (wn >> 26);
      const AT pn = tabA + ES * ((wn & 0xfffffu) + min(lane, (int)(wn >> 26) - 1));
      int dn[8];
#pragma unroll
      for (int t = 0; t < 8; ++t) dn[t] = c3_ldt<TT>(pn + t * sbn);
      int a0 = __viaddmin_s32(x[0], dd[0], C3_INF), a1 = __viaddmin_s32(x[1], dd[1], C3_INF);
#pragma unroll
      for (int t = 2; t < 8; t += 2) { a0 = __viaddmin_s32(x[t], dd[t], a0); a1 = __viaddmin_s32(x[t + 1], dd[t + 1], a1); }
      const int ma = (w >> 20) & 63, mb = w >> 26;
      if (ma > 8) {
#pragma unroll 1
         for (int t0 = 8; t0 < ma; t0 += 8) c3_rows8<TT, 32>(pa, ES * mb, v, t0, a0, a1);
      }
      const int acc = lane < mb ? min(a0, a1) : C3_INF;
      if (lane < mb) c3_sts(frow, acc);
      frow += mp4;
      v = acc;
      w = wn; wn = wn2; pa = pn;
#pragma unroll
      for (int t = 0; t < 8; ++t) dd[t] = dn[t];
   }
   __syncwarp();
   for (int l = lf + lane; l <= S; l += 32) {
      const int m = (int)(s.ew[l] >> 26);
      int mn = C3_INF;
      for (int j = 0; j < m; ++j) mn = min(mn, s.F[l * mp + j]);
      s.minF[l] = mn;
   }
   __syncwarp();
}

// backward layers [lb, 0]: B[l][i] = min_j (B[l+1][j] + D(c_l, c_{l+1})[i][j])   :1380-1399 / 1453-1472 / 1515-1539
// rows are stored by distance from the goal (slot S - l), so inserting a target never moves a row.  ew[l + 1] = edge l -> l + 1
template <typename TT> static __device__ void c3_bwd_scan(const C3Sm<TT> &s, int S, int lb, int lane)
{
   constexpr unsigned FULL = 0xffffffffu;
   typedef typename C3TT<TT>::addr AT;
   constexpr int ES = (int)sizeof(typename C3TT<TT>::elem);
   if (lb < 0) return;
   const int mp = s.mp;
   for (int l = lb - lane; l >= 0; l -= 32) s.ew[l + 1] = c3_edge_word(s, s.lcl[l], s.lcl[l + 1]);
   if (lane < 2) s.ew[-lane] = C3_EW_NONE;
   __syncwarp();
   lane = (int)c3_keep((unsigned)lane);
   const AT tabA = c3_tab_addr<TT>(s.tab);
   const unsigned mp4 = c3_keep(4u * (unsigned)mp);
   unsigned ewp = c3_keep(c3_saddr(s.ew + lb + 1)), brow = c3_keep(c3_saddr(s.Brev + (S - lb) * mp + lane));
   unsigned w = (unsigned)c3_lds(ewp), wn = (unsigned)c3_lds(ewp - 4);
   ewp -= 8;
   int v = lane < (int)(w >> 26) ? c3_lds(brow - mp4) : C3_INF;
   AT pa = tabA + ES * ((w & 0xfffffu) + min(lane, (int)((w >> 20) & 63) - 1) * (w >> 26));
   int dd[8];
#pragma unroll
   for (int t = 0; t < 8; ++t) dd[t] = c3_ldt<TT>(pa + t * ES);
#pragma unroll 2
   for (int l = lb; l >= 0; --l) {
      int x[8];
#pragma unroll
      for (int t = 0; t < 8; ++t) x[t] = __shfl_sync(FULL, v, t);
      const unsigned wn2 = (unsigned)c3_lds(ewp);
      ewp -= 4;
      const AT pn = tabA + ES * ((wn & 0xfffffu) + min(lane, (int)((wn >> 20) & 63) - 1) * (wn >> 26));
      int dn[8];
#pragma unroll
      for (int t = 0; t < 8; ++t) dn[t] = c3_ldt<TT>(pn + t * ES);
      int a0 = __viaddmin_s32(x[0], dd[0], C3_INF), a1 = __viaddmin_s32(x[1], dd[1], C3_INF);
#pragma unroll
      for (int t = 2; t < 8; t += 2) { a0 = __viaddmin_s32(x[t], dd[t], a0); a1 = __viaddmin_s32(x[t + 1], dd[t + 1], a1); }
      const int ma = (w >> 20) & 63, mb = w >> 26;
      if (mb > 8) {
#pragma unroll 1
         for (int t0 = 8; t0 < mb; t0 += 8) c3_rows8<TT, 32>(pa, ES, v, t0, a0, a1);
      }
      const int acc = lane < ma ? min(a0, a1) : C3_INF;
      if (lane < ma) c3_sts(brow, acc);
      brow += mp4;
      v = acc;
      w = wn; wn = wn2; pa = pn;
#pragma unroll
      for (int t = 0; t < 8; ++t) dd[t] = dn[t];
   }
   __syncwarp();
   for (int l = lb - lane; l >= 0; l -= 32) {
      const int m = (int)((s.ew[l + 1] >> 20) & 63);
      int mn = C3_INF;
      for (int j = 0; j < m; ++j) mn = min(mn, s.Brev[(S - l) * mp + j]);
      s.minBrev[S - l] = mn;
   }
   __syncwarp();
}

// pd[l] = sum_{l' < l} dmin(c_l', c_l'+1), l = 0 .. S  (one warp, chunks of 32 layers)
template <typename TT> static __device__ void c3_pd_scan(const C3Sm<TT> &s, int S, int lane)
{
   constexpr unsigned FULL = 0xffffffffu;
   int carry = 0;
   for (int l0 = 0; l0 <= S; l0 += 32) {
      const int l = l0 + lane;
      int x = 0;   // dmin of edge (l-1 -> l)
      if (l >= 1 && l <= S) x = s.dmin[s.lcl[l - 1] * s.C + s.lcl[l]];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
         const int t = __shfl_up_sync(FULL, x, d);
         if (lane >= d) x += t;
      }
      if (l <= S) s.pd[l] = carry + x;
      carry += __shfl_sync(FULL, x, 31);
   }
   __syncwarp();
}

// arr[lo .. hi] moves by one position (delta = +1 / -1); one warp
static __device__ __forceinline__ void c3_shift1(int *arr, int lo, int hi, int delta, int lane)
{
   if (delta > 0) {
      for (int top = hi; top >= lo; top -= 32) {       // chunks from the top
         const int k = top - lane;
         int v = 0;
         if (k >= lo) v = arr[k];
         __syncwarp();
         if (k >= lo) arr[k + 1] = v;
         __syncwarp();
      }
   } else {
      for (int bot = lo; bot <= hi; bot += 32) {
         const int k = bot + lane;
         int v = 0;
         if (k <= hi) v = arr[k];
         __syncwarp();
         if (k <= hi) arr[k - 1] = v;
         __syncwarp();
      }
   }
}

// ---- glibc rand() ring (math_common.cpp:21-28; random_r TYPE_3: x_i = x_{i-3} + x_{i-31}, output >> 1) -------------------
// ring[p & (C3_RING - 1)] = raw value number p of this call's stream; numbers 0 .. 30 are the state at entry.  One warp
// extends it 31 values at a time (three interleaved prefix sums: four shuffle-adds); every other thread tracks `gen`
// with the same arithmetic.
static __device__ __forceinline__ unsigned c3_gen_after(unsigned gen, unsigned need)
{
   return gen < need ? gen + 31u * ((need - gen + 30u) / 31u) : gen;
}
static __device__ __forceinline__ void c3_ring_fill(unsigned *ring, unsigned gen, unsigned need, int lane)
{
   constexpr unsigned FULL = 0xffffffffu, M = C3_RING - 1;
   while (gen < need) {
      unsigned a = 0;
      if (lane < 31) a = ring[(gen - 31 + lane) & M];
      if (lane < 3) a += ring[(gen - 3 + lane) & M];
#pragma unroll
      for (int d = 3; d < 32; d <<= 1) {               // x_{gen+k} = a_k + x_{gen+k-3}: inclusive scan with stride 3
         const unsigned t = __shfl_up_sync(FULL, a, d);
         if (lane >= d) a += t;
      }
      if (lane < 31) ring[(gen + lane) & M] = a;
      gen += 31;
      __syncwarp();
   }
}

// ---- lower bound of a proposal in O(1) (one thread): minF(s0) + dist_between + minB(sK) --------------------------------
// tryToMoveLowerBound vns_sop_path.cpp:799-930, tryToAddLowerBound :207-243, tryToExchangeLowerBound :573-633,
// tryToReplaceLowerBound :423-454; integers, so the reference's left-to-right sums equal these prefix differences
// Branch-free: the lanes of a warp hold different cases (up / down / from outside), and one warp executing the cases one
// after the other pays every case's chain of dependent loads.  Every case is  minF[s0] + sum of <= 4 cluster-pair minima +
// (pd[p1] - pd[p0]) + minBrev[S - sK]  with the pairs given as layer indices (-1 = the foreign cluster VNS[from]).
template <typename TT> static __device__ __forceinline__ int c3_lower_bound(const C3Sm<TT> &s, int T, int kind, int from, int to)
{
   const int S = T + 1, C = s.C;
   const int *lcl = s.lcl;
   const bool in = from < T;
   const int cfor = s.vns[from];
   const int lo = min(from, to), hi = max(from, to);
   const bool up = to > from;
   int s0, sK, p0 = 0, p1 = 0;
   int a0, b0, a1, b1, a2 = 0, b2 = 0, a3 = 0, b3 = 0;      // pairs (a_k -> b_k); pairs 2 and 3 may be absent
   bool has2 = false, has3 = false;
   if (kind == 1) {
      const int i = lo + 1, j = hi + 1;
      // in, up:   i-1, i+1 .. j-1, i, j        in, down: i-1, j, i .. j-1, j+1        out: to, c, to+1
      s0 = in ? i - 1 : to;
      sK = in ? (up ? j : j + 1) : to + 1;
      a0 = s0;                     b0 = in ? (up ? i + 1 : j) : -1;
      a1 = in ? (up ? j - 1 : j) : -1;  b1 = in ? i : to + 1;
      a2 = up ? i : j - 1;         b2 = up ? j : j + 1;       has2 = in;
      p0 = in ? (up ? i + 1 : i) : 0;  p1 = in ? j - 1 : 0;
   } else {
      const int i = lo, j = hi + 1;
      // in: i, j, i+2 .. j-1, i+1, j+1        out: to, c, to+2
      s0 = in ? i : to;
      sK = in ? j + 1 : to + 2;
      a0 = s0;                     b0 = in ? j : -1;
      a1 = in ? j : -1;            b1 = in ? i + 2 : to + 2;
      a2 = j - 1;                  b2 = i + 1;                has2 = in;
      a3 = i + 1;                  b3 = j + 1;                has3 = in;
      p0 = in ? i + 2 : 0;         p1 = in ? j - 1 : 0;
   }
   // layer -> cluster (absent pairs read layer 0: a valid address, their term is dropped)
   const int ca0 = lcl[a0], cb0 = b0 < 0 ? cfor : lcl[max(b0, 0)];
   const int ca1 = a1 < 0 ? cfor : lcl[max(a1, 0)], cb1 = lcl[b1];
   const int ca2 = lcl[has2 ? a2 : 0], cb2 = lcl[has2 ? b2 : 0];
   const int ca3 = lcl[has3 ? a3 : 0], cb3 = lcl[has3 ? b3 : 0];
   const int e0 = s.minF[s0], e1 = s.minBrev[S - sK], e2 = s.pd[p1], e3 = s.pd[p0];
   const int d0 = s.dmin[ca0 * C + cb0], d1 = s.dmin[ca1 * C + cb1], d2 = s.dmin[ca2 * C + cb2], d3 = s.dmin[ca3 * C + cb3];
   return (e0 + e1) + (e2 - e3) + (d0 + d1) + ((has2 ? d2 : 0) + (has3 ? d3 : 0));
}

// ---- one proposal by one sub-warp group of G lanes (lane = destination sample; mp <= G) ----------------------------------
// chain of K table blocks from F[s0] to B[sK] (tryToMove :932-1154, tryToAdd :245-304, tryToExchange :635-797,
// tryToReplace :456-513; CandCode of eval.cuh).  A warp advances its 32 / G proposals in lock step.  First every lane
// describes blocks of its group's chain and writes their words (offset | rows << 20 | columns << 26; bit 19: some group of
// the warp has more than eight rows in this step) to the group's list in shared memory; the step loop then is the
// re-scan's loop: SHFL x 8 -> next step's rows -> VIADDMNMX chain.  Integers: the reference's two closing associations
// give the same number.  Returns v[j] + B[sK][j] of this lane (INF-like beyond the last layer): the caller's acceptance
// test is monotone in the length, so it tests every lane's value instead of reducing them first.
// All 32 lanes must call; K = 0: idle group.
template <typename TT, int G>
static __device__ __forceinline__ int c3_eval(const C3Sm<TT> &s, int T, const CandCode &cc, int gl, unsigned *gw)
{
   constexpr unsigned FULL = 0xffffffffu;
   typedef typename C3TT<TT>::addr AT;
   constexpr int ES = (int)sizeof(typename C3TT<TT>::elem);
   const int S = T + 1, mp = s.mp, N = s.N, K = cc.K;
   int v = C3_INF, bK = C3_INF;
   if (K) {
      const int m0 = s.cinfo[s.lcl[cc.s0]].y, mK = s.cinfo[s.lcl[cc.sK]].y;
      v = gl < m0 ? s.F[cc.s0 * mp + gl] : C3_INF;
      bK = gl < mK ? s.Brev[(S - cc.sK) * mp + gl] : C3_INF;
   }
   const int Kmax = __reduce_max_sync(FULL, K);
   for (int e0 = 0; e0 < Kmax + 2; e0 += G) {
      const int k = e0 + gl + 1;
      unsigned w = C3_EW_NONE;
      int big = 0;
      if (k <= K) {
         const int la = cand_layer(cc, k - 1), lb = cand_layer(cc, k);
         const int2 cia = s.cinfo[la < 0 ? cc.foreign : s.lcl[la]], cib = s.cinfo[lb < 0 ? cc.foreign : s.lcl[lb]];
         w = (unsigned)(cia.x * N + cia.y * cib.x) | ((unsigned)cia.y << 20) | ((unsigned)cib.y << 26);
         big = cia.y > 8;
      }
      if (G > 8) {
#pragma unroll
         for (int sft = G; sft < 32; sft <<= 1) big |= __shfl_xor_sync(FULL, big, sft);
         w |= (unsigned)big << 19;
      }
      if (k <= Kmax + 2) gw[k] = w;
   }
   __syncwarp();
   const AT tabA = c3_tab_addr<TT>(s.tab);
   const int glk = (int)c3_keep((unsigned)gl);
   unsigned gwp = c3_keep(c3_saddr(gw + 1));
   unsigned w = (unsigned)c3_lds(gwp), wn = (unsigned)c3_lds(gwp + 4);
   gwp += 8;
   AT pa = tabA + ES * ((w & 0x7ffffu) + min(glk, (int)(w >> 26) - 1));
   int dd[8];
#pragma unroll
   for (int t = 0; t < 8; ++t) dd[t] = c3_ldt<TT>(pa + t * ES * (w >> 26));
#pragma unroll 2
   for (int k = 1; k <= Kmax; ++k) {
      int x[8];
#pragma unroll
      for (int t = 0; t < 8; ++t) x[t] = __shfl_sync(FULL, v, t, G);
      const unsigned wn2 = (unsigned)c3_lds(gwp);
      gwp += 4;
      const unsigned sbn = ES * (wn >> 26);
      const AT pn = tabA + ES * ((wn & 0x7ffffu) + min(glk, (int)(wn >> 26) - 1));
      int dn[8];
#pragma unroll
      for (int t = 0; t < 8; ++t) dn[t] = c3_ldt<TT>(pn + t * sbn);
      int a0 = __viaddmin_s32(x[0], dd[0], C3_INF), a1 = __viaddmin_s32(x[1], dd[1], C3_INF);
#pragma unroll
      for (int t = 2; t < 8; t += 2) { a0 = __viaddmin_s32(x[t], dd[t], a0); a1 = __viaddmin_s32(x[t + 1], dd[t + 1], a1); }
      const int mb = w >> 26;
      if (G > 8 && ((w >> 19) & 1)) {            // warp-uniform: rows 8 .. G - 1 for everybody (INF samples beyond a block)
#pragma unroll 1
         for (int t0 = 8; t0 < G; t0 += 8) c3_rows8<TT, G>(pa, ES * mb, v, t0, a0, a1);
      }
      if (k <= K) v = glk < mb ? min(a0, a1) : C3_INF;
      w = wn; wn = wn2; pa = pn;
#pragma unroll
      for (int t = 0; t < 8; ++t) dd[t] = dn[t];
   }
   return K ? min(v + bK, C3_INF) : C3_INF;          // lanes beyond the last layer hold INF + INF < 2^31
}

// insertRandom accept (vns_sop_solver.cpp:621-646): VNS[from] is re-inserted immediately before the element that was at
// index `to`; only the final state matters.  tg = lcl + 1.  exchangeRandom accept (:807-832):
// VNS[from] <-> VNS[to]; `to` is always in the path.  All threads; T (register copy) is updated; returns the new length.
template <typename TT> static __device__ int c3_apply(const C3Sm<TT> &s, int kind, int from, int to, int &T)
{
   const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
   int *tg = s.lcl + 1;
   int lf, lb;
   if (kind == 1) {
      const int c = s.vns[from];
      __syncthreads();                    // every thread holds c; nobody still evaluates
      // tg[tlo .. thi] and vns[vlo .. vhi] move by `delta`, then c lands at index `at`
      int tlo, thi, vlo, vhi, delta, at;
      if (from < T) {
         if (to > from) {                 // travels up to index to-1; layers from+1 .. to change
            tlo = vlo = from + 1; thi = vhi = to - 1; delta = -1; at = to - 1;
            lf = from + 1; lb = to;
         } else {                         // travels down to index `to`; layers to+1 .. from+1 change
            tlo = vlo = to; thi = vhi = from - 1; delta = +1; at = to;
            lf = to + 1; lb = from + 1;
         }
      } else {                            // from outside the path (from >= T >= to): one more target at index `to`
         tlo = vlo = to; thi = T; vhi = from - 1; delta = +1; at = to;      // tg's range includes the goal's entry
         T += 1;
         lf = to + 1; lb = to + 1;        // B rows of the old layers to+1 .. S keep their slots
      }
      if (warp < 2) {
         int *arr = warp == 0 ? tg : s.vns;
         c3_shift1(arr, warp == 0 ? tlo : vlo, warp == 0 ? thi : vhi, delta, lane);
         if (lane == 0) arr[at] = c;
      }
   } else {
      if (threadIdx.x == 0) {
         const int cf = s.vns[from], ct = s.vns[to];
         s.vns[to] = cf; s.vns[from] = ct;
         tg[to] = cf;
         if (from < T) tg[from] = ct;
      }
      const int lo = (from < T && from < to) ? from : to;
      const int hi = (from < T && from > to) ? from : to;
      lf = lo + 1; lb = hi + 1;
   }
   __syncthreads();
   const int S = T + 1;
   if (warp == 0) c3_fwd_scan(s, S, lf, lane);
   else if (warp == 1) c3_bwd_scan(s, S, lb, lane);
   else if (warp == 2) c3_pd_scan(s, S, lane);
   __syncthreads();
   return s.minF[S];                                                       // getPathLength :1660-1678
}

// ---- the local search ------------------------------------------------------------------------------------------------------
// g: generator state in canonical form at entry and exit.  ctl.T / L / ema are read at entry and written back at exit.
template <typename TT, int G>
static __device__ void c3_local_search(const C3Sm<TT> &s, C3Ctl &ctl, RngState &g, int kind, int num_changes, double budget, int skipflag)
{
   constexpr int NG = 32 / G, NGRP = C3_WARPS * NG;
   constexpr unsigned FULL = 0xffffffffu, M = C3_RING - 1;
   const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, gl = lane % G, gq = lane / G;
   const int nvns = ctl.nvns;
   const int gap = kind == 1 ? 1 : 2;
   // lengths and bounds are integers: x > budget <=> x > floor(budget)
   const int ibudget = budget >= 2147483647.0 ? 2147483647 : (budget < 0.0 ? -1 : (int)floor(budget));
   __syncthreads();
   if (!(kind == 1 ? (nvns >= 3) : (nvns >= 4 && ctl.T > 0))) return;    // guards :498 and :711 (uniform)
   if (warp == C3_WARPS - 1) {
      if (lane < 31) s.ring[lane] = (unsigned)g.r[(g.f + lane) % 31];    // oldest first: r[f] is x_{i-31}
      __syncwarp();
      c3_ring_fill(s.ring, 31u, 31u + C3_AHEAD, lane);
   }
   if (threadIdx.x < 4) ctl.first[threadIdx.x] = C3_NONE;
   unsigned pos = 31u, gen = c3_gen_after(31u, 31u + C3_AHEAD), pc = 0;
   int T = ctl.T, L = ctl.L, ema = ctl.ema, err = 0;
   long long n_prop = 0, n_skip = 0, n_acc = 0, n_round = 0, prof0 = 0, prof1 = 0, prof2 = 0, prof3 = 0;
   __syncthreads();
   int var = 0, empty_rounds = 0;
   while (var < num_changes) {
      const int S = T + 1, remaining = num_changes - var;
      const long long t0 = clock64();
      // ---- 1. this round's window: every warp derives all of it ----------------------------------------------------------
      const int target = min(max(2 * ema, 32), 32 * C3_MAXCH);
      int nch = target > 64 ? 4 : (target > 32 ? 2 : 1);
      while (nch > 1 && nch * 32 > remaining) nch >>= 1;
      int nvalid = 0, nsurv = 0, lastp = -1;
      const int tohi = kind == 1 ? T : T - 1;
#pragma unroll 1
      for (int ch = 0; ch < nch; ++ch) {
         const int p = ch * 32 + lane;
         const unsigned r1 = s.ring[(pos + 2 * p) & M] >> 1, r2 = s.ring[(pos + 2 * p + 1) & M] >> 1;
         const int from = (int)(((unsigned long long)r1 * (unsigned)nvns) >> 31);            // :507 / :723
         const int to = (int)(((unsigned long long)r2 * (unsigned)(tohi + 1)) >> 31);        // :508 / :724
         bool valid = abs(from - to) > gap;                                                 // :510-519 / :726-730
         const unsigned bv0 = __ballot_sync(FULL, valid);
         // the call ends after exactly num_changes proposals: later ones do not exist
         valid = valid && nvalid + __popc(bv0 & ((1u << lane) - 1)) < remaining;
         bool surv = valid;
         if (valid && skipflag) surv = !(c3_lower_bound(s, T, kind, from, to) > ibudget);   // :539,:574,:749,:779
         const unsigned bv = __ballot_sync(FULL, valid), bs = __ballot_sync(FULL, surv);
         if (surv) s.surv[nsurv + __popc(bs & ((1u << lane) - 1))] = p;     // every warp writes the same list
         if (lane == 0) { ctl.balv[ch] = bv; ctl.bals[ch] = bs; }
         nvalid += __popc(bv); nsurv += __popc(bs);
         if (bv) lastp = ch * 32 + 31 - __clz(bv);
      }
      __syncwarp();
      const long long t1 = clock64();
      // ---- 2. survivors in draw order, one group each; stop behind the first pass that accepts ---------------------------
      int fp = C3_NONE;
      for (int base = 0; base < nsurv; base += NGRP) {
         const unsigned cell = pc & 3;
         if (threadIdx.x == 0) ctl.first[(pc + 2) & 3] = C3_NONE;       // next used two barriers from here
         if (base + warp * NG < nsurv) {
            const int sidx = base + warp * NG + gq;
            const bool act = sidx < nsurv;
            const int p = act ? s.surv[sidx] : 0;
            const unsigned r1 = s.ring[(pos + 2 * p) & M] >> 1, r2 = s.ring[(pos + 2 * p + 1) & M] >> 1;
            const int from = (int)(((unsigned long long)r1 * (unsigned)nvns) >> 31);
            const int to = (int)(((unsigned long long)r2 * (unsigned)(tohi + 1)) >> 31);
            const bool in = from < T;
            const int cf = s.vns[from];
            int op = -1, a = 0;
            if (act) {
               if (kind == 1) { op = in ? VNSGPU_OP_MOVE : VNSGPU_OP_ADD; a = in ? from : cf; }
               else { op = in ? VNSGPU_OP_EXCHANGE : VNSGPU_OP_REPLACE; a = in ? from : cf; }
            }
            const CandCode cc = decode_cand(T, s.C, op, a, to);
            const int fin = c3_eval<TT, G>(s, T, cc, gl, s.gw + (warp * NG + gq) * s.gwcap);
            // integers: (L - 0.0) + (len - L) == len (:580,:583); len <= budget <=> len <= floor(budget); L - len > 1e-5 <=> L > len.
            // The test is monotone in the length: the group's minimum passes iff some lane's value does
            if (act) {
               double dR = 0.0;
               if (!in) dR = kind == 1 ? s.reward[cf]                                              // :581
                                       : 0.0 + (s.reward[cf] - s.reward[s.vns[to]]);              // :784
               const bool acc = cc.K > 0 && fin < C3_INF && fin <= ibudget && (dR > 0 || L > fin);   // :619,:805-806
               if (acc) atomicMin(&ctl.first[cell], p);
               if (cc.K == 0) ctl.bad = 1;
            }
         }
         __syncthreads();
         ++pc;
         fp = ctl.first[cell];
         if (fp != C3_NONE) break;
      }
      const long long t2 = clock64();
      // ---- 3. what the reference's sequential loop consumed ----------------------------------------------------------------
      const bool found = fp != C3_NONE;
      int used = nvalid, passed = nsurv;
      unsigned adv = lastp >= 0 ? 2u * (unsigned)(lastp + 1) : 64u * (unsigned)nch;   // no valid pair: all draws consumed
      int wfrom = 0, wto = 0;
      if (found) {
         used = 0; passed = 0;
#pragma unroll 1
         for (int ch = 0; ch <= (fp >> 5); ++ch) {
            const unsigned upto = ch < (fp >> 5) ? FULL : (2u << (fp & 31)) - 1u;
            used += __popc(ctl.balv[ch] & upto); passed += __popc(ctl.bals[ch] & upto);
         }
         adv = 2u * (unsigned)(fp + 1);
         const unsigned r1 = s.ring[(pos + 2 * fp) & M] >> 1, r2 = s.ring[(pos + 2 * fp + 1) & M] >> 1;
         wfrom = (int)(((unsigned long long)r1 * (unsigned)nvns) >> 31);
         wto = (int)(((unsigned long long)r2 * (unsigned)(tohi + 1)) >> 31);
      }
      pos += adv;
      var += used;
      n_prop += used; n_skip += used - passed; n_round += 1;
      if (found) { n_acc += 1; ema = (3 * ema + used + 3) / 4; }
      else ema = min(2 * max(ema, 16 * nch), 1 << 20);
      empty_rounds = used ? 0 : empty_rounds + 1;
      if (empty_rounds > VNS_MAX_DRAWS / 64) err = 3;
      const long long t3 = clock64();
      // ---- 4. the generator runs ahead while the move is applied -----------------------------------------------------------
      if (warp == C3_WARPS - 1) c3_ring_fill(s.ring, gen, pos + C3_AHEAD, lane);
      gen = c3_gen_after(gen, pos + C3_AHEAD);
      if (found) {
         L = c3_apply(s, kind, wfrom, wto, T);
         if (L != s.minBrev[T + 1]) err = 4;                              // addPoint's check :80-88 (integers: exact)
      } else
         __syncthreads();
      prof0 += t1 - t0; prof1 += t2 - t1; prof2 += t3 - t2; prof3 += clock64() - t3;
      if (err) break;
   }
   __syncthreads();
   if (warp == 0) {
      if (lane < 31) g.r[lane] = (int32_t)s.ring[(pos - 31 + lane) & M];
      if (lane == 0) {
         g.f = 0; g.b = 28; g.pad = 0;
         ctl.T = T; ctl.L = L; ctl.ema = ema;
         if (err && !ctl.err) ctl.err = err;
         ctl.st.proposals += n_prop; ctl.st.lb_skips += n_skip; ctl.st.evals += n_prop - n_skip; ctl.st.accepts += n_acc;
         ctl.st.rounds += n_round;
         ctl.prof[0] += prof0; ctl.prof[1] += prof1; ctl.prof[2] += prof2; ctl.prof[3] += prof3;
      }
   }
   __syncthreads();
}

// ---- global <-> shared path state ------------------------------------------------------------------------------------------
template <typename TT> static __device__ void c3_load(const InstDev &I, const PathDev &P, const C3Sm<TT> &s, C3Ctl &ctl)
{
   const int T = P.hdr[HDR_T], nvns = P.hdr[HDR_NVNS], S = T + 1, mp = I.mp;
   __syncthreads();
   for (int l = threadIdx.x; l <= S; l += blockDim.x) {
      s.lcl[l] = l == 0 ? I.start : (l == S ? I.goal : P.tg[l - 1]);
      s.minF[l] = (int)P.minF[l];
      s.minBrev[S - l] = (int)P.minB[l];
   }
   for (int k = threadIdx.x; k < nvns; k += blockDim.x) s.vns[k] = P.vns[k];
   for (int e = threadIdx.x; e < (S + 1) * mp; e += blockDim.x) {
      const int l = e / mp, j = e - l * mp;
      s.F[e] = (int)P.F[e];
      s.Brev[(S - l) * mp + j] = (int)P.B[e];
   }
   if (threadIdx.x == 0) { ctl.T = T; ctl.nvns = nvns; ctl.L = (int)P.hd[HD_LEN]; ctl.err = 0; ctl.bad = 0; }
   __syncthreads();
   if ((threadIdx.x >> 5) == 2) c3_pd_scan(s, S, threadIdx.x & 31);
   __syncthreads();
}

// entries beyond a layer's cluster size are never read by anyone and are not written back
template <typename TT> static __device__ void c3_store(const InstDev &I, const PathDev &P, const C3Sm<TT> &s, C3Ctl &ctl)
{
   __syncthreads();
   const int T = ctl.T, nvns = ctl.nvns, S = T + 1, mp = I.mp;
   for (int l = threadIdx.x; l <= S; l += blockDim.x) {
      if (l >= 1 && l <= T) P.tg[l - 1] = s.lcl[l];
      P.minF[l] = (double)s.minF[l];
      P.minB[l] = (double)s.minBrev[S - l];
   }
   for (int k = threadIdx.x; k < nvns; k += blockDim.x) P.vns[k] = s.vns[k];
   for (int e = threadIdx.x; e < (S + 1) * mp; e += blockDim.x) {
      const int l = e / mp, j = e - l * mp;
      if (j < s.cinfo[s.lcl[l]].y) {
         P.F[e] = (double)s.F[e];
         P.B[e] = (double)s.Brev[(S - l) * mp + j];
      }
   }
   __syncthreads();
   if (threadIdx.x == 0) {
      P.hdr[HDR_T] = T;
      P.hdr[HDR_NVNS] = nvns;
      P.hd[HD_LEN] = (double)s.minF[S];
      P.hd[HD_LEN_REAR] = (double)s.minBrev[S];
      P.hd[HD_REWARD] = path_reward(I, P.tg, T);                // :650,:835
      if (ctl.bad && !ctl.err) ctl.err = 3;
      if (ctl.err && !P.hdr[HDR_ERR]) P.hdr[HDR_ERR] = ctl.err;
   }
   __syncthreads();
}

// getReward (vns_sop_path.cpp:2016-2031) of the shared-memory state, one thread
template <typename TT> static __device__ double c3_reward(const InstDev &I, const C3Sm<TT> &s, int T)
{
   double r = 0;
   for (int t = 0; t < T; ++t) r += s.reward[s.lcl[t + 1]];
   const double rs = s.reward[I.start];
   if (rs > 0) {
      r += rs;
      if (I.start != I.goal) r += s.reward[I.goal];
   }
   return r;
}

// ---- shake + fitDOPtoBudget on the shared-memory state ---------------------------------------------------------------------
// pathInsert(.., 1) (kind 1, vns_sop_solver.cpp:931-999) / pathExchange(.., 1) (kind 2, :1001-1089): the vector splices of
// search.cuh's cta_shake (thread 0; O(|VNS|) integer moves in shared memory), then fitDOPtoBudget (:364-385) as forward
// layers only: tryToAdd(VNS[T], T) is the goal-row minimum of the path with VNS[T] appended (integers: exact).
template <typename TT> static __device__ void c3_shake(const InstDev &I, const C3Sm<TT> &s, C3Ctl &ctl, int kind, double budget, RngState &rng)
{
   __shared__ int s_first;
   const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
   __syncthreads();
   if (threadIdx.x == 0) {
      int T = ctl.T, nvns = ctl.nvns;
      int *vns = s.vns, *tg = s.lcl + 1, *seg = s.seg;
      const int sizeV = nvns - 1;
      int first = T + 2;   // lowest changed target index (for the forward re-scan)
      if (kind == 1 && sizeV >= 2) {
         int r[3];
         draw_distinct_sorted(rng, sizeV, 3, r);
         int fs = r[0], fe = r[1], to = r[2], up = 1;
         if (rng_int(rng, 0, 1) == 1) { to = r[0]; fs = r[1]; fe = r[2]; up = 0; }        // :951-956
         const int n = fe - fs + 1;
         for (int x = 0; x < n; ++x) seg[x] = vns[fs + x];
         first = min(fs, to);
         if (up) {                                                                      // :958-972
            vec_insert1(vns, nvns, to, seg, n);
            vec_erase1(vns, nvns, fs, fe);
            if (to < T) vec_insert1(tg, T, to, seg, n);
            if (fs < T) vec_erase1(tg, T, fs, min(fe, T - 1));
         } else {                                                                       // :974-990
            vec_erase1(vns, nvns, fs, fe);
            vec_insert1(vns, nvns, to, seg, n);
            if (fs < T) vec_erase1(tg, T, fs, min(fe, T - 1));
            if (to < T) vec_insert1(tg, T, to, seg, n);
         }
      } else if (kind == 2 && sizeV >= 3) {
         int r[4];
         draw_distinct_sorted(rng, sizeV, 4, r);
         int fs = r[0], fe = r[1], ts = r[2], te = r[3], up = 1;
         if (rng_int(rng, 0, 1) == 1) { fs = r[2]; fe = r[3]; ts = r[0]; te = r[1]; up = 0; }  // :1022-1028
         const int nf = fe - fs + 1, nt = te - ts + 1;
         int *vf = seg, *vt = seg + nf;
         for (int x = 0; x < nf; ++x) vf[x] = vns[fs + x];
         for (int x = 0; x < nt; ++x) vt[x] = vns[ts + x];
         first = min(fs, ts);
         if (up) {                                                                      // :1030-1051
            vec_erase1(vns, nvns, ts, te); vec_insert1(vns, nvns, ts, vf, nf);
            vec_erase1(vns, nvns, fs, fe); vec_insert1(vns, nvns, fs, vt, nt);
            if (ts < T) { vec_erase1(tg, T, ts, min(te, T - 1)); vec_insert1(tg, T, ts, vf, nf); }
            if (fs < T) { vec_erase1(tg, T, fs, min(fe, T - 1)); vec_insert1(tg, T, fs, vt, nt); }
         } else {                                                                       // :1053-1080
            vec_erase1(vns, nvns, fs, fe); vec_insert1(vns, nvns, fs, vt, nt);
            vec_erase1(vns, nvns, ts, te); vec_insert1(vns, nvns, ts, vf, nf);
            if (fs < T) { vec_erase1(tg, T, fs, min(fe, T - 1)); vec_insert1(tg, T, fs, vt, nt); }
            if (ts < T) { vec_erase1(tg, T, ts, min(te, T - 1)); vec_insert1(tg, T, ts, vf, nf); }
         }
      }
      tg[T] = I.goal;
      ctl.T = T; s_first = first;
   }
   __syncthreads();
   if (warp == 0) {
      int T = ctl.T;
      const int nvns = ctl.nvns, goal = I.goal;
      int *tg = s.lcl + 1;
      // forward DP of the spliced path (layers below the first changed target are still valid)
      c3_fwd_scan(s, T + 1, min(s_first + 1, T + 1), lane);
      // :366-368  pop the last target while the path is over budget
      while ((double)s.minF[T + 1] > budget && T > 0) {
         --T;
         if (lane == 0) tg[T] = goal;
         __syncwarp();
         c3_fwd_layer(s, T + 1, lane);
      }
      // :372-383  append VNS[T] while tryToAdd(VNS[T], T) fits
      while (T < nvns) {
         if (lane == 0) { tg[T] = s.vns[T]; tg[T + 1] = goal; }
         __syncwarp();
         c3_fwd_layer(s, T + 1, lane);
         c3_fwd_layer(s, T + 2, lane);
         if ((double)s.minF[T + 2] <= budget) { ++T; continue; }
         if (lane == 0) tg[T] = goal;
         __syncwarp();
         c3_fwd_layer(s, T + 1, lane);
         break;
      }
      if (lane == 0) ctl.T = T;
      __syncwarp();
   }
   __syncthreads();
   {  // full backward scan of the fitted path
      const int T = ctl.T, S = T + 1, mS = s.cinfo[I.goal].y;
      if (warp == 1) {
         if (lane < mS) s.Brev[lane] = 0;
         if (lane == 0) s.minBrev[0] = 0;
         __syncwarp();
         c3_bwd_scan(s, S, S - 1, lane);
      } else if (warp == 2)
         c3_pd_scan(s, S, lane);
      __syncthreads();
      if (threadIdx.x == 0) {
         ctl.L = s.minF[S];
         if (s.minF[S] != s.minBrev[S] && !ctl.err) ctl.err = 4;
      }
      __syncthreads();
   }
}

// ---- kernels ----------------------------------------------------------------------------------------------------------------
template <typename TT, int G>
__global__ void __launch_bounds__(C3_THREADS) k_chain3_ls(C3Env E, PathDev P, int kind, int num_changes, double budget, int skipflag,
                                                          RngState *rng, LsStats *stats, long long *prof)
{
   extern __shared__ __align__(16) unsigned char smem_c[];
   __shared__ C3Ctl ctl;
   __shared__ RngState g;
   const C3Sm<TT> s = c3_carve<TT>(smem_c, E.I, E.tab_g);
   if (threadIdx.x == 0) {
      g = *rng;
      ctl.st = LsStats{0, 0, 0, 0, 0}; ctl.prof[0] = ctl.prof[1] = ctl.prof[2] = ctl.prof[3] = 0; ctl.ema = 16;
   }
   c3_load_tables(E, s);
   c3_load(E.I, P, s, ctl);
   c3_local_search<TT, G>(s, ctl, g, kind, num_changes, budget, skipflag);
   c3_store(E.I, P, s, ctl);
   if (threadIdx.x == 0) {
      *rng = g; *stats = ctl.st;
      for (int k = 0; k < 4; ++k) prof[k] += ctl.prof[k];
   }
}

// VnsSopSolver::iterate (vns_sop_solver.cpp:82-260).  tour lives in global memory (it changes only on improvements); the
// working path of a VNS step (actualVNSGOPPath, :195) exists only in shared memory.
template <typename TT, int G>
static __device__ void c3_vns_body(const C3Env &E, const PathDev &tour, const VnsParamsDev &prm, RngState *rng, VnsResultDev *res,
                                   double *impr_reward, double *impr_length, double *impr_ms, long long *prof)
{
   extern __shared__ __align__(16) unsigned char smem_c[];
   __shared__ C3Ctl ctl;
   __shared__ RngState g;
   __shared__ int s_stop, s_improved;
   const InstDev &I = E.I;
   double *vbuf_all = reinterpret_cast<double *>(smem_c);       // greedy scratch (start of the run only): under the table
   const C3Sm<TT> s = c3_carve<TT>(smem_c, I, E.tab_g);
   if (threadIdx.x == 0) {
      g = *rng; s_stop = 0;
      ctl.st = LsStats{0, 0, 0, 0, 0}; ctl.prof[0] = ctl.prof[1] = ctl.prof[2] = ctl.prof[3] = 0; ctl.err = 0; ctl.bad = 0; ctl.ema = 16;
   }
   __syncthreads();
   const double t0 = now_ms();
   cta_update(I, tour, tour.hdr[HDR_T]);
   if (prm.greedy) cta_greedy(I, tour, prm.budget, vbuf_all);            // :127-129
   __syncthreads();
   c3_load_tables(E, s);
   const int num_avail = tour.hdr[HDR_NVNS];                            // :115
   int last_impr = 1, it = 0, n_impr = 0;                               // :145
   double t_last = now_ms() - t0;
   if (threadIdx.x == 0 && impr_reward && n_impr < prm.max_improvements) {   // :150-154
      impr_reward[n_impr] = tour.hd[HD_REWARD]; impr_length[n_impr] = tour.hd[HD_LEN]; impr_ms[n_impr] = t_last;
   }
   n_impr = 1;
   __syncthreads();
   bool stop = false;
   while (!stop) {
      ++it;
      if (it >= prm.num_iterations) stop = true;                         // :178
      if (it - last_impr >= prm.num_iterations_unimproved) stop = true;  // :182
      int k = 1;
      while (k <= 2) {
         if (threadIdx.x == 0) {
            if (prm.max_time_ms > 0 && now_ms() - t0 >= prm.max_time_ms) s_stop = 1;   // :189-193
            if (prm.stop_flag) {      // multi-start race: not part of the reference's loop
               if (tour.hd[HD_REWARD] >= prm.target_reward) atomicExch(prm.stop_flag, 1);
               if (*(volatile int *)prm.stop_flag) s_stop = 2;
            }
         }
         __syncthreads();
         if (s_stop == 2) break;
         c3_load(I, tour, s, ctl);                                       // :195-197
         const double r0 = tour.hd[HD_REWARD], l0 = (double)ctl.L;
         c3_shake(I, s, ctl, k, prm.budget, g);                          // :202
         c3_local_search<TT, G>(s, ctl, g, k, num_avail * num_avail, prm.budget, prm.skip);   // :207,:346
         if (threadIdx.x == 0) {
            const double r1 = c3_reward(I, s, ctl.T), l1 = (double)ctl.L;
            s_improved = (r1 > r0 || (r1 == r0 && l1 < l0));             // :213
         }
         __syncthreads();
         if (ctl.err) { stop = true; break; }
         if (s_improved) {
            last_impr = it;
            t_last = now_ms() - t0;
            c3_store(I, tour, s, ctl);                                   // :227-228
            if (threadIdx.x == 0) {
               if (impr_reward && n_impr < prm.max_improvements) {
                  impr_reward[n_impr] = tour.hd[HD_REWARD]; impr_length[n_impr] = tour.hd[HD_LEN]; impr_ms[n_impr] = t_last;
               }
               // :229-237 sanity check of the improved tour
               const double calc = path_walk_back(I, tour, tour.hdr[HDR_T], nullptr, nullptr);
               const double nl = tour.hd[HD_LEN];
               if (calc > prm.budget || nl > prm.budget || calc - nl > 0.01) tour.hdr[HDR_ERR] = 4;
            }
            __syncthreads();
            ++n_impr;
            if (tour.hdr[HDR_ERR]) { stop = true; break; }
            k = 1;                                                       // :241
         } else {
            ++k;                                                         // :243
         }
      }
      __syncthreads();
      if (s_stop) stop = true;
   }
   __syncthreads();
   if (threadIdx.x == 0) {
      res->num_iters = it; res->num_iters_last_impr = last_impr; res->n_improvements = n_impr;
      res->status = tour.hdr[HDR_ERR] ? tour.hdr[HDR_ERR] : (ctl.err ? ctl.err : (ctl.bad ? 3 : 0));
      res->reward = tour.hd[HD_REWARD]; res->length = tour.hd[HD_LEN];
      res->time_ms = now_ms() - t0; res->time_last_impr_ms = t_last;
      res->proposals = ctl.st.proposals; res->evals = ctl.st.evals; res->accepts = ctl.st.accepts;
      *rng = g;
      if (prof)
         for (int q = 0; q < 4; ++q) prof[q] += ctl.prof[q];
   }
   __syncthreads();
}

template <typename TT, int G>
__global__ void __launch_bounds__(C3_THREADS) k_chain3_vns_run(C3Env E, PathDev tour, VnsParamsDev prm, RngState *rng, VnsResultDev *res,
                                                               double *impr_reward, double *impr_length, double *impr_ms, long long *prof)
{
   c3_vns_body<TT, G>(E, tour, prm, rng, res, impr_reward, impr_length, impr_ms, prof);
}

// Multi-start (the reference's --batch=N, algorithm.cc:120-161): one thread block per run, each with its own seed and
// tour; no communication between runs.  Same slot layout as the second form's (chain2.cuh C2RunSlot).
struct C3RunSlot {
   PathDev tour;
   RngState rng;
   VnsResultDev res;
};

template <typename TT, int G>
__global__ void __launch_bounds__(C3_THREADS) k_chain3_vns_multi(C3Env E, C3RunSlot *slots, VnsParamsDev prm)
{
   C3RunSlot &sl = slots[blockIdx.x];
   c3_vns_body<TT, G>(E, sl.tour, prm, &sl.rng, &sl.res, nullptr, nullptr, nullptr, nullptr);
}
