// Trajectory-exact search chain, second form (exact int32 tables): insertRandom (vns_sop_solver.cpp:469-684) and
// exchangeRandom (:686-864) with the whole search state in shared memory, run by ONE thread block or by a GRID of
// cooperating thread blocks that hold replicas of the state.
//
// What changed against chain.cuh (which stays the IEEE double form):
//   * lower bound FIRST.  The reference skips a proposal whose lower bound exceeds the budget (:539,:574,:749,:779) -
//     70-75 % of all proposals on the shipped instances.  One THREAD per proposal computes that bound in O(1) from the
//     per-layer minima and prefix sums of the cluster-pair minima along the path (integers: the summation order is
//     free), the survivors are compacted in draw order and only they are evaluated - the first form evaluated every
//     proposal of its window with a sub-warp group each.
//   * a round looks at up to 256 draw pairs per thread block and adapts that number to the running accept rate; the
//     survivors are evaluated in passes of (warps x groups) with an early exit behind the first pass that accepts.
//   * several thread blocks (cooperative launch): block r owns the r-th slice of 256 draw pairs of the round's window.
//     The glibc stream x_i = x_{i-3} + x_{i-31} is linear, so block r jumps to its slice with one 31 x 31 matrix-vector
//     product (host-built powers of the companion matrix, mod 2^32).  Every block publishes {first accepted pair,
//     proposal / skip counts, RNG state behind it} in a slot; ONE grid barrier per round; every block then resolves the
//     same winner - the first slice with an accept - and applies the move to its own replica.  No other communication.
//   * evaluation: block descriptors are software-pipelined one round of G steps ahead and each describing lane
//     prefetches its block into L1, so the dependent chain of a MOVE / EXCHANGE no longer pays an L2 round trip per step
//   * re-scan after an accept: the next layer's table rows are loaded while the current layer is reduced; per-layer
//     minima come out of the same pass; the prefix sums of the cluster-pair minima are rebuilt by a third warp meanwhile
//   * shake (pathInsert / pathExchange) and fitDOPtoBudget run on the shared-memory state too: the working path of a
//     VNS step never exists in global memory.
#pragma once
#include "common.cuh"
#include "eval.cuh"
#include "search.cuh"
#include "chain2_tree.cuh"

#define C2_THREADS 512
#define C2_WARPS (C2_THREADS / 32)
#define C2_PAIRS 256                       // draw pairs per slice (= per thread block and round)
#define C2_RING (31 + 31 * 17 + 2)         // raw values: state (31) + 17 chunks of 31 >= 2 * C2_PAIRS draws
#define C2_INF VNS_INT_INF

struct C2Slot {            // what one slice reports per round (global memory when several blocks cooperate)
   int first_pair;         // local pair index of the first accepted proposal, -1: none
   int from, to;
   int nprop_total, nprop_upto;   // proposals (valid pairs) in the slice / up to and including the accepted one
   int nskip_total, nskip_upto;   // of which skipped by the lower bound
   int pad;
   unsigned st_acc[32];    // raw RNG state (x_{p-31} .. x_{p-1}) just behind the accepted pair's draws
   unsigned st_end[32];    // ... behind the slice's last valid pair (behind the whole slice when it has none)
};

struct C2Grid {            // null bar: single block
   unsigned long long *bar;
   C2Slot *slots;          // [2][ncta]
   const unsigned *jump;   // [ncta][31][32]: jump[r][j][t] = coefficient of s[j] in state t after 2 * C2_PAIRS * r draws
   int *gctl;              // [4] control words of the VNS loop (stop flag broadcast)
   int ncta;
};

struct C2Ctl {
   int T, nvns, L, err;
   int bad;                       // a proposal decoded to no chain (impossible by construction); block-local, so it never
                                  // steers control flow - it is folded into err when the state is stored
   int ema;                       // running proposals per accept (window sizing)
   int first[2];                  // survivor index of the first accept (atomicMin), one cell per pass parity
   int wcnt[C2_WARPS], wval[C2_WARPS], wlast[C2_WARPS], wsrv[C2_WARPS];
   int nvalid, nsurv, last_valid_pair;
   int win_from, win_to, win_found, cons_props, cons_skips;
   unsigned ring[C2_RING];
   int sfrom[C2_PAIRS], sto[C2_PAIRS], spair[C2_PAIRS], sord[C2_PAIRS];   // survivors in draw order
   unsigned newstate[32];
   LsStats st;
   long long prof[4];
};

struct C2Sm {
   int *F, *Brev;            // R x mp
   int *minF, *minBrev, *pd; // R
   int *lcl, *vns, *seg;     // 2R, 2R, alias of Brev
   const int2 *cinfo;        // C: {first sample, size}; shared memory when it fits, else global
   int2 *cinfo_sm;
   const C2Tree *tr;         // product tree of this block (chain2_tree.cuh); nlev = 0: none
};

// F, Brev: R x mpp ints each (mpp = max(mp, 2): the shake's scratch of 2R ints lives in Brev, which is rebuilt after every
// shake) - in shared memory when they fit, else in this block's slice of a global buffer (L1 / L2 resident).  Always in
// shared memory: minF, minBrev, pd R ints each; lcl, vns 2R each (the shakes insert a range before erasing one,
// vns_sop_solver.cpp:958-990).
__host__ __device__ inline size_t c2_rows_ints(int C, int mp)
{
   const size_t R = C + 3, mpp = mp < 2 ? 2 : mp;
   return 2 * R * mpp;
}
__host__ __device__ inline size_t c2_smem_bytes(int C, int mp, bool cinfo_in_smem, bool rows_in_smem)
{
   const size_t R = C + 3;
   return ((sizeof(int) * ((rows_in_smem ? c2_rows_ints(C, mp) : 0) + 7 * R) + 15) & ~(size_t)15) + (cinfo_in_smem ? sizeof(int2) * (size_t)C : 0) + 16;
}

// rows: this block's F / Brev in global memory, or null (shared)
static __device__ __forceinline__ C2Sm c2_carve(unsigned char *base, int C, int mp, const int2 *g_cinfo, bool cinfo_in_smem, int *rows)
{
   const size_t R = C + 3, mpp = mp < 2 ? 2 : mp;
   C2Sm s;
   int *small = reinterpret_cast<int *>(base);
   if (rows) s.F = rows;
   else { s.F = small; small += 2 * R * mpp; }
   s.Brev = s.F + R * mpp;
   s.minF = small;
   s.minBrev = s.minF + R;
   s.pd = s.minBrev + R;
   s.lcl = s.pd + R;
   s.vns = s.lcl + 2 * R;
   s.seg = s.Brev;        // 2R ints of shake scratch
   size_t o = (sizeof(int) * ((rows ? 0 : 2 * R * mpp) + 7 * R) + 15) & ~(size_t)15;
   s.cinfo_sm = cinfo_in_smem ? reinterpret_cast<int2 *>(base + o) : nullptr;
   s.cinfo = cinfo_in_smem ? s.cinfo_sm : g_cinfo;
   s.tr = nullptr;
   return s;
}

struct C2Env {             // per-launch constants
   InstDev I;
   const int *Dv;          // int32 table
   const int *dmin;        // C x C cluster-pair minima as int32
   const int2 *g_cinfo;    // C
   C2Grid grid;
   int cinfo_in_smem;
   int *tree_base;         // product-tree replicas, one per thread block (null: no tree)
   long long tree_stride;  // ints per replica
   int *rows_base;         // F / Brev replicas, one per thread block (null: they live in shared memory)
   long long rows_stride;
};

static __device__ __forceinline__ C2Sm c2_carve_env(unsigned char *base, const C2Env &E)
{
   return c2_carve(base, E.I.C, E.I.mp, E.g_cinfo, E.cinfo_in_smem != 0, E.rows_base ? E.rows_base + (long long)blockIdx.x * E.rows_stride : nullptr);
}

// the block's tree descriptor (thread 0 of every kernel fills it before the first barrier)
static __device__ __forceinline__ void c2_tree_init(C2Tree &tr, const C2Env &E)
{
   int nlev = 0;
   c2_tree_layout(E.I.C + 2, tr.lvl_off, &nlev);
   tr.pitch = E.I.mp * E.I.mp;
   tr.nodes = E.tree_base ? E.tree_base + (long long)blockIdx.x * E.tree_stride : nullptr;
   tr.nlev = tr.nodes ? nlev : 0;
}

// ---- grid barrier (cooperative launch: all blocks resident) ----------------------------------------------------------
static __device__ __forceinline__ void c2_grid_sync(const C2Grid &G, unsigned long long &epoch)
{
   __syncthreads();
   if (threadIdx.x == 0) {
      __threadfence();
      epoch += (unsigned long long)G.ncta;
      atomicAdd(G.bar, 1ull);
      unsigned long long v;
      do {
         asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(G.bar) : "memory");
      } while (v < epoch);
      __threadfence();
   }
   __syncthreads();
}

// ---- one DP layer by one warp on the shared-memory state (lane = sample; mp <= 32) -------------------------------------
// F[l] = F[l-1] (x) D(c_{l-1}, c_l)   vns_sop_path.cpp:1353-1373 / 1426-1448 / 1488-1511
static __device__ __forceinline__ void c2_fwd_layer(const C2Env &E, const C2Sm &s, int l, int lane)
{
   constexpr unsigned FULL = 0xffffffffu;
   const int mp = E.I.mp;
   const int2 cia = s.cinfo[s.lcl[l - 1]], cib = s.cinfo[s.lcl[l]];
   const int v = lane < cia.y ? s.F[(l - 1) * mp + lane] : C2_INF;
   const int *blk = E.Dv + ((long long)cia.x * E.I.N + (long long)cia.y * cib.x) + min(lane, cib.y - 1);
   int acc = C2_INF;
   for (int i = 0; i < cia.y; ++i) acc = __viaddmin_s32(__shfl_sync(FULL, v, i), __ldg(blk + i * cib.y), acc);
   acc = lane < cib.y ? acc : C2_INF;
   if (lane < cib.y) s.F[l * mp + lane] = acc;
   const int mn = __reduce_min_sync(FULL, acc);
   if (lane == 0) s.minF[l] = mn;
   __syncwarp();
}

// forward layers [lf, S] (lf >= 1), pipelined: the first eight rows of the NEXT layer's block are loaded while the
// current layer is reduced; per-layer minima in the same pass
static __device__ void c2_fwd_scan(const C2Env &E, const C2Sm &s, int S, int lf, int lane)
{
   constexpr unsigned FULL = 0xffffffffu;
   if (lf > S) return;
   const int mp = E.I.mp, N = E.I.N;
   const int *Dv = E.Dv;
   int2 cia = s.cinfo[s.lcl[lf - 1]], cib = s.cinfo[s.lcl[lf]];
   int v = lane < cia.y ? s.F[(lf - 1) * mp + lane] : C2_INF;
   int dd[8];
   {
      const int *blk = Dv + ((long long)cia.x * N + (long long)cia.y * cib.x) + min(lane, cib.y - 1);
#pragma unroll
      for (int t = 0; t < 8; ++t) dd[t] = __ldg(blk + min(t, cia.y - 1) * cib.y);
   }
   for (int l = lf; l <= S; ++l) {
      int2 cin = cib;
      int dn[8];
      if (l < S) {
         cin = s.cinfo[s.lcl[l + 1]];
         const int *nb = Dv + ((long long)cib.x * N + (long long)cib.y * cin.x) + min(lane, cin.y - 1);
#pragma unroll
         for (int t = 0; t < 8; ++t) dn[t] = __ldg(nb + min(t, cib.y - 1) * cin.y);
      } else {
#pragma unroll
         for (int t = 0; t < 8; ++t) dn[t] = 0;
      }
      int a0 = C2_INF, a1 = C2_INF;
#pragma unroll
      for (int t = 0; t < 8; t += 2) {
         a0 = __viaddmin_s32(__shfl_sync(FULL, v, min(t, cia.y - 1)), dd[t], a0);
         a1 = __viaddmin_s32(__shfl_sync(FULL, v, min(t + 1, cia.y - 1)), dd[t + 1], a1);
      }
      if (cia.y > 8) {
         const int *blk = Dv + ((long long)cia.x * N + (long long)cia.y * cib.x) + min(lane, cib.y - 1);
         for (int i = 8; i < cia.y; ++i) a0 = __viaddmin_s32(__shfl_sync(FULL, v, i), __ldg(blk + i * cib.y), a0);
      }
      const int acc = lane < cib.y ? min(a0, a1) : C2_INF;
      if (lane < cib.y) s.F[l * mp + lane] = acc;
      const int mn = __reduce_min_sync(FULL, acc);
      if (lane == 0) s.minF[l] = mn;
      v = acc;
      cia = cib; cib = cin;
#pragma unroll
      for (int t = 0; t < 8; ++t) dd[t] = dn[t];
   }
   __syncwarp();
}

// backward layers [lb, 0]: B[l][i] = min_j (B[l+1][j] + D(c_l, c_{l+1})[i][j])   :1380-1399 / 1453-1472 / 1515-1539
// rows are stored by distance from the goal (slot S - l)
static __device__ void c2_bwd_scan(const C2Env &E, const C2Sm &s, int S, int lb, int lane)
{
   constexpr unsigned FULL = 0xffffffffu;
   if (lb < 0) return;
   const int mp = E.I.mp, N = E.I.N;
   const int *Dv = E.Dv;
   int2 cib = s.cinfo[s.lcl[lb + 1]], cia = s.cinfo[s.lcl[lb]];
   int v = lane < cib.y ? s.Brev[(S - (lb + 1)) * mp + lane] : C2_INF;
   int dd[8];
   {
      const int *blk = Dv + ((long long)cia.x * N + (long long)cia.y * cib.x) + min(lane, cia.y - 1) * cib.y;
#pragma unroll
      for (int t = 0; t < 8; ++t) dd[t] = __ldg(blk + min(t, cib.y - 1));
   }
   for (int l = lb; l >= 0; --l) {
      int2 cip = cia;
      int dn[8];
      if (l > 0) {
         cip = s.cinfo[s.lcl[l - 1]];
         const int *nb = Dv + ((long long)cip.x * N + (long long)cip.y * cia.x) + min(lane, cip.y - 1) * cia.y;
#pragma unroll
         for (int t = 0; t < 8; ++t) dn[t] = __ldg(nb + min(t, cia.y - 1));
      } else {
#pragma unroll
         for (int t = 0; t < 8; ++t) dn[t] = 0;
      }
      int a0 = C2_INF, a1 = C2_INF;
#pragma unroll
      for (int t = 0; t < 8; t += 2) {
         a0 = __viaddmin_s32(__shfl_sync(FULL, v, min(t, cib.y - 1)), dd[t], a0);
         a1 = __viaddmin_s32(__shfl_sync(FULL, v, min(t + 1, cib.y - 1)), dd[t + 1], a1);
      }
      if (cib.y > 8) {
         const int *blk = Dv + ((long long)cia.x * N + (long long)cia.y * cib.x) + min(lane, cia.y - 1) * cib.y;
         for (int j = 8; j < cib.y; ++j) a0 = __viaddmin_s32(__shfl_sync(FULL, v, j), __ldg(blk + j), a0);
      }
      const int acc = lane < cia.y ? min(a0, a1) : C2_INF;
      if (lane < cia.y) s.Brev[(S - l) * mp + lane] = acc;
      const int mn = __reduce_min_sync(FULL, acc);
      if (lane == 0) s.minBrev[S - l] = mn;
      v = acc;
      cib = cia; cia = cip;
#pragma unroll
      for (int t = 0; t < 8; ++t) dd[t] = dn[t];
   }
   __syncwarp();
}

// pd[l] = sum_{l' < l} dmin(c_l', c_l'+1), l = 0 .. S  (one warp, chunks of 32 layers)
static __device__ void c2_pd_scan(const C2Env &E, const C2Sm &s, int S, int lane)
{
   constexpr unsigned FULL = 0xffffffffu;
   int carry = 0;
   for (int l0 = 0; l0 <= S; l0 += 32) {
      const int l = l0 + lane;
      int x = 0;   // x = dmin of edge (l-1 -> l)
      if (l >= 1 && l <= S) x = __ldg(E.dmin + (size_t)s.lcl[l - 1] * E.I.C + s.lcl[l]);
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

// F / B rows through the product tree, ONE tree step per layer.  Forward: F[l] = F[l - 2^z] (x) node(z, .), z = ctz(l) -
// the row it starts from is either below lf (still valid) or has a larger ctz, so the layers are processed in phases
// z = nlev .. 0, all layers of a phase in parallel (one sub-warp group each).  Backward: B[l] = node(z, l >> z) (x)
// B[l + 2^z] in the same phases, except for the at most nlev + 1 layers whose aligned node would overhang the goal
// (l + 2^z > S: the boundaries p_0 = 0 < p_1 < .. of the binary expansion of S); those are a sequential chain
// B[p_i] = node (x) B[p_{i+1}] done first.  Work: S tree steps instead of S x 2 log2(S); depth nlev + 1 barriers.
// Warps 0 .. W/2-1 take the forward layers, the others the backward ones (the last warp first rebuilds the prefix sums
// of the cluster-pair minima).
template <int G, int NR, bool TR>
static __device__ __forceinline__ void c2_row_step(const C2Env &E, const C2Sm &s, int S, bool on, int l, int z, int gl)
{
   // forward (TR = false): row l from row l - 2^z;  backward (TR = true): row l from row l + 2^z
   constexpr unsigned FULL = 0xffffffffu;
   const int mp = E.I.mp;
   const int src = TR ? l + (1 << z) : l - (1 << z), nd = (TR ? l : src) >> z;
   int v[NR], mo = 0;
   const int m0 = on ? s.cinfo[s.lcl[src]].y : 0;
   const int *row = TR ? s.Brev + (S - src) * mp : s.F + src * mp;
#pragma unroll
   for (int r = 0; r < NR; ++r) v[r] = gl + r * G < m0 ? row[gl + r * G] : C2_INF;
   c2_chain<G, NR, TR>(on ? 1 : 0, gl, [&](int) { return c2_tree_mat(*s.tr, E.Dv, E.I.N, s.cinfo, s.lcl, z, nd); }, v, mo);
   int mn = C2_INF;
   int *out = TR ? s.Brev + (S - l) * mp : s.F + l * mp;
#pragma unroll
   for (int r = 0; r < NR; ++r) {
      const int j = gl + r * G;
      if (on && j < mo) { out[j] = v[r]; mn = min(mn, v[r]); }
   }
#pragma unroll
   for (int sft = G / 2; sft > 0; sft >>= 1) mn = min(mn, __shfl_xor_sync(FULL, mn, sft, G));
   if (on && gl == 0) { if (TR) s.minBrev[S - l] = mn; else s.minF[l] = mn; }
}

template <int G, int NR>
static __device__ void c2_rows_by_tree(const C2Env &E, const C2Sm &s, int S, int lf, int lb)
{
   constexpr int NG = 32 / G, HW = C2_WARPS / 2;
   const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, gl = lane % G, gq = lane / G;
   const int nlev = s.tr->nlev;
   if (warp == C2_WARPS - 1) c2_pd_scan(E, s, S, lane);
   if (warp == HW) {
      // overhanging layers p_i <= lb, nearest to the goal first: p_i = S with its lowest i set bits cleared
      int hi = S;                                   // p_{i+1}
      while (hi > 0) {
         const int z = __ffs(hi) - 1, p = hi - (1 << z);
         if (p <= lb) { c2_row_step<G, NR, true>(E, s, S, gq == 0, p, z, gl); __syncwarp(); }
         hi = p;
      }
   }
   __syncthreads();
   for (int z = nlev; z >= 0; --z) {
      // layers of this phase: odd multiples of 2^z
      if (warp < HW) {
         int q0 = (lf + (1 << z) - 1) >> z;        // first multiple of 2^z >= lf
         q0 |= 1;
         const int cnt = (q0 << z) <= S ? (((S >> z) - q0) >> 1) + 1 : 0;
         for (int base = warp * NG; base < cnt; base += HW * NG) {
            const int t = base + gq;
            c2_row_step<G, NR, false>(E, s, S, t < cnt, (q0 + 2 * t) << z, z, gl);
         }
      } else {
         // l = q 2^z, q odd, l <= lb, l + 2^z <= S
         const int qmax = min(lb >> z, (S >> z) - 1);
         const int cnt = qmax >= 1 ? ((qmax - 1) >> 1) + 1 : 0;
         for (int base = (warp - HW) * NG; base < cnt; base += (C2_WARPS - HW) * NG) {
            const int t = base + gq;
            c2_row_step<G, NR, true>(E, s, S, t < cnt, (1 + 2 * t) << z, z, gl);
         }
      }
      __syncthreads();
   }
}

// all threads.  Changed edges of the path: [e1lo, e1hi) and [e2lo, e2hi) (product tree).  Without a tree: forward layers
// [lf, S] on warp 0, backward [lb, 0] on warp 1, prefix sums on warp 2
template <int G, int NR>
static __device__ void c2_rescan(const C2Env &E, const C2Sm &s, C2Ctl &ctl, int T, int lf, int lb, int e1lo, int e1hi, int e2lo, int e2hi)
{
   const int S = T + 1;
   const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
   __syncthreads();
   if (s.tr->nlev > 0) {
      c2_tree_rebuild(*s.tr, E.Dv, E.I.N, s.cinfo, s.lcl, S, e1lo, min(e1hi, S), e2lo, min(e2hi, S));
      c2_rows_by_tree<G, NR>(E, s, S, lf, lb);
   } else {
      if (warp == 0) c2_fwd_scan(E, s, S, lf, lane);
      else if (warp == 1) c2_bwd_scan(E, s, S, lb, lane);
      else if (warp == 2) c2_pd_scan(E, s, S, lane);
   }
   __syncthreads();
   if (threadIdx.x == 0) {
      ctl.L = s.minF[S];                                                   // getPathLength :1660-1678
      if (s.minF[S] != s.minBrev[S] && !ctl.err) ctl.err = 4;              // addPoint's check :80-88 (integers: exact)
   }
   __syncthreads();
}

// arr[lo .. hi] moves by one position (delta = +1 / -1); all threads call
static __device__ __forceinline__ void c2_shift1(int *arr, int lo, int hi, int delta)
{
   if (delta > 0) {
      for (int top = hi; top >= lo; top -= C2_THREADS) {       // chunks from the top
         const int k = top - (int)threadIdx.x;
         int v = 0;
         if (k >= lo) v = arr[k];
         __syncthreads();
         if (k >= lo) arr[k + 1] = v;
         __syncthreads();
      }
   } else {
      for (int bot = lo; bot <= hi; bot += C2_THREADS) {
         const int k = bot + (int)threadIdx.x;
         int v = 0;
         if (k <= hi) v = arr[k];
         __syncthreads();
         if (k <= hi) arr[k - 1] = v;
         __syncthreads();
      }
   }
}

// insertRandom accept (vns_sop_solver.cpp:621-646): VNS[from] is re-inserted immediately before the element that was
// at index `to`; only the final state matters.  tg = lcl + 1.
template <int G, int NR>
static __device__ void c2_apply_insert(const C2Env &E, const C2Sm &s, C2Ctl &ctl, int from, int to)
{
   const int T = ctl.T;
   int *tg = s.lcl + 1;
   const int c = s.vns[from];
   __syncthreads();
   if (from < T) {
      if (to > from) {        // travels up to index to-1; layers from+1 .. to change
         c2_shift1(s.vns, from + 1, to - 1, -1);
         c2_shift1(tg, from + 1, to - 1, -1);
         if (threadIdx.x == 0) { s.vns[to - 1] = c; tg[to - 1] = c; }
         c2_rescan<G, NR>(E, s, ctl, T, from + 1, to, from, to + 1, 0, 0);
      } else {                // travels down to index `to`; layers to+1 .. from+1 change
         c2_shift1(s.vns, to, from - 1, +1);
         c2_shift1(tg, to, from - 1, +1);
         if (threadIdx.x == 0) { s.vns[to] = c; tg[to] = c; }
         c2_rescan<G, NR>(E, s, ctl, T, to + 1, from + 1, to, from + 2, 0, 0);
      }
   } else {                   // from outside the path (from >= T >= to): one more target at index `to`
      c2_shift1(s.vns, to, from - 1, +1);
      c2_shift1(tg, to, T, +1);               // includes the goal's entry tg[T] = lcl[T+1]
      if (threadIdx.x == 0) { s.vns[to] = c; tg[to] = c; ctl.T = T + 1; }
      // B rows of the old layers to+1 .. S keep their slots (distance from the goal is unchanged)
      c2_rescan<G, NR>(E, s, ctl, T + 1, to + 1, to + 1, to, T + 2, 0, 0);   // every edge behind the new target moves
   }
}

// exchangeRandom accept (:807-832): VNS[from] <-> VNS[to]; `to` is always in the path
template <int G, int NR>
static __device__ void c2_apply_exchange(const C2Env &E, const C2Sm &s, C2Ctl &ctl, int from, int to)
{
   const int T = ctl.T;
   int *tg = s.lcl + 1;
   __syncthreads();
   if (threadIdx.x == 0) {
      const int cf = s.vns[from], ct = s.vns[to];
      s.vns[to] = cf; s.vns[from] = ct;
      tg[to] = cf;
      if (from < T) tg[from] = ct;
   }
   const int lo = (from < T && from < to) ? from : to;
   const int hi = (from < T && from > to) ? from : to;
   c2_rescan<G, NR>(E, s, ctl, T, lo + 1, hi + 1, lo, lo + 2, hi, hi + 2);   // layers lo + 1 and hi + 1 changed
}

// ---- glibc rand() stream (math_common.cpp:21-28; random_r TYPE_3: x_i = x_{i-3} + x_{i-31}, output >> 1) ----------------
// ring[0..31) = the state (oldest first), ring[31 + k] = draw k of the slice.  Warp 0 only.
static __device__ __forceinline__ void c2_ring_fill(C2Ctl &ctl, int ndraws, int lane)
{
   constexpr unsigned FULL = 0xffffffffu;
   for (int gen = 31; gen < 31 + ndraws; gen += 31) {
      unsigned a = 0;
      if (lane < 31) a = ctl.ring[gen - 31 + lane];
      if (lane < 3) a += ctl.ring[gen - 3 + lane];
      // x_{gen+k} = a_k + x_{gen+k-3}: inclusive scan with stride 3
#pragma unroll
      for (int d = 3; d < 32; d <<= 1) {
         const unsigned t = __shfl_up_sync(FULL, a, d);
         if (lane >= d) a += t;
      }
      if (lane < 31) ctl.ring[gen + lane] = a;
      __syncwarp();
   }
}

// state of the stream 2 * C2_PAIRS * r draws ahead: one matrix-vector product mod 2^32 (r >= 1)
static __device__ __forceinline__ void c2_ring_jump(C2Ctl &ctl, const RngState &g, const unsigned *__restrict__ jump_r, int lane)
{
   unsigned acc = 0;
   if (lane < 31) {
#pragma unroll 4
      for (int j = 0; j < 31; ++j) acc += __ldg(jump_r + j * 32 + lane) * (unsigned)g.r[(g.f + j) % 31];
      ctl.ring[lane] = acc;
   }
   __syncwarp();
}

// ---- chain of a proposal: up to two explicit blocks, a run of path edges [ra, rb) walked through the product tree, up to
//      two explicit blocks; closed against B[sK].  Layers; -1 = the foreign cluster (ADD / REPLACE) ------------------------
struct C2Code {
   int K, s0, sK, foreign;
   int npre, npost, ra, rb, rcnt;
   int pa0, pb0, pa1, pb1, qa0, qb0, qa1, qb1;
};

// insertRandom: in-path -> tryToMove(from, to) (vns_sop_path.cpp:932-1154), else tryToAdd(VNS[from], to) (:245-304);
// exchangeRandom: in-path -> tryToExchange(from, to) (:635-797), else tryToReplace(VNS[from], to) (:456-513)
static __device__ __forceinline__ C2Code c2_decode(int nlev, int T, int kind, bool in, int from, int to, int cf)
{
   C2Code c;
   c.K = 0; c.s0 = c.sK = 0; c.foreign = cf; c.npre = c.npost = 0; c.ra = c.rb = 0; c.rcnt = 0;
   c.pa0 = c.pb0 = c.pa1 = c.pb1 = c.qa0 = c.qb0 = c.qa1 = c.qb1 = 0;
   if (!in) {
      c.s0 = to; c.sK = kind == 1 ? to + 1 : to + 2;
      c.npre = 2; c.pa0 = to; c.pb0 = -1; c.pa1 = -1; c.pb1 = c.sK;
   } else if (kind == 1) {
      const int i = min(from, to) + 1, j = max(from, to) + 1;
      if (to > from) {   // up:   i-1, i+1 .. j-1, i, j
         c.s0 = i - 1; c.sK = j;
         c.npre = 1; c.pa0 = i - 1; c.pb0 = i + 1;
         c.ra = i + 1; c.rb = j - 1;
         c.npost = 2; c.qa0 = j - 1; c.qb0 = i; c.qa1 = i; c.qb1 = j;
      } else {           // down: i-1, j, i .. j-1, j+1
         c.s0 = i - 1; c.sK = j + 1;
         c.npre = 2; c.pa0 = i - 1; c.pb0 = j; c.pa1 = j; c.pb1 = i;
         c.ra = i; c.rb = j - 1;
         c.npost = 1; c.qa0 = j - 1; c.qb0 = j + 1;
      }
   } else {              // i, j, i+2 .. j-1, i+1, j+1
      const int i = min(from, to), j = max(from, to) + 1;
      c.s0 = i; c.sK = j + 1;
      c.npre = 2; c.pa0 = i; c.pb0 = j; c.pa1 = j; c.pb1 = i + 2;
      c.ra = i + 2; c.rb = j - 1;
      c.npost = 2; c.qa0 = j - 1; c.qb0 = i + 1; c.qa1 = i + 1; c.qb1 = j + 1;
   }
   int l, n;
   c.rcnt = c.rb > c.ra ? c2_decomp(nlev, c.ra, c.rb, -1, &l, &n) : 0;
   c.K = c.npre + c.rcnt + c.npost;
   return c;
}

// block k (1 <= k <= K) of the chain
static __device__ __forceinline__ C2Mat c2_code_block(const C2Env &E, const C2Sm &s, const C2Code &c, int k)
{
   int la, lb;
   if (k <= c.npre) { la = k == 1 ? c.pa0 : c.pa1; lb = k == 1 ? c.pb0 : c.pb1; }
   else if (k <= c.npre + c.rcnt) {
      int l = 0, n = 0;
      c2_decomp(s.tr->nlev, c.ra, c.rb, k - c.npre - 1, &l, &n);
      return c2_tree_mat(*s.tr, E.Dv, E.I.N, s.cinfo, s.lcl, l, n);
   } else {
      const int q = k - c.npre - c.rcnt;
      la = q == 1 ? c.qa0 : c.qa1; lb = q == 1 ? c.qb0 : c.qb1;
   }
   const int2 cia = s.cinfo[la < 0 ? c.foreign : s.lcl[la]], cib = s.cinfo[lb < 0 ? c.foreign : s.lcl[lb]];
   C2Mat m;
   m.p = E.Dv + ((long long)cia.x * E.I.N + (long long)cia.y * cib.x);
   m.ma = cia.y; m.mb = cib.y;
   return m;
}

// ---- lower bound of a proposal in O(1) (one thread): minF(s0) + dist_between + minB(sK) --------------------------------
// tryToMoveLowerBound vns_sop_path.cpp:799-930, tryToAddLowerBound :207-243, tryToExchangeLowerBound :573-633,
// tryToReplaceLowerBound :423-454; integers, so the reference's left-to-right sums equal these prefix differences
static __device__ __forceinline__ int c2_lower_bound(const C2Env &E, const C2Sm &s, int T, int kind, int from, int to)
{
   const int S = T + 1, C = E.I.C;
   const int *lcl = s.lcl;
   auto dm = [&](int ca, int cb) { return __ldg(E.dmin + (size_t)ca * C + cb); };
   const bool in = from < T;
   if (kind == 1) {
      if (in) {
         const int i = min(from, to) + 1, j = max(from, to) + 1;
         if (to > from)   // up: i-1, i+1 .. j-1, i, j
            return s.minF[i - 1] + dm(lcl[i - 1], lcl[i + 1]) + (s.pd[j - 1] - s.pd[i + 1]) + dm(lcl[j - 1], lcl[i]) + dm(lcl[i], lcl[j]) +
                   s.minBrev[S - j];
         // down: i-1, j, i .. j-1, j+1
         return s.minF[i - 1] + dm(lcl[i - 1], lcl[j]) + dm(lcl[j], lcl[i]) + (s.pd[j - 1] - s.pd[i]) + dm(lcl[j - 1], lcl[j + 1]) +
                s.minBrev[S - (j + 1)];
      }
      const int c = s.vns[from];
      return s.minF[to] + dm(lcl[to], c) + dm(c, lcl[to + 1]) + s.minBrev[S - (to + 1)];
   }
   if (in) {   // i, j, i+2 .. j-1, i+1, j+1
      const int i = min(from, to), j = max(from, to) + 1;
      return s.minF[i] + dm(lcl[i], lcl[j]) + dm(lcl[j], lcl[i + 2]) + (s.pd[j - 1] - s.pd[i + 2]) + dm(lcl[j - 1], lcl[i + 1]) +
             dm(lcl[i + 1], lcl[j + 1]) + s.minBrev[S - (j + 1)];
   }
   const int c = s.vns[from];
   return s.minF[to] + dm(lcl[to], c) + dm(c, lcl[to + 2]) + s.minBrev[S - (to + 2)];
}

// ---- the local search ------------------------------------------------------------------------------------------------
// g: RNG state in canonical form at entry and exit (all blocks hold the same).  Every block runs the same control flow
// on its replica; only the evaluation of a round's window is divided.
template <int G, int NR>
static __device__ void c2_local_search(const C2Env &E, const C2Sm &s, C2Ctl &ctl, RngState &g, int kind, int num_changes, double budget,
                                       int skipflag, unsigned long long &epoch, unsigned &round_no)
{
   constexpr int NG = 32 / G, NGRP = C2_WARPS * NG;
   constexpr unsigned FULL = 0xffffffffu;
   const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, gl = lane % G, gq = lane / G;
   const int nvns = ctl.nvns, mp = E.I.mp;
   const int gap = kind == 1 ? 1 : 2;
   const bool multi = E.grid.bar != nullptr;
   const int ncta = multi ? E.grid.ncta : 1, rank = multi ? (int)blockIdx.x : 0;
   __shared__ C2Slot myslot;
   __syncthreads();
   const bool enabled = kind == 1 ? (nvns >= 3) : (nvns >= 4 && ctl.T > 0);   // guards :498 and :711
   int var = 0, empty_rounds = 0;
   while (enabled && var < num_changes) {
      const int T = ctl.T, S = T + 1;
      const int remaining = num_changes - var;
      const long long t0 = clock64();
      // ---- window of this round (a function of replicated state: the same in every block) -------------------------
      const int target = min(max(2 * ctl.ema, 32), C2_PAIRS * ncta);
      int nactive = 1, P = C2_PAIRS;
      if (target <= C2_PAIRS || remaining < 2 * C2_PAIRS) {
         P = 32;
         while (P < target && P < C2_PAIRS) P <<= 1;
      } else
         nactive = min(min(ncta, (target + C2_PAIRS - 1) / C2_PAIRS), remaining / C2_PAIRS);
      const bool active = rank < nactive;
      // ---- 1. draws of this block's slice -> proposals -> lower-bound filter -> survivors in draw order ---------------
      if (active && warp == 0) {
         if (rank == 0) {
            if (lane < 31) ctl.ring[lane] = (unsigned)g.r[(g.f + lane) % 31];
            __syncwarp();
         } else
            c2_ring_jump(ctl, g, E.grid.jump + (size_t)rank * 31 * 32, lane);
         c2_ring_fill(ctl, 2 * P, lane);
      }
      __syncthreads();
      if (active) {
         const int p = threadIdx.x;
         bool valid = false, surv = false;
         int from = 0, to = 0;
         if (p < P) {
            const unsigned r1 = ctl.ring[31 + 2 * p] >> 1, r2 = ctl.ring[31 + 2 * p + 1] >> 1;
            const int tohi = kind == 1 ? T : T - 1;
            from = (int)(((unsigned long long)r1 * (unsigned)nvns) >> 31);            // :507 / :723
            to = (int)(((unsigned long long)r2 * (unsigned)(tohi + 1)) >> 31);        // :508 / :724
            valid = abs(from - to) > gap;                                             // :510-519 / :726-730
         }
         const unsigned balv = __ballot_sync(FULL, valid);
         if (lane == 0) ctl.wcnt[warp] = __popc(balv);
         __syncthreads();
         int ord = __popc(balv & ((1u << lane) - 1));
         for (int w = 0; w < warp; ++w) ord += ctl.wcnt[w];
         // the call ends after exactly num_changes proposals: later ones do not exist
         valid = valid && ord < remaining;
         if (valid) {
            surv = true;
            if (skipflag) surv = !((double)c2_lower_bound(E, s, T, kind, from, to) > budget);   // :539,:574,:749,:779
         }
         const unsigned bals = __ballot_sync(FULL, surv), balv2 = __ballot_sync(FULL, valid);
         if (lane == 0) { ctl.wsrv[warp] = __popc(bals); ctl.wval[warp] = __popc(balv2); ctl.wlast[warp] = 31 - __clz(balv2); }
         __syncthreads();
         int si = __popc(bals & ((1u << lane) - 1));
         for (int w = 0; w < warp; ++w) si += ctl.wsrv[w];
         if (surv) { ctl.sfrom[si] = from; ctl.sto[si] = to; ctl.spair[si] = p; ctl.sord[si] = ord; }
         if (threadIdx.x == 0) {
            int nv = 0, ns = 0, lastp = -1;
            for (int w = 0; w < C2_WARPS; ++w) {
               const int cnt = ctl.wval[w];
               nv += cnt; ns += ctl.wsrv[w];
               if (cnt) lastp = w * 32 + ctl.wlast[w];
            }
            ctl.nvalid = nv; ctl.nsurv = ns; ctl.last_valid_pair = lastp; ctl.first[0] = ctl.first[1] = 0x7fffffff;
         }
      } else if (threadIdx.x == 0) { ctl.nvalid = 0; ctl.nsurv = 0; ctl.last_valid_pair = -1; ctl.first[0] = ctl.first[1] = 0x7fffffff; }
      __syncthreads();
      const long long t1 = clock64();
      // ---- 2. evaluate the survivors in draw order, one sub-warp group each; stop behind the first pass that accepts -------
      const int nsurv = ctl.nsurv;
      for (int base = 0, pass = 0; base < nsurv; base += NGRP, ++pass) {
         if (base + warp * NG < nsurv) {
            const int sidx = base + warp * NG + gq;
            const bool act = sidx < nsurv;
            const int from = act ? ctl.sfrom[sidx] : 0, to = act ? ctl.sto[sidx] : 0;
            const bool in = from < T;
            const int cf = s.vns[from];
            const C2Code cc = act ? c2_decode(s.tr->nlev, T, kind, in, from, to, cf) : C2Code{};
            const int K = cc.K;
            int v[NR];
            {
               const int m0 = K ? s.cinfo[s.lcl[cc.s0]].y : 0;
#pragma unroll
               for (int r = 0; r < NR; ++r) v[r] = gl + r * G < m0 ? s.F[cc.s0 * mp + gl + r * G] : C2_INF;
            }
            int mK = 0;
            c2_chain<G, NR, false>(K, gl, [&](int k) { return c2_code_block(E, s, cc, k); }, v, mK);
            int fin = C2_INF;
#pragma unroll
            for (int r = 0; r < NR; ++r) fin = min(fin, (K && gl + r * G < mK) ? v[r] + s.Brev[(S - cc.sK) * mp + gl + r * G] : C2_INF);
#pragma unroll
            for (int sft = G / 2; sft > 0; sft >>= 1) fin = min(fin, __shfl_xor_sync(FULL, fin, sft, G));
            if (act && gl == 0) {
               const double L = (double)ctl.L;
               double dR = 0.0;
               if (!in) dR = kind == 1 ? __ldg(E.I.reward + cf)                                          // :581
                                       : 0.0 + (__ldg(E.I.reward + cf) - __ldg(E.I.reward + s.vns[to]));   // :784
               const double len = (double)fin;     // integers: (L - 0.0) + (len - L) == len  (:580,:583)
               const bool acc = K > 0 && len <= budget && (dR > 0 || (L - len) > VNS_MIN_CHANGE_EPS);       // :619,:805-806
               if (acc) atomicMin(&ctl.first[pass & 1], sidx);
               if (K == 0) ctl.bad = 1;
            }
         }
         __syncthreads();
         // the cell of this pass's parity is next written two passes on, behind another barrier: no reader is still here
         if (ctl.first[pass & 1] != 0x7fffffff) break;
      }
      const long long t2 = clock64();
      // ---- 3. this slice's report; all slices -> the round's winner -----------------------------------------------------
      if (warp == 0) {
         const int fs = min(ctl.first[0], ctl.first[1]);
         const bool found = fs != 0x7fffffff;
         const int fp = found ? ctl.spair[fs] : -1;
         if (lane == 0) {
            myslot.first_pair = fp;
            myslot.from = found ? ctl.sfrom[fs] : 0; myslot.to = found ? ctl.sto[fs] : 0;
            myslot.nprop_total = ctl.nvalid; myslot.nskip_total = ctl.nvalid - ctl.nsurv;
            myslot.nprop_upto = found ? ctl.sord[fs] + 1 : 0; myslot.nskip_upto = found ? ctl.sord[fs] - fs : 0;
         }
         // the state behind pair x (draws 0 .. 2x+1 consumed) is ring[2 (x + 1) .. 2 (x + 1) + 31)
         const int endp = ctl.last_valid_pair >= 0 ? ctl.last_valid_pair : P - 1;
         if (lane < 31) {
            myslot.st_acc[lane] = found ? ctl.ring[2 * (fp + 1) + lane] : 0u;
            myslot.st_end[lane] = active ? ctl.ring[2 * (endp + 1) + lane] : 0u;
         }
         __syncwarp();
      }
      __syncthreads();
      const C2Slot *slots = &myslot;
      if (multi) {
         if (active) {
            C2Slot *dst = E.grid.slots + (size_t)(round_no & 1) * ncta + rank;
            for (int k = threadIdx.x; k < (int)(sizeof(C2Slot) / 4); k += blockDim.x) reinterpret_cast<int *>(dst)[k] = reinterpret_cast<const int *>(&myslot)[k];
         }
         c2_grid_sync(E.grid, epoch);
         slots = E.grid.slots + (size_t)(round_no & 1) * ncta;
      }
      ++round_no;
      if (warp == 0) {
         // first slice with an accept wins; proposals / skips consumed = whole slices before it + its own part
         int winner = -1, props = 0, skips = 0, lastv = -1;
         for (int q0 = 0; q0 < nactive && winner < 0; q0 += 32) {
            const int q = q0 + lane;
            int fp = -1, npt = 0, nst = 0;
            if (q < nactive) {
               const int *sl = reinterpret_cast<const int *>(slots + q);
               fp = multi ? __ldcg(sl + 0) : sl[0];
               npt = multi ? __ldcg(sl + 3) : sl[3];
               nst = multi ? __ldcg(sl + 5) : sl[5];
            }
            const unsigned hit = __ballot_sync(FULL, fp >= 0);
            const int upto = hit ? __ffs(hit) - 1 : 32;          // lanes below `upto` are whole slices
            const unsigned hasv = __ballot_sync(FULL, npt > 0) & (upto >= 32 ? FULL : ((1u << upto) - 1));
            int pp = lane < upto ? npt : 0, ss = lane < upto ? nst : 0;
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) { pp += __shfl_xor_sync(FULL, pp, d); ss += __shfl_xor_sync(FULL, ss, d); }
            props += pp; skips += ss;
            if (hasv) lastv = q0 + 31 - __clz(hasv);
            if (hit) winner = q0 + upto;
         }
         const C2Slot *ws = slots + (winner >= 0 ? winner : (lastv >= 0 ? lastv : nactive - 1));
         const int *wi = reinterpret_cast<const int *>(ws);
         auto rd = [&](int k) { return multi ? __ldcg(wi + k) : wi[k]; };
         if (winner >= 0) { props += rd(4); skips += rd(6); }
         if (lane < 31) ctl.newstate[lane] = (unsigned)rd((winner >= 0 ? 8 : 40) + lane);
         if (lane == 0) {
            ctl.win_found = winner >= 0; ctl.win_from = rd(1); ctl.win_to = rd(2);
            ctl.cons_props = props; ctl.cons_skips = skips;
         }
         __syncwarp();
         if (lane < 31) g.r[lane] = (int32_t)ctl.newstate[lane];
         if (lane == 0) {
            g.f = 0; g.b = 28;
            ctl.st.proposals += props; ctl.st.lb_skips += skips; ctl.st.evals += props - skips; ctl.st.rounds += 1;
            if (winner >= 0) { ctl.st.accepts += 1; ctl.ema = (3 * ctl.ema + props + 3) / 4; }
            else ctl.ema = min(2 * max(ctl.ema, P * nactive / 2), 1 << 20);
         }
      }
      __syncthreads();
      const int used = ctl.cons_props;
      var += used;
      const long long t3 = clock64();
      empty_rounds = used ? 0 : empty_rounds + 1;
      if (empty_rounds > VNS_MAX_DRAWS / 64 && threadIdx.x == 0) ctl.err = 3;
      if (ctl.win_found) {
         if (kind == 1) c2_apply_insert<G, NR>(E, s, ctl, ctl.win_from, ctl.win_to);
         else c2_apply_exchange<G, NR>(E, s, ctl, ctl.win_from, ctl.win_to);
      }
      __syncthreads();
      if (threadIdx.x == 0) {
         ctl.prof[0] += t1 - t0; ctl.prof[1] += t2 - t1; ctl.prof[2] += t3 - t2; ctl.prof[3] += clock64() - t3;
      }
      if (ctl.err) break;
   }
   __syncthreads();
}

// ---- global <-> shared path state ------------------------------------------------------------------------------------------
// cg: read through L2 (another block wrote the path)
static __device__ void c2_load(const C2Env &E, const PathDev &P, const C2Sm &s, C2Ctl &ctl, bool cg)
{
   const InstDev &I = E.I;
   auto ldi = [&](const int *p) { return cg ? __ldcg(p) : *p; };
   auto ldd = [&](const double *p) { return cg ? __ldcg(p) : *p; };
   const int T = ldi(P.hdr + HDR_T), nvns = ldi(P.hdr + HDR_NVNS), S = T + 1, mp = I.mp;
   __syncthreads();
   if (s.cinfo_sm)
      for (int c = threadIdx.x; c < I.C; c += blockDim.x) s.cinfo_sm[c] = make_int2(I.off[c], I.m[c]);
   for (int l = threadIdx.x; l <= S; l += blockDim.x) {
      s.lcl[l] = l == 0 ? I.start : (l == S ? I.goal : ldi(P.tg + l - 1));
      s.minF[l] = (int)ldd(P.minF + l);
      s.minBrev[S - l] = (int)ldd(P.minB + l);
   }
   for (int k = threadIdx.x; k < nvns; k += blockDim.x) s.vns[k] = ldi(P.vns + k);
   for (int e = threadIdx.x; e < (S + 1) * mp; e += blockDim.x) {
      const int l = e / mp, j = e - l * mp;
      s.F[e] = (int)ldd(P.F + e);
      s.Brev[(S - l) * mp + j] = (int)ldd(P.B + e);
   }
   if (threadIdx.x == 0) { ctl.T = T; ctl.nvns = nvns; ctl.L = (int)ldd(P.hd + HD_LEN); ctl.err = 0; ctl.bad = 0; }
   __syncthreads();
   if ((threadIdx.x >> 5) == 2) c2_pd_scan(E, s, S, threadIdx.x & 31);
   __syncthreads();
   if (s.tr->nlev > 0) c2_tree_rebuild(*s.tr, E.Dv, E.I.N, s.cinfo, s.lcl, S, 0, S, 0, 0);
}

// entries beyond a layer's cluster size are never read by anyone and are not written back
static __device__ void c2_store(const C2Env &E, const PathDev &P, const C2Sm &s, C2Ctl &ctl)
{
   const InstDev &I = E.I;
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
      __threadfence();
   }
   __syncthreads();
}

// getReward (vns_sop_path.cpp:2016-2031) of the shared-memory state, one thread
static __device__ double c2_reward(const C2Env &E, const C2Sm &s, int T)
{
   double r = 0;
   for (int t = 0; t < T; ++t) r += __ldg(E.I.reward + s.lcl[t + 1]);
   const double rs = __ldg(E.I.reward + E.I.start);
   if (rs > 0) {
      r += rs;
      if (E.I.start != E.I.goal) r += __ldg(E.I.reward + E.I.goal);
   }
   return r;
}

// ---- shake + fitDOPtoBudget on the shared-memory state ---------------------------------------------------------------------
// pathInsert(.., 1) (kind 1, vns_sop_solver.cpp:931-999) / pathExchange(.., 1) (kind 2, :1001-1089): the vector splices of
// search.cuh's cta_shake (thread 0; O(|VNS|) integer moves in shared memory), then fitDOPtoBudget (:364-385) as forward
// layers only: tryToAdd(VNS[T], T) is the goal-row minimum of the path with VNS[T] appended (integers: exact).
static __device__ void c2_shake(const C2Env &E, const C2Sm &s, C2Ctl &ctl, int kind, double budget, RngState &rng)
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
      tg[T] = E.I.goal;
      ctl.T = T; s_first = first;
   }
   __syncthreads();
   if (warp == 0) {
      int T = ctl.T;
      const int nvns = ctl.nvns, goal = E.I.goal;
      int *tg = s.lcl + 1;
      // forward DP of the spliced path (layers below the first changed target are still valid)
      c2_fwd_scan(E, s, T + 1, min(s_first + 1, T + 1), lane);
      // :366-368  pop the last target while the path is over budget
      while ((double)s.minF[T + 1] > budget && T > 0) {
         --T;
         if (lane == 0) tg[T] = goal;
         __syncwarp();
         c2_fwd_layer(E, s, T + 1, lane);
      }
      // :372-383  append VNS[T] while tryToAdd(VNS[T], T) fits
      while (T < nvns) {
         if (lane == 0) { tg[T] = s.vns[T]; tg[T + 1] = goal; }
         __syncwarp();
         c2_fwd_layer(E, s, T + 1, lane);
         c2_fwd_layer(E, s, T + 2, lane);
         if ((double)s.minF[T + 2] <= budget) { ++T; continue; }
         if (lane == 0) tg[T] = goal;
         __syncwarp();
         c2_fwd_layer(E, s, T + 1, lane);
         break;
      }
      if (lane == 0) ctl.T = T;
      __syncwarp();
   }
   __syncthreads();
   {  // full backward scan of the fitted path
      const int T = ctl.T, S = T + 1, mS = s.cinfo[E.I.goal].y;
      if (warp == 1) {
         if (lane < mS) s.Brev[lane] = 0;
         if (lane == 0) s.minBrev[0] = 0;
         __syncwarp();
         c2_bwd_scan(E, s, S, S - 1, lane);
      } else if (warp == 2)
         c2_pd_scan(E, s, S, lane);
      __syncthreads();
      if (threadIdx.x == 0) {
         ctl.L = s.minF[S];
         if (s.minF[S] != s.minBrev[S] && !ctl.err) ctl.err = 4;
      }
      __syncthreads();
      if (s.tr->nlev > 0) c2_tree_rebuild(*s.tr, E.Dv, E.I.N, s.cinfo, s.lcl, S, 0, S, 0, 0);
   }
}

// ---- kernels ----------------------------------------------------------------------------------------------------------------
template <int G, int NR>
__global__ void __launch_bounds__(C2_THREADS) k_chain2_ls(C2Env E, PathDev P, int kind, int num_changes, double budget, int skipflag,
                                                          RngState *rng, LsStats *stats, long long *prof)
{
   extern __shared__ __align__(16) unsigned char smem_c[];
   __shared__ C2Ctl ctl;
   __shared__ RngState g;
   __shared__ C2Tree tree;
   C2Sm s = c2_carve_env(smem_c, E);
   s.tr = &tree;
   if (threadIdx.x == 0) {
      c2_tree_init(tree, E);
      g = *rng;
      ctl.st = LsStats{0, 0, 0, 0, 0}; ctl.prof[0] = ctl.prof[1] = ctl.prof[2] = ctl.prof[3] = 0; ctl.ema = 16;
   }
   __syncthreads();
   unsigned long long epoch = 0;
   unsigned round_no = 0;
   c2_load(E, P, s, ctl, false);
   c2_local_search<G, NR>(E, s, ctl, g, kind, num_changes, budget, skipflag, epoch, round_no);
   // every block has read the path before block 0 overwrites it: the blocks left the last round's barrier after c2_load
   if (E.grid.bar != nullptr) c2_grid_sync(E.grid, epoch);
   if (blockIdx.x == 0) {
      c2_store(E, P, s, ctl);
      if (threadIdx.x == 0) {
         *rng = g; *stats = ctl.st;
         for (int k = 0; k < 4; ++k) prof[k] += ctl.prof[k];
      }
   }
}

// VnsSopSolver::iterate (vns_sop_solver.cpp:82-260).  tour lives in global memory (it changes only on improvements); the
// working path of a VNS step (actualVNSGOPPath, :195) exists only as the blocks' shared-memory replicas.
template <int G, int NR>
static __device__ void c2_vns_body(const C2Env &E, const PathDev &tour, const VnsParamsDev &prm, RngState *rng, VnsResultDev *res,
                                   double *impr_reward, double *impr_length, double *impr_ms, long long *prof)
{
   extern __shared__ __align__(16) unsigned char smem_c[];
   __shared__ C2Ctl ctl;
   __shared__ RngState g;
   __shared__ int s_stop, s_improved;
   const InstDev &I = E.I;
   const bool multi = E.grid.bar != nullptr;
   const bool lead = !multi || blockIdx.x == 0;
   double *vbuf_all = reinterpret_cast<double *>(smem_c);       // greedy scratch (start of the run only): overlaps the state
   __shared__ C2Tree tree;
   C2Sm s = c2_carve_env(smem_c, E);
   s.tr = &tree;
   unsigned long long epoch = 0;
   unsigned round_no = 0;
   if (threadIdx.x == 0) {
      c2_tree_init(tree, E);
      g = *rng; s_stop = 0;
      ctl.st = LsStats{0, 0, 0, 0, 0}; ctl.prof[0] = ctl.prof[1] = ctl.prof[2] = ctl.prof[3] = 0; ctl.err = 0; ctl.bad = 0; ctl.ema = 16;
   }
   __syncthreads();
   const double t0 = now_ms();
   if (lead) {
      cta_update(I, tour, tour.hdr[HDR_T]);
      if (prm.greedy) cta_greedy(I, tour, prm.budget, vbuf_all);         // :127-129
      if (threadIdx.x == 0) __threadfence();
   }
   if (multi) c2_grid_sync(E.grid, epoch);
   const int num_avail = multi ? __ldcg(tour.hdr + HDR_NVNS) : tour.hdr[HDR_NVNS];   // :115
   int last_impr = 1, it = 0, n_impr = 0;                               // :145
   double t_last = now_ms() - t0;
   if (lead && threadIdx.x == 0 && impr_reward && n_impr < prm.max_improvements) {   // :150-154
      impr_reward[n_impr] = tour.hd[HD_REWARD]; impr_length[n_impr] = tour.hd[HD_LEN]; impr_ms[n_impr] = t_last;
   }
   n_impr = 1;
   int tour_err = 0;
   __syncthreads();
   bool stop = false;
   unsigned step_no = 0;
   while (!stop) {
      ++it;
      if (it >= prm.num_iterations) stop = true;                         // :178
      if (it - last_impr >= prm.num_iterations_unimproved) stop = true;  // :182
      int k = 1;
      while (k <= 2) {
         // the time limit (:189-193) and the race flag are read by the lead block and broadcast: every block must take
         // the same branch
         if (lead && threadIdx.x == 0) {
            if (prm.max_time_ms > 0 && now_ms() - t0 >= prm.max_time_ms) s_stop = 1;
            if (prm.stop_flag) {      // multi-start race: not part of the reference's loop
               if (tour.hd[HD_REWARD] >= prm.target_reward) atomicExch(prm.stop_flag, 1);
               if (*(volatile int *)prm.stop_flag) s_stop = 2;
            }
            if (multi) { E.grid.gctl[step_no & 1] = s_stop; __threadfence(); }
         }
         if (multi) {
            c2_grid_sync(E.grid, epoch);
            if (threadIdx.x == 0) s_stop = __ldcg(E.grid.gctl + (step_no & 1));
         }
         ++step_no;
         __syncthreads();
         if (s_stop == 2) break;
         c2_load(E, tour, s, ctl, multi);                                // :195-197
         const double r0 = multi ? __ldcg(tour.hd + HD_REWARD) : tour.hd[HD_REWARD], l0 = (double)ctl.L;
         c2_shake(E, s, ctl, k, prm.budget, g);                          // :202
         c2_local_search<G, NR>(E, s, ctl, g, k, num_avail * num_avail, prm.budget, prm.skip, epoch, round_no);   // :207,:346
         if (threadIdx.x == 0) {
            const double r1 = c2_reward(E, s, ctl.T), l1 = (double)ctl.L;
            s_improved = (r1 > r0 || (r1 == r0 && l1 < l0));             // :213
         }
         __syncthreads();
         if (ctl.err) { stop = true; break; }
         if (s_improved) {
            last_impr = it;
            t_last = now_ms() - t0;
            if (multi) c2_grid_sync(E.grid, epoch);                      // nobody still reads the old tour
            if (lead) {
               c2_store(E, tour, s, ctl);                                // :227-228
               if (threadIdx.x == 0) {
                  if (impr_reward && n_impr < prm.max_improvements) {
                     impr_reward[n_impr] = tour.hd[HD_REWARD]; impr_length[n_impr] = tour.hd[HD_LEN]; impr_ms[n_impr] = t_last;
                  }
                  // :229-237 sanity check of the improved tour
                  const double calc = path_walk_back(I, tour, tour.hdr[HDR_T], nullptr, nullptr);
                  const double nl = tour.hd[HD_LEN];
                  if (calc > prm.budget || nl > prm.budget || calc - nl > 0.01) tour.hdr[HDR_ERR] = 4;
                  __threadfence();
               }
               __syncthreads();
            }
            ++n_impr;
            if (multi) c2_grid_sync(E.grid, epoch);
            tour_err = multi ? __ldcg(tour.hdr + HDR_ERR) : tour.hdr[HDR_ERR];
            if (tour_err) { stop = true; break; }
            k = 1;                                                       // :241
         } else {
            ++k;                                                         // :243
         }
      }
      __syncthreads();
      if (s_stop) stop = true;
   }
   __syncthreads();
   if (lead && threadIdx.x == 0) {
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

template <int G, int NR>
__global__ void __launch_bounds__(C2_THREADS) k_chain2_vns_run(C2Env E, PathDev tour, VnsParamsDev prm, RngState *rng, VnsResultDev *res,
                                                               double *impr_reward, double *impr_length, double *impr_ms, long long *prof)
{
   c2_vns_body<G, NR>(E, tour, prm, rng, res, impr_reward, impr_length, impr_ms, prof);
}

// Multi-start (the reference's --batch=N, algorithm.cc:120-161): one thread block per run, each with its own seed and
// tour; no communication between runs.
struct C2RunSlot {
   PathDev tour;
   RngState rng;
   VnsResultDev res;
};

template <int G, int NR>
__global__ void __launch_bounds__(C2_THREADS) k_chain2_vns_multi(C2Env E, C2RunSlot *slots, VnsParamsDev prm)
{
   C2RunSlot &sl = slots[blockIdx.x];
   c2_vns_body<G, NR>(E, sl.tour, prm, &sl.rng, &sl.res, nullptr, nullptr, nullptr, nullptr);
}
