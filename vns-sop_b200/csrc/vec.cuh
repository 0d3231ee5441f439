// Endpoint vectors (integer tables only): every tryTo* as  min_{s,t} A[s] + M[s][t] + B[t].
//
// The reference already shares work between candidates through the prefix / suffix vectors F[l], B[l]
// (vns_sop_path.cpp:1426-1539).  With integer entries min-plus products are exactly associative, so the
// same idea extends to the two ends of every candidate's chain: they depend only on (cluster, layer) or on
// (path layer, layer), not on the candidate.  Per path state two passes store (lcl = layer -> cluster, S = T + 1):
//   k_vec_uw, one warp per (layer l, cluster c) - one block step each:
//     U(c,l) = F[l] (x) D(lcl[l],c)          over c's samples: "in c right after layer l"      l = 0 .. S-1
//     W(c,l) = D(c,lcl[l]) (x) B[l]          over c's samples: "leave c into layer l"          l = 1 .. S
//     FI[l]  = F[l] as int
//   k_vec_pairs, one warp per (path layer x, layer l), cx = lcl[x] - one block step each from U / W:
//     InsF(x,l) = U(cx,l-1) (x) D(cx,lcl[l])        at layer l      cx inserted before l
//     InsB(x,l) = D(lcl[l-1],cx) (x) W(cx,l)        at layer l-1
//     RepF(x,l) = U(cx,l-1) (x) D(cx,lcl[l+1])      at layer l+1    layer l replaced by cx
//     RepB(x,l) = D(lcl[l-1],cx) (x) W(cx,l+1)      at layer l-1
// plus the matching sums of cluster-pair minima for the lower bounds.  Then
//   tryToAdd(c,idx)      = U(c,idx) . W(c,idx+1)                      (vns_sop_path.cpp:245-304)   no block at all
//   tryToReplace(c,idx)  = U(c,idx) . W(c,idx+2)                      (:456-513)
//   tryToRemove(idx)     = FI[idx] . W(lcl[idx],idx+2)                (:515-535)
//   tryToMove up   (i=a+1 -> before j=b+1)  = U(lcl[i+1],i-1) (x) M(i+1,j-1) . InsB(i,j)      (:932-1154)
//   tryToMove down (j=a+1 -> before i=b+1)  = InsF(j,i) (x) M(i,j-1) . W(lcl[j-1],j+1)
//   tryToExchange  (layers p=i+1, q=j)      = RepF(q,p) (x) M(p+1,q-1) . RepB(p,q)            (:635-797)
// where "." is min_s x[s] + y[s], M(a,b) is the segment product of eval.cuh (a table block when b == a+1,
// absent when b == a).  One candidate = at most ONE block step.  Lengths, lower bounds and flags are
// bit-identical to the block-by-block evaluation (tests/test_gpu_parity.py runs all three forms).
// Build cost per path state: 2 C (T+2) + 4 T (T+2) block steps (the first version did 8 C (T+2)); the U pass
// streams whole row bands of the table (consecutive warps read consecutive blocks).
#pragma once
#include "eval.cuh"

struct VecCache {
   const int *vec;   // U, W: ((k * LS + l) * C + c) * mp, k = 0 / 1; FI: row l; then InsF, InsB, RepF, RepB: ((fam * LS + x) * LS + l) * mp
   const int *lbs;   // insLB[C][LS], repLB[C][LS], remLB[LS]
   int LS;           // layer stride (>= S + 1)
};

__host__ __device__ inline size_t vec_uw_ints(int C, int LS, int mp) { return (size_t)2 * LS * C * mp + (size_t)LS * mp; }
__host__ __device__ inline size_t vec_ints(int C, int LS, int mp) { return vec_uw_ints(C, LS, mp) + (size_t)4 * LS * LS * mp; }
__host__ __device__ inline size_t vec_lb_ints(int C, int LS) { return (size_t)(2 * C + 1) * LS; }

__device__ __forceinline__ unsigned vec_uw(const VecCache &vc, int C, int mp, int k, int c, int l)   // k: 0 U, 1 W
{
   return (unsigned)(((k * vc.LS + l) * C + c) * mp);
}
__device__ __forceinline__ unsigned vec_fi(const VecCache &vc, int C, int mp, int l) { return (unsigned)((2 * vc.LS * C + l) * mp); }
__device__ __forceinline__ unsigned vec_pair(const VecCache &vc, int C, int mp, int fam, int x, int l)   // fam 0..3 = InsF, InsB, RepF, RepB
{
   return (unsigned)(2 * vc.LS * C + vc.LS) * (unsigned)mp + (unsigned)(((fam * vc.LS + x) * vc.LS + l) * mp);
}

// A warp stages a whole table block in its slice of shared memory with cp.async (fire and forget): every block a warp
// needs is in flight at once and costs ONE memory round trip; the min-plus loops then run out of shared memory.
// Always 16-byte copies, whatever the block's alignment: the copy starts at the 16-byte line below the block and the
// data sits at the same offset inside the (16-byte aligned) slot - up to 15 bytes of neighbouring entries come along on
// either side (inside the table's allocation: it starts 256-byte aligned and ends with slack).  Returns the block.
// TT = int (int32 table) or unsigned short (uint16 table: half the bytes).  Slot size: vec_slot_bytes.
__host__ __device__ inline size_t vec_slot_bytes(int mp, size_t elem) { return (((size_t)mp * mp * elem + 15) & ~(size_t)15) + 32; }

// k_vec_pairs: a staged block + the four endpoint vectors it is multiplied with (pitch: mp rounded up to 4 ints)
__host__ __device__ inline size_t vec_pairs_slot_bytes(int mp, size_t elem) { return vec_slot_bytes(mp, elem) + (size_t)16 * ((mp + 3) & ~3); }

template <typename TT>
__device__ __forceinline__ const TT *warp_stage(char *slot, const TT *__restrict__ src, int n, int lane)
{
   const unsigned shift = (unsigned)((size_t)src & 15);
   const char *s0 = reinterpret_cast<const char *>(src) - shift;
   const unsigned d0 = (unsigned)__cvta_generic_to_shared(slot);
   const int nb = (int)((shift + (unsigned)n * (unsigned)sizeof(TT) + 15u) & ~15u);
   for (int e = lane * 16; e < nb; e += 512) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d0 + e), "l"(s0 + e));
   return reinterpret_cast<const TT *>(slot + shift);
}
__device__ __forceinline__ void warp_stage_wait()
{
   asm volatile("cp.async.commit_group;");
   asm volatile("cp.async.wait_group 0;");
   __syncwarp();
}
// lane = output index; v (lane-held, INF beyond the source size) times the staged ma x mb block
template <typename TT>
__device__ __forceinline__ int vec_fwd(const TT *blk, int ma, int mb, int v, int lane)
{
   int acc = VNS_INT_INF;
   const int j = min(lane, mb - 1);
#pragma unroll 4
   for (int s = 0; s < ma; ++s) acc = __viaddmin_s32(__shfl_sync(0xffffffffu, v, s), (int)blk[s * mb + j], acc);
   return lane < mb ? acc : VNS_INT_INF;
}
// lane = source index; the staged ma x mb block times v (lane-held over the mb destination samples).  Lane s walks its
// row DIAGONALLY (column (s + t) mod mb at step t): with mb = 32 the 32 lanes then hit 32 different banks, where the
// straight order (every lane the same column: stride mb) is a 32-way bank conflict.  min is order-free: same result.
template <typename TT>
__device__ __forceinline__ int vec_bwd(const TT *blk, int ma, int mb, int v, int lane)
{
   int acc = VNS_INT_INF;
   const int s = min(lane, ma - 1);
   const TT *row = blk + s * mb;
   int j = s % mb;
#pragma unroll 4
   for (int t = 0; t < mb; ++t) {
      acc = __viaddmin_s32(__shfl_sync(0xffffffffu, v, j), (int)row[j], acc);
      j = j + 1 == mb ? 0 : j + 1;
   }
   return lane < ma ? acc : VNS_INT_INF;
}

// per-layer description of the path written by k_seg_offsets (eval.cuh): cluster, first sample, size
struct LayerMeta {
   const int *cum, *lcl, *loff;
   __device__ __forceinline__ int cl(int l) const { return __ldg(lcl + l); }
   __device__ __forceinline__ int off(int l) const { return __ldg(loff + l); }
   __device__ __forceinline__ int m(int l) const { return __ldg(cum + l + 1) - __ldg(cum + l); }
};
__device__ __forceinline__ LayerMeta layer_meta(const InstDev &I, const int *meta)
{
   const int R = I.rows + 3;
   return LayerMeta{meta, meta + 4 * R, meta + 5 * R};
}

// F and B rows as int32 (fbi: F rows, then B rows from LS * mp on): the build kernels read every row once per cluster, and a
// double -> int conversion costs them an FP64-pipe instruction each time (the hottest stall of k_vec_uw_packed in ncu r5h)
static __global__ void __launch_bounds__(256) k_rows_to_int(InstDev I, PathDev P, int *__restrict__ fbi, int LS)
{
   const int n = (P.hdr[HDR_T] + 2) * I.mp;
   for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x) {
      fbi[e] = (int)P.F[e];
      fbi[LS * I.mp + e] = (int)P.B[e];
   }
}

// The lower-bound sums of k_vec_uw / k_vec_uw_packed as a kernel of their own, one thread per (cluster c, layer l), l fastest:
// inside the build kernels a whole warp issued the three gathers and their 64-bit indices for one lane in ten (ncu r5k: 14 %
// of k_vec_uw_packed's instructions).  Large instances only (the build kernels skip the sums when they get int32 rows).
static __global__ void __launch_bounds__(256) k_vec_lbs(InstDev I, PathDev P, const int *__restrict__ meta, const int *__restrict__ dmin,
                                                        int *__restrict__ lbs, int LS)
{
   const int T = P.hdr[HDR_T], S = T + 1, C = I.C;
   const LayerMeta lm = layer_meta(I, meta);
   const long long n = (long long)C * S;
   for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) {
      const int c = (int)(e / S), l = (int)(e - (long long)c * S) + 1;            // l = 1 .. S
      const int a = lm.cl(l - 1), cl = lm.cl(l), nx = l <= S - 1 ? lm.cl(l + 1) : cl;
      const int din = __ldg(dmin + (size_t)a * C + c);
      lbs[(size_t)c * LS + l] = din + __ldg(dmin + (size_t)c * C + cl);                               // insert c before l
      if (l <= S - 1) lbs[(size_t)(C + c) * LS + l] = din + __ldg(dmin + (size_t)c * C + nx);         // c replaces l
      if (c == 0 && l <= S - 1) lbs[(size_t)2 * C * LS + l] = __ldg(dmin + (size_t)a * C + nx);       // l removed
   }
}

// one warp per (l, c), l strided over grid.y: consecutive warps read consecutive blocks of row band lcl[l] (U) and write
// consecutive vectors.  Needs mp <= 32.  Dynamic shared memory: 2 slots (vec_slot_bytes) per warp.
template <typename TT>
__global__ void __launch_bounds__(256) k_vec_uw(InstDev I, PathDev P, const TT *__restrict__ Dv, const int *__restrict__ meta,
                                                int *__restrict__ vec, int *__restrict__ lbs, int LS, const int *__restrict__ fbi,
                                                const int *__restrict__ dmin)
{
   extern __shared__ int smem_i[];
   const int T = P.hdr[HDR_T], S = T + 1, C = I.C, mp = I.mp;
   const int lane = threadIdx.x & 31;
   const int c = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
   if (c >= C) return;
   const LayerMeta lm = layer_meta(I, meta);
   VecCache vc{vec, lbs, LS};
   const size_t slot = vec_slot_bytes(mp, sizeof(TT));
   char *sU = reinterpret_cast<char *>(smem_i) + (threadIdx.x >> 5) * 2 * slot, *sW = sU + slot;
   const int mc = __ldg(I.m + c), oc = __ldg(I.off + c);
   // T is only known on the device: grid.y is a fixed number of layer lanes, every warp strides over the layers
   for (int l = blockIdx.y; l <= S; l += gridDim.y) {
      const int cl = lm.cl(l), ml = lm.m(l), ol = lm.off(l);
      const bool doU = l <= S - 1, doW = l >= 1, sym = I.symmetric != 0;
      // symmetric tables: D(c, cl) is the transpose of D(cl, c) - the column strip of W comes out of the same row band
      const TT *bU = nullptr, *bW = nullptr;
      if (doU || sym) bU = warp_stage(sU, Dv + ((long long)ol * I.N + (long long)ml * oc), ml * mc, lane);     // D(cl, c)
      if (doW && !sym) bW = warp_stage(sW, Dv + ((long long)oc * I.N + (long long)mc * ol), mc * ml, lane);    // D(c, cl)
      // fbi: the rows as int32 (k_rows_to_int, large instances), else converted here
      const int f = (doU && lane < ml) ? (fbi ? __ldg(fbi + l * mp + lane) : (int)P.F[(size_t)l * mp + lane]) : VNS_INT_INF;
      const int bv = (doW && lane < ml) ? (fbi ? __ldg(fbi + (LS + l) * mp + lane) : (int)P.B[(size_t)l * mp + lane]) : VNS_INT_INF;
      int lb_ins = 0, lb_rep = 0, lb_rem = 0;
      if (doW && lane == 0 && !fbi) {   // lower-bound sums (integers: order-free); large instances: k_vec_lbs
         const int a = lm.cl(l - 1), nx = l <= S - 1 ? lm.cl(l + 1) : cl;
         const int din = __ldg(dmin + (size_t)a * C + c);
         lb_ins = din + __ldg(dmin + (size_t)c * C + cl);
         lb_rep = din + __ldg(dmin + (size_t)c * C + nx);
         lb_rem = __ldg(dmin + (size_t)a * C + nx);
      }
      warp_stage_wait();
      if (doU) {   // U(c,l) = F[l] (x) D(cl, c)
         const int u = vec_fwd(bU, ml, mc, f, lane);
         if (lane < mp) vec[vec_uw(vc, C, mp, 0, c, l) + lane] = u;
         if (c == 0 && lane < mp) vec[vec_fi(vc, C, mp, l) + lane] = f;
      }
      if (doW) {   // W(c,l) = D(c, cl) (x) B[l]
         const int wv = sym ? vec_fwd(bU, ml, mc, bv, lane) : vec_bwd(bW, mc, ml, bv, lane);
         if (lane < mp) vec[vec_uw(vc, C, mp, 1, c, l) + lane] = wv;
         if (lane == 0 && !fbi) {
            lbs[(size_t)c * LS + l] = lb_ins;                                   // insert c before l
            if (l <= S - 1) lbs[(size_t)(C + c) * LS + l] = lb_rep;             // c replaces l
            if (c == 0 && l <= S - 1) lbs[(size_t)2 * C * LS + l] = lb_rem;     // l removed
         }
      }
      __syncwarp();   // the staged blocks are overwritten by the next layer
   }
}

// Small clusters on a symmetric table: PK = 32 / SLOT clusters per warp (SLOT = 8, 10 or 16 lanes each, >= mp).  The blocks
// D(cl, c0 .. c0 + PK - 1) of consecutive clusters are ONE contiguous run of the block-major row band, staged with a single
// warp_stage; group g = lane / SLOT works on cluster c0 + g, lane j = lane % SLOT on its sample j.  With clusters of 10
// samples the one-cluster-per-warp form has a third of its lanes busy (20 k-node instance: 0.66 ms for 0.5 GB).
template <typename TT, int SLOT>
__global__ void __launch_bounds__(256) k_vec_uw_packed(InstDev I, PathDev P, const TT *__restrict__ Dv, const int *__restrict__ meta,
                                                       int *__restrict__ vec, int *__restrict__ lbs, int LS, const int *__restrict__ fbi,
                                                       const int *__restrict__ dmin)
{
   extern __shared__ int smem_i[];
   constexpr int PK = 32 / SLOT;
   const int T = P.hdr[HDR_T], S = T + 1, C = I.C, mp = I.mp;
   const int lane = threadIdx.x & 31, g = lane / SLOT, j = lane - g * SLOT;
   const int c0 = (int)(((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * PK);
   if (c0 >= C) return;
   const LayerMeta lm = layer_meta(I, meta);
   VecCache vc{vec, lbs, LS};
   const size_t slot = PK * vec_slot_bytes(mp, sizeof(TT));
   char *sU = reinterpret_cast<char *>(smem_i) + (threadIdx.x >> 5) * slot;
   const int c = c0 + g;
   const bool cok = g < PK && c < C;
   const int mc = cok ? __ldg(I.m + c) : 0, oc = cok ? __ldg(I.off + c) : 0;
   const int oc0 = __ldg(I.off + c0), oend = __ldg(I.off + min(c0 + PK, C));       // samples of the warp's clusters: [oc0, oend)
   for (int l = blockIdx.y; l <= S; l += gridDim.y) {
      const int cl = lm.cl(l), ml = lm.m(l), ol = lm.off(l);
      const bool doU = l <= S - 1, doW = l >= 1;
      const TT *run = warp_stage(sU, Dv + ((long long)ol * I.N + (long long)ml * oc0), ml * (oend - oc0), lane);   // D(cl, c0 ..)
      // fbi: the rows as int32 (k_rows_to_int, large instances), else converted here
      const int f = (doU && lane < ml) ? (fbi ? __ldg(fbi + l * mp + lane) : (int)P.F[(size_t)l * mp + lane]) : VNS_INT_INF;
      const int bv = (doW && lane < ml) ? (fbi ? __ldg(fbi + (LS + l) * mp + lane) : (int)P.B[(size_t)l * mp + lane]) : VNS_INT_INF;
      int lb_ins = 0, lb_rep = 0, lb_rem = 0;
      if (doW && cok && j == 0 && !fbi) {   // lower-bound sums (integers: order-free); large instances: k_vec_lbs
         const int a = lm.cl(l - 1), nx = l <= S - 1 ? lm.cl(l + 1) : cl;
         const int din = __ldg(dmin + (size_t)a * C + c);
         lb_ins = din + __ldg(dmin + (size_t)c * C + cl);
         lb_rep = din + __ldg(dmin + (size_t)c * C + nx);
         lb_rem = __ldg(dmin + (size_t)a * C + nx);
      }
      warp_stage_wait();
      // this group's block, this lane's column; idle groups walk the start of the run (their results are not stored)
      const TT *blk = cok ? run + ml * (oc - oc0) + min(j, max(mc, 1) - 1) : run;
      int u = VNS_INT_INF, wv = VNS_INT_INF;
      // both products walk the same staged entry (symmetric table).  The kernel is issue-bound (ncu r5h: 376 instructions per
      // (warp, layer), 20 of them add-mins): a layer of SLOT samples - the common case - runs a fixed, unrolled trip count
      if (ml == SLOT) {
#pragma unroll
         for (int s_ = 0; s_ < SLOT; ++s_, blk += mc) {
            const int d = (int)*blk;
            u = __viaddmin_s32(__shfl_sync(0xffffffffu, f, s_), d, u);
            wv = __viaddmin_s32(__shfl_sync(0xffffffffu, bv, s_), d, wv);
         }
      } else {
         for (int s_ = 0; s_ < ml; ++s_, blk += mc) {
            const int d = (int)*blk;
            u = __viaddmin_s32(__shfl_sync(0xffffffffu, f, s_), d, u);
            wv = __viaddmin_s32(__shfl_sync(0xffffffffu, bv, s_), d, wv);
         }
      }
      if (cok && j < mp) {
         if (doU) {
            vec[vec_uw(vc, C, mp, 0, c, l) + j] = j < mc ? u : VNS_INT_INF;
            if (c == 0) vec[vec_fi(vc, C, mp, l) + j] = f;      // c == 0 is group 0: lane == j holds F[l][j]
         }
         if (doW) vec[vec_uw(vc, C, mp, 1, c, l) + j] = j < mc ? wv : VNS_INT_INF;
      }
      if (doW && cok && j == 0 && !fbi) {
         lbs[(size_t)c * LS + l] = lb_ins;                                   // insert c before l
         if (l <= S - 1) lbs[(size_t)(C + c) * LS + l] = lb_rep;             // c replaces l
         if (c == 0 && l <= S - 1) lbs[(size_t)2 * C * LS + l] = lb_rem;     // l removed
      }
      __syncwarp();   // the staged run is overwritten by the next layer
   }
}

// one warp per (x, l), x = 1 .. T (path layer, cluster cx; strided over grid.y), l = 1 .. S (strided over the warps of grid.x:
// the host sizes grid.x from T as it last read it, the result does not depend on it).  Two table blocks, four products.
template <typename TT>
__global__ void __launch_bounds__(256) k_vec_pairs(InstDev I, PathDev P, const TT *__restrict__ Dv, const int *__restrict__ meta,
                                                   int *__restrict__ vec, int LS)
{
   extern __shared__ int smem_i[];
   const int T = P.hdr[HDR_T], S = T + 1, C = I.C, mp = I.mp;
   const int lane = threadIdx.x & 31;
   const int wl0 = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), nwx = (int)((gridDim.x * blockDim.x) >> 5);
   const LayerMeta lm = layer_meta(I, meta);
   VecCache vc{vec, nullptr, LS};
   const size_t slot = vec_pairs_slot_bytes(mp, sizeof(TT)), voff = vec_slot_bytes(mp, sizeof(TT));
   char *s1 = reinterpret_cast<char *>(smem_i) + (threadIdx.x >> 5) * 2 * slot, *s2 = s1 + slot;
   for (int wl = wl0; wl <= S; wl += nwx) {
   if (I.symmetric) {
      // one block D(cx, lcl[l]) per warp, l = 0 .. S: it is also D(lcl[l], cx) transposed, so all four products are
      // forward steps through it:  InsF(x,l), RepF(x,l-1) from U;  InsB(x,l+1), RepB(x,l+1) from W
      const int l = wl;
      const int ml = lm.m(l), ofl = lm.off(l);
      // double-buffered over x: the block and the four vectors of the NEXT x are in flight (cp.async into the warp's second
      // slot, plain loads into registers) while the current x is reduced - a warp's iterations were one exposed HBM round
      // trip each (16 % of the HBM peak on the 20 k-node state).  After that the kernel was issue-bound (ncu r5h: 367
      // instructions per (x, l), 40 of them add-mins): the parts of the indices that do not depend on x are computed once per
      // layer, the product walks two block rows per trip with a stepped pointer.
      const unsigned xstr = (unsigned)LS * (unsigned)mp, fstr = (unsigned)LS * xstr, pbase = (unsigned)(2 * LS * C + LS) * (unsigned)mp;
      const bool on1 = l >= 1, on2 = l >= 2, on3 = l <= S - 1, on4 = l <= S - 2;
      // vec_uw(k, cx, ll) = ((k LS + ll) C + cx) mp;   vec_pair(fam, x, ll) = pbase + fam fstr + x xstr + ll mp
      const unsigned ld1 = (unsigned)(max(l - 1, 0) * C) * mp + lane, ld2 = (unsigned)(max(l - 2, 0) * C) * mp + lane;
      const unsigned ld3 = (unsigned)((LS + min(l + 1, S)) * C) * mp + lane, ld4 = (unsigned)((LS + min(l + 2, S)) * C) * mp + lane;
      const unsigned st1 = pbase + (unsigned)l * mp + lane, st2 = pbase + 2 * fstr + (unsigned)max(l - 1, 0) * mp + lane;
      const unsigned st3 = pbase + fstr + (unsigned)(l + 1) * mp + lane, st4 = pbase + 3 * fstr + (unsigned)(l + 1) * mp + lane;
      // the four vectors go to shared memory with the block (4-byte cp.async per lane) and are read back four entries at a
      // time as broadcast loads: the shuffles that handed them round (4 per block row) had the kernel on the shared-memory /
      // shuffle pipe (ncu r5h: LSU pipe 60 %, short-scoreboard stalls 3.4 warps per issue)
      const int VP = (mp + 3) & ~3;
      struct XIn { int mx; const TT *b; const int *v; };
      auto issue = [&](int x, char *slot_) {
         XIn q;
         const int cx = lm.cl(x), ox = lm.off(x);
         q.mx = lm.m(x);
         q.b = warp_stage(slot_, Dv + ((long long)ox * I.N + (long long)q.mx * ofl), q.mx * ml, lane);
         q.v = reinterpret_cast<const int *>(slot_ + voff);
         if (lane < q.mx) {
            const unsigned cb = (unsigned)cx * mp, d0 = (unsigned)__cvta_generic_to_shared(slot_ + voff) + 4 * lane;
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d0), "l"(vec + ld1 + cb) : "memory");
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d0 + 4 * VP), "l"(vec + ld2 + cb) : "memory");
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d0 + 8 * VP), "l"(vec + ld3 + cb) : "memory");
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d0 + 12 * VP), "l"(vec + ld4 + cb) : "memory");
         }
         asm volatile("cp.async.commit_group;" ::: "memory");
         return q;
      };
      int x = blockIdx.y + 1, par = 0;
      if (x > T) return;      // no x for this grid.y lane, whatever the layer
      const int j = min(lane, ml - 1);
      const bool st_on = lane < mp, in_l = lane < ml;
      XIn cur = issue(x, s1);
      for (; x <= T; x += gridDim.y) {
         const int xn = x + gridDim.y;
         XIn nxt = cur;
         if (xn <= T) {
            nxt = issue(xn, par ? s1 : s2);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
         } else
            asm volatile("cp.async.wait_group 0;" ::: "memory");
         __syncwarp();
         const TT *bp = cur.b + j;
         const int *v = cur.v;
         const int mx = cur.mx;
         // vectors whose layer does not exist (on1 .. on4 false) are read all the same - from the row next to it - and their
         // results are not stored
         int a1 = VNS_INT_INF, a2 = VNS_INT_INF, a3 = VNS_INT_INF, a4 = VNS_INT_INF;
         int s = 0;
         for (; s + 3 < mx; s += 4, bp += 4 * ml) {
            const int4 U1 = *reinterpret_cast<const int4 *>(v + s), U2 = *reinterpret_cast<const int4 *>(v + VP + s);
            const int4 W1 = *reinterpret_cast<const int4 *>(v + 2 * VP + s), W2 = *reinterpret_cast<const int4 *>(v + 3 * VP + s);
            const int d0 = (int)bp[0], d1 = (int)bp[ml], d2 = (int)bp[2 * ml], d3 = (int)bp[3 * ml];
            a1 = __viaddmin_s32(U1.x, d0, a1); a2 = __viaddmin_s32(U2.x, d0, a2); a3 = __viaddmin_s32(W1.x, d0, a3); a4 = __viaddmin_s32(W2.x, d0, a4);
            a1 = __viaddmin_s32(U1.y, d1, a1); a2 = __viaddmin_s32(U2.y, d1, a2); a3 = __viaddmin_s32(W1.y, d1, a3); a4 = __viaddmin_s32(W2.y, d1, a4);
            a1 = __viaddmin_s32(U1.z, d2, a1); a2 = __viaddmin_s32(U2.z, d2, a2); a3 = __viaddmin_s32(W1.z, d2, a3); a4 = __viaddmin_s32(W2.z, d2, a4);
            a1 = __viaddmin_s32(U1.w, d3, a1); a2 = __viaddmin_s32(U2.w, d3, a2); a3 = __viaddmin_s32(W1.w, d3, a3); a4 = __viaddmin_s32(W2.w, d3, a4);
         }
         for (; s < mx; ++s, bp += ml) {
            const int d0 = (int)bp[0];
            a1 = __viaddmin_s32(v[s], d0, a1); a2 = __viaddmin_s32(v[VP + s], d0, a2);
            a3 = __viaddmin_s32(v[2 * VP + s], d0, a3); a4 = __viaddmin_s32(v[3 * VP + s], d0, a4);
         }
         if (st_on) {
            const unsigned xb = (unsigned)x * xstr;
            if (on1) vec[st1 + xb] = in_l ? a1 : VNS_INT_INF;
            if (on2) vec[st2 + xb] = in_l ? a2 : VNS_INT_INF;
            if (on3) vec[st3 + xb] = in_l ? a3 : VNS_INT_INF;
            if (on4) vec[st4 + xb] = in_l ? a4 : VNS_INT_INF;
         }
         __syncwarp();   // this x's slot is the target of the staging two iterations on
         cur = nxt;
         par ^= 1;
      }
      continue;
   }
   const int l = wl + 1;
   if (l > S) break;
   const int ml = lm.m(l), ma = lm.m(l - 1), ofl = lm.off(l), ofa = lm.off(l - 1);
   for (int x = blockIdx.y + 1; x <= T; x += gridDim.y) {
      const int cx = lm.cl(x), mx = lm.m(x), ox = lm.off(x);
      const TT *b1 = warp_stage(s1, Dv + ((long long)ox * I.N + (long long)mx * ofl), mx * ml, lane);          // D(cx, lcl[l])
      const TT *b2 = warp_stage(s2, Dv + ((long long)ofa * I.N + (long long)ma * ox), ma * mx, lane);          // D(lcl[l-1], cx)
      auto ld = [&](unsigned off, bool on) { return (on && lane < mx) ? __ldg(vec + off + lane) : VNS_INT_INF; };
      auto put = [&](unsigned off, int v) { if (lane < mp) vec[off + lane] = v; };
      const int u1 = ld(vec_uw(vc, C, mp, 0, cx, l - 1), true), u2 = ld(vec_uw(vc, C, mp, 0, cx, max(l - 2, 0)), l >= 2);
      const int w1 = ld(vec_uw(vc, C, mp, 1, cx, l), true), w2 = ld(vec_uw(vc, C, mp, 1, cx, min(l + 1, S)), l <= S - 1);
      warp_stage_wait();
      put(vec_pair(vc, C, mp, 0, x, l), vec_fwd(b1, mx, ml, u1, lane));                       // InsF(x,l)   from U(cx,l-1)
      if (l >= 2) put(vec_pair(vc, C, mp, 2, x, l - 1), vec_fwd(b1, mx, ml, u2, lane));       // RepF(x,l-1) from U(cx,l-2)
      put(vec_pair(vc, C, mp, 1, x, l), vec_bwd(b2, ma, mx, w1, lane));                       // InsB(x,l)   from W(cx,l)
      if (l <= S - 1) put(vec_pair(vc, C, mp, 3, x, l), vec_bwd(b2, ma, mx, w2, lane));       // RepB(x,l)   from W(cx,l+1)
      __syncwarp();   // the staged blocks are overwritten by the next x
   }
   }
}

// decoded candidate in endpoint-vector form
struct VecCode {
   int ok;            // 0: indices out of range
   unsigned offA, offB;
   int mA, mB;        // sizes of A and B
   int la, lb;        // block la -> lb between them when lb > la (path layers)
   int lbsum;         // dist_between of the lower bound
   int s0, sK;        // layers whose minF / minB enter the lower bound
   int nolb;
};

// lcl: layer -> cluster, lm(l): size of layer l, cm(c): size of cluster c, pd(l): prefix sums of dmin along the path
template <typename LM, typename CM, typename PD>
__device__ __forceinline__ VecCode decode_cand_vec(const VecCache &vc, int T, int C, int mp, const int *lcl, LM lm, CM cm, PD pd, int op,
                                                   int a, int b)
{
   VecCode v;
   v.ok = 0; v.offA = v.offB = 0; v.mA = v.mB = 0; v.la = v.lb = 0; v.lbsum = 0; v.s0 = v.sK = 0; v.nolb = 0;
   const int *insLB = vc.lbs, *repLB = vc.lbs + (size_t)C * vc.LS, *remLB = vc.lbs + (size_t)2 * C * vc.LS;
   if (op == VNSGPU_OP_ADD) {
      if (b < 0 || b > T || a < 0 || a >= C) return v;
      v.ok = 1; v.mA = v.mB = cm(a); v.offA = vec_uw(vc, C, mp, 0, a, b); v.offB = vec_uw(vc, C, mp, 1, a, b + 1);
      v.lbsum = __ldg(insLB + (size_t)a * vc.LS + b + 1); v.s0 = b; v.sK = b + 1;
   } else if (op == VNSGPU_OP_REPLACE) {
      if (b < 0 || b >= T || a < 0 || a >= C) return v;
      v.ok = 1; v.mA = v.mB = cm(a); v.offA = vec_uw(vc, C, mp, 0, a, b); v.offB = vec_uw(vc, C, mp, 1, a, b + 2);
      v.lbsum = __ldg(repLB + (size_t)a * vc.LS + b + 1); v.s0 = b; v.sK = b + 2;
   } else if (op == VNSGPU_OP_REMOVE) {
      if (a < 0 || a >= T) return v;
      v.ok = 1; v.mA = v.mB = lm(a); v.offA = vec_fi(vc, C, mp, a); v.offB = vec_uw(vc, C, mp, 1, lcl[a], a + 2);
      v.s0 = a; v.sK = a + 2; v.nolb = 1;
   } else if (op == VNSGPU_OP_MOVE) {
      if (a >= T || a < 0 || b > T || b < 0 || abs(b - a) < 2) return v;
      const int i = min(a, b) + 1, j = max(a, b) + 1;
      v.ok = 1;
      if (b > a) {   // layer i moves up, in front of layer j
         v.la = i + 1; v.lb = j - 1; v.offA = vec_uw(vc, C, mp, 0, lcl[i + 1], i - 1); v.offB = vec_pair(vc, C, mp, 1, i, j);
         v.lbsum = __ldg(remLB + i) + (pd(j - 1) - pd(i + 1)) + __ldg(insLB + (size_t)lcl[i] * vc.LS + j);
         v.s0 = i - 1; v.sK = j;
      } else {       // layer j moves down, in front of layer i
         v.la = i; v.lb = j - 1; v.offA = vec_pair(vc, C, mp, 0, j, i); v.offB = vec_uw(vc, C, mp, 1, lcl[j - 1], j + 1);
         v.lbsum = __ldg(insLB + (size_t)lcl[j] * vc.LS + i) + (pd(j - 1) - pd(i)) + __ldg(remLB + j);
         v.s0 = i - 1; v.sK = j + 1;
      }
      v.mA = lm(v.la); v.mB = lm(v.lb);
   } else if (op == VNSGPU_OP_EXCHANGE) {
      if (a >= T || a < 0 || b >= T || b < 0 || abs(a - b) < 2) return v;
      const int p = min(a, b) + 1, q = max(a, b) + 1;   // the two exchanged layers
      v.ok = 1; v.la = p + 1; v.lb = q - 1;
      v.offA = vec_pair(vc, C, mp, 2, q, p); v.offB = vec_pair(vc, C, mp, 3, p, q);
      v.lbsum = __ldg(repLB + (size_t)lcl[q] * vc.LS + p) + (pd(q - 1) - pd(p + 1)) + __ldg(repLB + (size_t)lcl[p] * vc.LS + q);
      v.s0 = p - 1; v.sK = q + 1;
      v.mA = lm(v.la); v.mB = lm(v.lb);
   }
   return v;
}

// PF: the table does not fit the caches (synthetic scaled instances) - while an iteration's loads are in flight the next
// candidate's block rows and vectors are prefetched into L1 (compiled in only there: on cache-resident tables the extra
// registers cost up to 10 %).  An earlier form for these tables issued all G row loads of a block step before the first
// min (128 registers, 2 CTAs / SM): 1.48 ms against 1.17 ms for this one on the 64 k-sample DOP table (ncu r2c).
template <int G, int NR, bool PF>
__global__ void __launch_bounds__(VNS_EVAL_THREADS, NR > 2 ? 3 : 4)  // CTAs / SM: 4 (64 registers), 3 with four registers per vector
k_eval_vec(InstDev I, const int *__restrict__ Dv, PathDev P, SegCache seg, VecCache vc, long long n, long long idx0,
           const void *__restrict__ cands, const unsigned *__restrict__ perm, double budget, int skipflag,
           double *__restrict__ len, double *__restrict__ lb, uint8_t *__restrict__ flags, unsigned long long *d_first,
           int fmt, int *__restrict__ leni)
{
   // fmt & 3: 0 = vnsgpu_cand (16 B), 2 = vnsgpu_pack_cand32 words (4 B) read in place; leni: lengths as int32 (may be
   // null), with fmt & 4 as result words: length | flags << 29, all ones for bad indices (vnsgpu_eval_batch_packed32r)
   extern __shared__ int smem_i[];
   constexpr unsigned FULL = 0xffffffffu;
   constexpr int INF = VNS_INT_INF;
   const int T = P.hdr[HDR_T], C = I.C, mp = I.mp;
   int *lcl = smem_i;
   int2 *cinfo = reinterpret_cast<int2 *>(smem_i + ((T + 2 + 1) & ~1));
   for (int l = threadIdx.x; l < T + 2; l += blockDim.x) lcl[l] = layer_cluster(I, P.tg, T, l);
   for (int c = threadIdx.x; c < C; c += blockDim.x) cinfo[c] = make_int2(I.off[c], I.m[c]);
   long long *s_rowoff = reinterpret_cast<long long *>(cinfo + C);
   int *s_cum = reinterpret_cast<int *>(s_rowoff + (T + 2)), *s_pd = s_cum + (T + 3);
   for (int l = threadIdx.x; l < T + 3; l += blockDim.x) s_cum[l] = seg.cum[l];
   for (int l = threadIdx.x; l < T + 2; l += blockDim.x) { s_pd[l] = seg.pd[l]; s_rowoff[l] = seg.rowoff[l]; }
   __syncthreads();

   // A warp takes 32 candidates at a time.  Phase 1: LANE-parallel - lane L fetches and decodes candidate
   // base + L (offsets of its two vectors and of its block, sizes, lower-bound sum) and, by xor-shuffles over
   // its iteration-mates, the warp-wide loop bounds of the iteration it will be evaluated in.  Phase 2: G
   // iterations; in iteration `it` the NG = 32 / G groups evaluate candidates it * NG + g, fetching their plan
   // from the lane that decoded it (4 shuffles), and hand the result back to that lane.  Phase 3:
   // lane-parallel epilogue (bound, flags, stores, first accepted index) for all 32 candidates at once.
   constexpr int NG = 32 / G;
   const int lane = threadIdx.x & 31, gl = lane % G, g = lane / G;
   const double L = P.hd[HD_LEN];
   const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
   const long long warp0 = (((long long)blockIdx.x * blockDim.x) >> 5) + (threadIdx.x >> 5);
   unsigned long long best = ULLONG_MAX;

   for (long long base = warp0 * 32; base < n; base += nwarps * 32) {
      // ---- phase 1 -----------------------------------------------------------------------------------
      const long long slot = base + lane;
      long long kidx = slot;
      if (perm && slot < n) kidx = __ldg(perm + slot);
      int4 cd = make_int4(-1, 0, 0, -1);
      if (slot < n) {
         if ((fmt & 3) == 2) {
            const unsigned w = __ldg(reinterpret_cast<const unsigned *>(cands) + kidx);
            cd.x = (int)(w & 7); cd.y = (int)((w >> 3) & 0x3fff); cd.z = (int)((w >> 17) & 0x3fff); cd.w = (w >> 31) ? cd.y : -1;
         } else
            cd = __ldg(reinterpret_cast<const int4 *>(cands) + kidx);
      }
      if (cd.w >= C) cd.x = -1;                   // gain cluster out of range: a bad descriptor like any other bad index
      const VecCode vcd = decode_cand_vec(vc, T, C, mp, lcl, [&](int l) { return s_cum[l + 1] - s_cum[l]; },
                                          [&](int c) { return cinfo[c].y; }, [&](int l) { return s_pd[l]; }, cd.x, cd.y, cd.z);
      long long p_rb = 0;
      unsigned p_pack;
      {
         const int mA = vcd.mA, mB = vcd.mB;
         const bool blk = vcd.ok && vcd.lb > vcd.la;
         if (blk) {   // block la -> lb: a table block when adjacent, else the segment product M(la, lb)
            if (vcd.lb == vcd.la + 1) {
               const int2 cia = cinfo[lcl[vcd.la]], cib = cinfo[lcl[vcd.lb]];
               p_rb = (long long)cia.x * I.N + (long long)cia.y * cib.x;
            } else
               p_rb = s_rowoff[vcd.la] + (long long)mA * (s_cum[vcd.lb] - s_cum[vcd.la + 2]);
         }
         int tmax = blk ? mA : 1, nvmax = blk ? mB : 1, anyblk = blk;
#pragma unroll
         for (int s = 1; s < NG; s <<= 1) {       // iteration-mates = lanes that differ in the low log2(NG) bits
            tmax = max(tmax, __shfl_xor_sync(FULL, tmax, s));
            nvmax = max(nvmax, __shfl_xor_sync(FULL, nvmax, s));
            anyblk |= __shfl_xor_sync(FULL, anyblk, s);
         }
         p_pack = (unsigned)mA | ((unsigned)mB << 6) | ((unsigned)tmax << 12) | ((unsigned)nvmax << 18) | ((unsigned)blk << 24) |
                  ((unsigned)anyblk << 25);
      }
      int myfin = INF;
      // ---- phase 2 -----------------------------------------------------------------------------------
#pragma unroll 2
      for (int it = 0; it < G; ++it) {
         const int c = it * NG + g;
         const unsigned offA = __shfl_sync(FULL, vcd.offA, c), offB = __shfl_sync(FULL, vcd.offB, c);
         const int *blkp = Dv + __shfl_sync(FULL, p_rb, c);
         const unsigned pk = __shfl_sync(FULL, p_pack, c);
         const int mA = pk & 63, mB = (pk >> 6) & 63, tmaxAll = (pk >> 12) & 63, nvmax = (pk >> 18) & 63;
         const bool blk = (pk >> 24) & 1;
         int vA[NR], vB[NR];
#pragma unroll
         for (int r = 0; r < NR; ++r) {
            const int j = gl + r * G;
            vA[r] = j < mA ? __ldg(vc.vec + (offA + (unsigned)j)) : INF;      // one 32-bit add, then base + 4 * index
            vB[r] = j < mB ? __ldg(vc.vec + (offB + (unsigned)j)) : INF;
         }
         if (PF && it + 1 < G) {
            // tables beyond the caches: while this iteration's loads are in flight, pull the NEXT candidate's block rows
            // (one prefetch per lane = one row) and its two vectors into L1 - an iteration is otherwise one exposed
            // memory round trip after the other
            const int cn = c + NG;
            const unsigned pkn = __shfl_sync(FULL, p_pack, cn);
            const long long rbn = __shfl_sync(FULL, p_rb, cn);
            const unsigned oAn = __shfl_sync(FULL, vcd.offA, cn), oBn = __shfl_sync(FULL, vcd.offB, cn);
            if ((pkn >> 24) & 1) {
               const int mAn = pkn & 63, mBn = (pkn >> 6) & 63;
#pragma unroll
               for (int r = 0; r < NR; ++r)
                  if (gl + r * G < mAn) asm volatile("prefetch.global.L1 [%0];" ::"l"(Dv + rbn + (long long)((gl + r * G) * mBn)));
            }
            if (gl == 0) {
               asm volatile("prefetch.global.L1 [%0];" ::"l"(vc.vec + oAn));
               asm volatile("prefetch.global.L1 [%0];" ::"l"(vc.vec + oBn));
            }
         }
         if ((pk >> 25) & 1) {                                 // warp-uniform: some group of this iteration has a block
            const int nvalid = blk ? mB : 1;
            int acc[NR];
#pragma unroll
            for (int rd = 0; rd < NR; ++rd) {
               acc[rd] = INF;
               if (rd == 0 || nvmax > rd * G) {                 // warp-uniform
                  const unsigned cbase = (unsigned)min(gl + rd * G, nvalid - 1);
#pragma unroll
                  for (int rs = 0; rs < NR; ++rs) {
                     if (rs == 0 || tmaxAll > rs * G) {         // warp-uniform
                        const int tm = min(tmaxAll - rs * G, G);
                        const int opnd = vA[rs];
                        {
                           // eight unrolled steps at a time, rows unclamped - per step one
                           // IMAD.WIDE (row address), LDG, SHFL with an immediate lane and VIADDMNMX, the eight loads
                           // in flight together.  Rows beyond this group's block pair a sample that is INF (vA beyond
                           // mA) with whatever follows the block: table entries, products or the zeroed rest of the
                           // allocation, all in [0, 2^29), so INF + it neither wraps nor wins the min
                           const char *prow = reinterpret_cast<const char *>(blkp + (cbase + (unsigned)(rs * G * nvalid)));
                           const unsigned stride = 4u * (unsigned)nvalid;      // bytes per row
#pragma unroll
                           for (int t0 = 0; t0 < G; t0 += 8) {
                              if (t0 == 0 || t0 < tm) {                // warp-uniform
                                 int dd[8];
#pragma unroll
                                 for (int t = 0; t < 8; ++t) dd[t] = __ldg(reinterpret_cast<const int *>(prow + (unsigned long long)stride * (unsigned)(t0 + t)));
#pragma unroll
                                 for (int t = 0; t < 8; ++t) acc[rd] = __viaddmin_s32(__shfl_sync(FULL, opnd, t0 + t, G), dd[t], acc[rd]);
                              }
                           }
                        }
                     }
                  }
               }
            }
#pragma unroll
            for (int r = 0; r < NR; ++r) vA[r] = blk ? (gl + r * G < mB ? acc[r] : INF) : vA[r];
         }
         int fin = INF;
#pragma unroll
         for (int r = 0; r < NR; ++r) fin = min(fin, gl + r * G < mB ? vA[r] + vB[r] : INF);
#pragma unroll
         for (int s = G / 2; s > 0; s >>= 1) fin = min(fin, __shfl_xor_sync(FULL, fin, s, G));
         // candidate it * NG + g' was evaluated by group g' = lane % NG: hand it to its owner lane
         const int got = __shfl_sync(FULL, fin, (lane % NG) * G);
         myfin = (lane / NG == it) ? got : myfin;
      }
      // ---- phase 3 -----------------------------------------------------------------------------------
      if (slot < n) {
         double l, bound;
         unsigned f = 0;
         if (!vcd.ok) {
            l = CUDART_NAN; bound = CUDART_NAN; f = VNSGPU_F_BAD;
         } else {
            l = (double)myfin;
            bound = vcd.nolb ? l : P.minF[vcd.s0] + (double)vcd.lbsum + P.minB[vcd.sK];
            double dR = 0.0, lm = l;
            if (cd.x == VNSGPU_OP_ADD) {
               dR = cd.w >= 0 ? __ldg(I.reward + cd.w) : 0.0;      // vns_sop_solver.cpp:581
               lm = (L - 0.0) + (l - L);                           // :580,:583
            } else if (cd.x == VNSGPU_OP_REPLACE) {
               dR = cd.w >= 0 ? 0.0 + (__ldg(I.reward + cd.w) - __ldg(I.reward + lcl[cd.z + 1])) : 0.0;   // :784
            }
            if (lm <= budget) f |= VNSGPU_F_FEASIBLE;
            if (lm <= budget && (dR > 0 || (L - lm) > VNS_MIN_CHANGE_EPS)) f |= VNSGPU_F_ACCEPT;          // :619,:805
            if (bound > budget && !vcd.nolb) f |= VNSGPU_F_LB_SKIP;                                  // :539,...
         }
         if (len) len[kidx] = l;
         if (leni) leni[kidx] = (fmt & 4) ? (vcd.ok ? (int)((unsigned)myfin | (f << 29)) : -1) : (vcd.ok ? myfin : INT_MIN);
         if (lb) lb[kidx] = bound;
         if (flags) flags[kidx] = (uint8_t)f;
         if ((f & VNSGPU_F_ACCEPT) && !(skipflag && (f & VNSGPU_F_LB_SKIP)) && (unsigned long long)(kidx + idx0) < best)
            best = (unsigned long long)(kidx + idx0);
      }
   }
   if (d_first && best != ULLONG_MAX) atomicMin(d_first, best);
}


// ---- staged form for tables beyond the caches with small clusters (mp <= 16) ------------------------------------------
// k_eval_vec<.,.,PF> gathers every candidate's block and vectors with plain loads + L1 prefetches: on the 20 k-node table
// (blocks of 400 B, vectors of 40 B, 5 GB of table + products + vectors) it is a chain of exposed HBM round trips (ncu
// r2i: 23 % of the measured copy bandwidth, eligible warps 0.8 / cycle).  Here lane L owns candidate base + L through
// the memory phase: it issues the copies of ITS block and two vectors as (at most three) bulk copies - cp.async.bulk,
// completion on the warp's mbarrier - into its own shared-memory slot (start rounded down to the 16-byte line, the data
// keeps its offset inside the slot), so one warp has 32 candidates' bytes in flight from 3 instructions per lane and 8
// warps per SM keep ~140 KB outstanding; after one wait the groups evaluate out of shared memory.  Slots start zeroed and only ever receive table / product / vector words (all in [0, INF]), so the
// unclamped rows of the block product pair INF with numbers that neither wrap nor win.
__host__ __device__ inline int vecst_slot_ints(int mp)
{
   const int v = ((mp * 4 + 15) & ~15) + 16, b = ((mp * mp * 4 + 15) & ~15) + 16;      // bytes: vector, block (+ alignment slack)
   return (2 * v + b) / 4 + 4;
}

template <int G, int NR>
__global__ void __launch_bounds__(VNS_EVAL_THREADS)
k_eval_vec_st(InstDev I, const int *__restrict__ Dv, PathDev P, SegCache seg, VecCache vc, long long n, long long idx0,
              const void *__restrict__ cands, const unsigned *__restrict__ perm, double budget, int skipflag,
              double *__restrict__ len, double *__restrict__ lb, uint8_t *__restrict__ flags, unsigned long long *d_first,
              int fmt, int *__restrict__ leni, int meta_ints)
{
   extern __shared__ __align__(16) int smem_i[];
   constexpr unsigned FULL = 0xffffffffu;
   constexpr int INF = VNS_INT_INF;
   const int T = P.hdr[HDR_T], C = I.C, mp = I.mp;
   int *lcl = smem_i;
   int2 *cinfo = reinterpret_cast<int2 *>(smem_i + ((T + 2 + 1) & ~1));
   for (int l = threadIdx.x; l < T + 2; l += blockDim.x) lcl[l] = layer_cluster(I, P.tg, T, l);
   for (int c = threadIdx.x; c < C; c += blockDim.x) cinfo[c] = make_int2(I.off[c], I.m[c]);
   long long *s_rowoff = reinterpret_cast<long long *>(cinfo + C);
   int *s_cum = reinterpret_cast<int *>(s_rowoff + (T + 2)), *s_pd = s_cum + (T + 3);
   for (int l = threadIdx.x; l < T + 3; l += blockDim.x) s_cum[l] = seg.cum[l];
   for (int l = threadIdx.x; l < T + 2; l += blockDim.x) { s_pd[l] = seg.pd[l]; s_rowoff[l] = seg.rowoff[l]; }
   // staging slots behind the layer maps: [warp][candidate lane][slot]
   const int slot_ints = vecst_slot_ints(mp);
   const int vbytes = ((mp * 4 + 15) & ~15) + 16;
   int *stage_all = smem_i + meta_ints;
   for (int e = threadIdx.x; e < (int)(blockDim.x >> 5) * 32 * slot_ints; e += blockDim.x) stage_all[e] = 0;
   // one mbarrier per warp: every lane arrives once per batch with the bytes of its own copies (expect_tx); the bulk copies
   // complete on it (complete_tx), so the phase ends when all 32 lanes' bytes have landed
   __shared__ __align__(8) unsigned long long s_bar[VNS_EVAL_THREADS / 32];
   const unsigned bar = (unsigned)__cvta_generic_to_shared(&s_bar[threadIdx.x >> 5]);
   if ((threadIdx.x & 31) == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], 32;" ::"r"(bar));
   asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // the zeroed slots and the barrier, before the async proxy writes
   __syncthreads();
   int *wstage = stage_all + (threadIdx.x >> 5) * 32 * slot_ints;
   unsigned parity = 0;

   constexpr int NG = 32 / G;
   const int lane = threadIdx.x & 31, gl = lane % G, g = lane / G;
   const double L = P.hd[HD_LEN];
   const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
   const long long warp0 = (((long long)blockIdx.x * blockDim.x) >> 5) + (threadIdx.x >> 5);
   unsigned long long best = ULLONG_MAX;

   for (long long base = warp0 * 32; base < n; base += nwarps * 32) {
      // ---- phase 1: lane-parallel fetch + decode ------------------------------------------------------------------
      const long long slot = base + lane;
      long long kidx = slot;
      if (perm && slot < n) kidx = __ldg(perm + slot);
      int4 cd = make_int4(-1, 0, 0, -1);
      if (slot < n) {
         if ((fmt & 3) == 2) {
            const unsigned w = __ldg(reinterpret_cast<const unsigned *>(cands) + kidx);
            cd.x = (int)(w & 7); cd.y = (int)((w >> 3) & 0x3fff); cd.z = (int)((w >> 17) & 0x3fff); cd.w = (w >> 31) ? cd.y : -1;
         } else
            cd = __ldg(reinterpret_cast<const int4 *>(cands) + kidx);
      }
      if (cd.w >= C) cd.x = -1;
      const VecCode vcd = decode_cand_vec(vc, T, C, mp, lcl, [&](int l) { return s_cum[l + 1] - s_cum[l]; },
                                          [&](int c) { return cinfo[c].y; }, [&](int l) { return s_pd[l]; }, cd.x, cd.y, cd.z);
      long long p_rb = 0;
      const int mA = vcd.mA, mB = vcd.mB;
      const bool blk = vcd.ok && vcd.lb > vcd.la;
      if (blk) {
         if (vcd.lb == vcd.la + 1) {
            const int2 cia = cinfo[lcl[vcd.la]], cib = cinfo[lcl[vcd.lb]];
            p_rb = (long long)cia.x * I.N + (long long)cia.y * cib.x;
         } else
            p_rb = s_rowoff[vcd.la] + (long long)mA * (s_cum[vcd.lb] - s_cum[vcd.la + 2]);
      }
      // ---- memory phase: this lane's candidate -> this lane's slot ---------------------------------------------------
      char *myslot = reinterpret_cast<char *>(wstage + lane * slot_ints);
      unsigned shifts = 0;
      {
         const unsigned d0 = (unsigned)__cvta_generic_to_shared(myslot);
         // 16-byte aligned enclosing range of [src, src + nints): start rounded down, length rounded up; the data keeps its
         // offset inside the slot
         const int *srcA = vc.vec + vcd.offA, *srcB = vc.vec + vcd.offB, *srcM = Dv + p_rb;
         const unsigned shA = (unsigned)((size_t)srcA & 15), shB = (unsigned)((size_t)srcB & 15), shM = (unsigned)((size_t)srcM & 15);
         const unsigned nbA = vcd.ok ? ((shA + (unsigned)mA * 4u + 15u) & ~15u) : 0u, nbB = vcd.ok ? ((shB + (unsigned)mB * 4u + 15u) & ~15u) : 0u;
         const unsigned nbM = blk ? ((shM + (unsigned)(mA * mB) * 4u + 15u) & ~15u) : 0u;
         asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(nbA + nbB + nbM) : "memory");
         if (nbA) asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                               ::"r"(d0), "l"(reinterpret_cast<const char *>(srcA) - shA), "r"(nbA), "r"(bar) : "memory");
         if (nbB) asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                               ::"r"(d0 + vbytes), "l"(reinterpret_cast<const char *>(srcB) - shB), "r"(nbB), "r"(bar) : "memory");
         if (nbM) asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                               ::"r"(d0 + 2 * vbytes), "l"(reinterpret_cast<const char *>(srcM) - shM), "r"(nbM), "r"(bar) : "memory");
         shifts = shA | (shB << 4) | (shM << 8);
      }
      unsigned p_pack;
      {
         int tmax = blk ? mA : 1, nvmax = blk ? mB : 1, anyblk = blk;
#pragma unroll
         for (int sft = 1; sft < NG; sft <<= 1) {
            tmax = max(tmax, __shfl_xor_sync(FULL, tmax, sft));
            nvmax = max(nvmax, __shfl_xor_sync(FULL, nvmax, sft));
            anyblk |= __shfl_xor_sync(FULL, anyblk, sft);
         }
         p_pack = (unsigned)mA | ((unsigned)mB << 6) | ((unsigned)tmax << 12) | ((unsigned)nvmax << 18) | ((unsigned)blk << 24) |
                  ((unsigned)anyblk << 25);
      }
      {
         unsigned done = 0;
         while (!done)
            asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
         parity ^= 1;
      }
      int myfin = INF;
      // ---- phase 2: G iterations, group g evaluates candidate it * NG + g out of shared memory -----------------------------
#pragma unroll 2
      for (int it = 0; it < G; ++it) {
         const int c = it * NG + g;
         const unsigned pk = __shfl_sync(FULL, p_pack, c), sh = __shfl_sync(FULL, shifts, c);
         const int mAc = pk & 63, mBc = (pk >> 6) & 63, tmaxAll = (pk >> 12) & 63, nvmax = (pk >> 18) & 63;
         const bool blkc = (pk >> 24) & 1;
         const char *cs = reinterpret_cast<const char *>(wstage + c * slot_ints);
         const int *sA = reinterpret_cast<const int *>(cs + (sh & 15)), *sB = reinterpret_cast<const int *>(cs + vbytes + ((sh >> 4) & 15));
         const int *sM = reinterpret_cast<const int *>(cs + 2 * vbytes + ((sh >> 8) & 15));
         int vA[NR], vB[NR];
#pragma unroll
         for (int r = 0; r < NR; ++r) {
            const int j = gl + r * G;
            vA[r] = j < mAc ? sA[j] : INF;
            vB[r] = j < mBc ? sB[j] : INF;
         }
         if ((pk >> 25) & 1) {                                 // warp-uniform: some group of this iteration has a block
            const int nvalid = blkc ? mBc : 1;
            int acc[NR];
#pragma unroll
            for (int rd = 0; rd < NR; ++rd) {
               acc[rd] = INF;
               if (rd == 0 || nvmax > rd * G) {                 // warp-uniform
                  const int cb = min(gl + rd * G, nvalid - 1);
#pragma unroll
                  for (int rs = 0; rs < NR; ++rs) {
                     if (rs == 0 || tmaxAll > rs * G) {         // warp-uniform
                        const int tm = min(tmaxAll - rs * G, G);
                        const int opnd = vA[rs];
                        const int *prow = sM + cb + rs * G * nvalid;
#pragma unroll
                        for (int t0 = 0; t0 < G; t0 += 8) {
                           if (t0 == 0 || t0 < tm) {            // warp-uniform
                              int dd[8];
#pragma unroll
                              for (int t = 0; t < 8; ++t) dd[t] = blkc ? prow[nvalid * min(t0 + t, max(mAc - rs * G, 1) - 1)] : 0;
#pragma unroll
                              for (int t = 0; t < 8; ++t) acc[rd] = __viaddmin_s32(__shfl_sync(FULL, opnd, min(t0 + t, max(mAc - rs * G, 1) - 1), G), dd[t], acc[rd]);
                           }
                        }
                     }
                  }
               }
            }
#pragma unroll
            for (int r = 0; r < NR; ++r) vA[r] = blkc ? (gl + r * G < mBc ? acc[r] : INF) : vA[r];
         }
         int fin = INF;
#pragma unroll
         for (int r = 0; r < NR; ++r) fin = min(fin, gl + r * G < mBc ? vA[r] + vB[r] : INF);
#pragma unroll
         for (int sft = G / 2; sft > 0; sft >>= 1) fin = min(fin, __shfl_xor_sync(FULL, fin, sft, G));
         const int got = __shfl_sync(FULL, fin, (lane % NG) * G);
         myfin = (lane / NG == it) ? got : myfin;
      }
      __syncwarp();      // the slots are overwritten by the next batch: order this batch's reads before the async proxy's writes
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      // ---- phase 3: lane-parallel epilogue ----------------------------------------------------------------------------------
      if (slot < n) {
         double l, bound;
         unsigned f = 0;
         if (!vcd.ok) {
            l = CUDART_NAN; bound = CUDART_NAN; f = VNSGPU_F_BAD;
         } else {
            l = (double)myfin;
            bound = vcd.nolb ? l : P.minF[vcd.s0] + (double)vcd.lbsum + P.minB[vcd.sK];
            double dR = 0.0, lm = l;
            if (cd.x == VNSGPU_OP_ADD) {
               dR = cd.w >= 0 ? __ldg(I.reward + cd.w) : 0.0;      // vns_sop_solver.cpp:581
               lm = (L - 0.0) + (l - L);                           // :580,:583
            } else if (cd.x == VNSGPU_OP_REPLACE) {
               dR = cd.w >= 0 ? 0.0 + (__ldg(I.reward + cd.w) - __ldg(I.reward + lcl[cd.z + 1])) : 0.0;   // :784
            }
            if (lm <= budget) f |= VNSGPU_F_FEASIBLE;
            if (lm <= budget && (dR > 0 || (L - lm) > VNS_MIN_CHANGE_EPS)) f |= VNSGPU_F_ACCEPT;          // :619,:805
            if (bound > budget && !vcd.nolb) f |= VNSGPU_F_LB_SKIP;                                  // :539,...
         }
         if (len) len[kidx] = l;
         if (leni) leni[kidx] = (fmt & 4) ? (vcd.ok ? (int)((unsigned)myfin | (f << 29)) : -1) : (vcd.ok ? myfin : INT_MIN);
         if (lb) lb[kidx] = bound;
         if (flags) flags[kidx] = (uint8_t)f;
         if ((f & VNSGPU_F_ACCEPT) && !(skipflag && (f & VNSGPU_F_LB_SKIP)) && (unsigned long long)(kidx + idx0) < best)
            best = (unsigned long long)(kidx + idx0);
      }
   }
   if (d_first && best != ULLONG_MAX) atomicMin(d_first, best);
}
