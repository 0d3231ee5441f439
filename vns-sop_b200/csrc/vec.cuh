// Endpoint vectors (integer tables only): every tryTo* as  min_{s,t} A[s] + M[s][t] + B[t].
//
// The reference already shares work between candidates through the prefix / suffix vectors F[l], B[l]
// (vns_sop_path.cpp:1426-1539).  With integer entries min-plus products are exactly associative, so the
// same idea extends to the two ends of every candidate's chain: they depend only on (cluster, layer), not on
// the candidate, and there are C x (T+2) of them against millions of candidates per batch.  Per path state
// one pass (k_vec_build) stores, for every cluster c and layer l (lcl = layer -> cluster, S = T + 1):
//   InsF(c,l) = F[l-1] (x) D(lcl[l-1],c) (x) D(c,lcl[l])              at layer l      c inserted before l
//   InsB(c,l) = D(lcl[l-1],c) (x) D(c,lcl[l]) (x) B[l]                at layer l-1
//   RepF(c,l) = F[l-1] (x) D(lcl[l-1],c) (x) D(c,lcl[l+1])            at layer l+1    layer l replaced by c
//   RepB(c,l) = D(lcl[l-1],c) (x) D(c,lcl[l+1]) (x) B[l+1]            at layer l-1
//   RemF(l)   = F[l-1] (x) D(lcl[l-1],lcl[l+1])                       at layer l+1    layer l removed
//   RemB(l)   = D(lcl[l-1],lcl[l+1]) (x) B[l+1]                       at layer l-1
// plus int copies FI / BI of F / B and the matching sums of cluster-pair minima for the lower bounds.  Then
//   tryToAdd(c,idx)      = FI[idx] . InsB(c,idx+1)                    (vns_sop_path.cpp:245-304)
//   tryToReplace(c,idx)  = FI[idx] . RepB(c,idx+1)                    (:456-513)
//   tryToRemove(idx)     = FI[idx] . RemB(idx+1)                      (:515-535)
//   tryToMove up   (i=a+1 -> before j=b+1)  = RemF(i) (x) M(i+1,j-1) . InsB(lcl[i],j)      (:932-1154)
//   tryToMove down (j=a+1 -> before i=b+1)  = InsF(lcl[j],i) (x) M(i,j-1) . RemB(j)
//   tryToExchange  (layers p=i+1, q=j)      = RepF(lcl[q],p) (x) M(p+1,q-1) . RepB(lcl[p],q)  (:635-797)
// where "." is min_s x[s] + y[s], M(a,b) is the segment product of eval.cuh (a table block when b == a+1,
// absent when b == a).  One candidate = at most ONE block step.  Lengths, lower bounds and flags are
// bit-identical to the block-by-block evaluation (tests/test_gpu_parity.py runs all three forms).
#pragma once
#include "eval.cuh"

struct VecCache {
   const int *vec;   // ((fam * C + c) * LS + l) * mp, fam 0..3 = InsF, InsB, RepF, RepB; then rows RemF, RemB, FI, BI
   const int *lbs;   // insLB[C][LS], repLB[C][LS], remLB[LS]
   int LS;           // layer stride (>= S + 1)
};

__device__ __forceinline__ unsigned vec_off(const VecCache &vc, int C, int mp, int fam, int c, int l)
{
   return (unsigned)(((fam * C + c) * vc.LS + l) * mp);
}
__device__ __forceinline__ unsigned vec_row(const VecCache &vc, int C, int mp, int k, int l)   // k: 0 RemF, 1 RemB, 2 FI, 3 BI
{
   return (unsigned)(((4 * C + k) * vc.LS + l) * mp);
}
__host__ __device__ inline size_t vec_ints(int C, int LS, int mp) { return (size_t)(4 * C + 4) * LS * mp; }
__host__ __device__ inline size_t vec_lb_ints(int C, int LS) { return (size_t)(2 * C + 1) * LS; }

// lane = output index; v (lane-held, INF beyond the source size) times block (ca -> cb)
__device__ __forceinline__ int vec_fwd(const InstDev &I, const int *__restrict__ Dv, int v, int ca, int cb, int lane)
{
   const int ma = I.m[ca], mb = I.m[cb];
   const int *blk = Dv + ((long long)I.off[ca] * I.N + (long long)ma * I.off[cb]);
   int acc = VNS_INT_INF;
   const int j = min(lane, mb - 1);
   for (int s = 0; s < ma; ++s) acc = __viaddmin_s32(__shfl_sync(0xffffffffu, v, s), __ldg(blk + s * mb + j), acc);
   return lane < mb ? acc : VNS_INT_INF;
}
// lane = source index; block (ca -> cb) times v (lane-held over cb's samples)
__device__ __forceinline__ int vec_bwd(const InstDev &I, const int *__restrict__ Dv, int v, int ca, int cb, int lane)
{
   const int ma = I.m[ca], mb = I.m[cb];
   const int *blk = Dv + ((long long)I.off[ca] * I.N + (long long)ma * I.off[cb]);
   int acc = VNS_INT_INF;
   const int s = min(lane, ma - 1);
   for (int j = 0; j < mb; ++j) acc = __viaddmin_s32(__shfl_sync(0xffffffffu, v, j), __ldg(blk + s * mb + j), acc);
   return lane < ma ? acc : VNS_INT_INF;
}

// one warp per (c, l); row c == C holds RemF / RemB / FI / BI.  Needs mp <= 32.
__global__ void __launch_bounds__(256) k_vec_build(InstDev I, PathDev P, const int *__restrict__ Dv, int *__restrict__ vec,
                                                   int *__restrict__ lbs, int LS)
{
   const int T = P.hdr[HDR_T], S = T + 1, C = I.C, mp = I.mp;
   const int lane = threadIdx.x & 31;
   const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
   if (w >= (long long)(C + 1) * (S + 1)) return;
   const int c = (int)(w / (S + 1)), l = (int)(w % (S + 1));
   VecCache vc{vec, lbs, LS};
   auto cl = [&](int x) { return layer_cluster(I, P.tg, T, x); };
   auto Fv = [&](int x) { return lane < I.m[cl(x)] ? (int)P.F[(size_t)x * mp + lane] : VNS_INT_INF; };
   auto Bv = [&](int x) { return lane < I.m[cl(x)] ? (int)P.B[(size_t)x * mp + lane] : VNS_INT_INF; };
   auto dm = [&](int a, int b) { return (int)I.dmin[(size_t)a * C + b]; };
   auto put = [&](unsigned off, int v) { if (lane < mp) vec[off + lane] = v; };
   if (c == C) {
      put(vec_row(vc, C, mp, 2, l), Fv(l));
      put(vec_row(vc, C, mp, 3, l), Bv(l));
      if (l >= 1 && l <= S - 1) {
         const int a = cl(l - 1), b = cl(l + 1);
         put(vec_row(vc, C, mp, 0, l), vec_fwd(I, Dv, Fv(l - 1), a, b, lane));
         put(vec_row(vc, C, mp, 1, l), vec_bwd(I, Dv, Bv(l + 1), a, b, lane));
         if (lane == 0) lbs[(size_t)2 * C * LS + l] = dm(a, b);
      }
      return;
   }
   if (l < 1) return;
   const int a = cl(l - 1);
   {  // insert c between layers l-1 and l
      const int b = cl(l);
      put(vec_off(vc, C, mp, 0, c, l), vec_fwd(I, Dv, vec_fwd(I, Dv, Fv(l - 1), a, c, lane), c, b, lane));
      put(vec_off(vc, C, mp, 1, c, l), vec_bwd(I, Dv, vec_bwd(I, Dv, Bv(l), c, b, lane), a, c, lane));
      if (lane == 0) lbs[(size_t)c * LS + l] = dm(a, c) + dm(c, b);
   }
   if (l <= S - 1) {   // c replaces layer l
      const int b = cl(l + 1);
      put(vec_off(vc, C, mp, 2, c, l), vec_fwd(I, Dv, vec_fwd(I, Dv, Fv(l - 1), a, c, lane), c, b, lane));
      put(vec_off(vc, C, mp, 3, c, l), vec_bwd(I, Dv, vec_bwd(I, Dv, Bv(l + 1), c, b, lane), a, c, lane));
      if (lane == 0) lbs[(size_t)(C + c) * LS + l] = dm(a, c) + dm(c, b);
   }
}

// decoded candidate in endpoint-vector form
struct VecCode {
   int ok;            // 0: indices out of range
   unsigned offA, offB;
   int la, lb;        // layers of A and B (block la -> lb when lb > la)
   int lbsum;         // dist_between of the lower bound
   int s0, sK;        // layers whose minF / minB enter the lower bound
   int nolb;
};

template <typename PD>
__device__ __forceinline__ VecCode decode_cand_vec(const VecCache &vc, int T, int C, int mp, const int *lcl, PD pd, int op, int a, int b)
{
   VecCode v;
   v.ok = 0; v.offA = v.offB = 0; v.la = v.lb = 0; v.lbsum = 0; v.s0 = v.sK = 0; v.nolb = 0;
   const int *insLB = vc.lbs, *repLB = vc.lbs + (size_t)C * vc.LS, *remLB = vc.lbs + (size_t)2 * C * vc.LS;
   if (op == VNSGPU_OP_ADD) {
      if (b < 0 || b > T || a < 0 || a >= C) return v;
      v.ok = 1; v.la = v.lb = b; v.offA = vec_row(vc, C, mp, 2, b); v.offB = vec_off(vc, C, mp, 1, a, b + 1);
      v.lbsum = __ldg(insLB + (size_t)a * vc.LS + b + 1); v.s0 = b; v.sK = b + 1;
   } else if (op == VNSGPU_OP_REPLACE) {
      if (b < 0 || b >= T || a < 0 || a >= C) return v;
      v.ok = 1; v.la = v.lb = b; v.offA = vec_row(vc, C, mp, 2, b); v.offB = vec_off(vc, C, mp, 3, a, b + 1);
      v.lbsum = __ldg(repLB + (size_t)a * vc.LS + b + 1); v.s0 = b; v.sK = b + 2;
   } else if (op == VNSGPU_OP_REMOVE) {
      if (a < 0 || a >= T) return v;
      v.ok = 1; v.la = v.lb = a; v.offA = vec_row(vc, C, mp, 2, a); v.offB = vec_row(vc, C, mp, 1, a + 1);
      v.s0 = a; v.sK = a + 2; v.nolb = 1;
   } else if (op == VNSGPU_OP_MOVE) {
      if (a >= T || a < 0 || b > T || b < 0 || abs(b - a) < 2) return v;
      const int i = min(a, b) + 1, j = max(a, b) + 1;
      v.ok = 1;
      if (b > a) {   // layer i moves up, in front of layer j
         v.la = i + 1; v.lb = j - 1; v.offA = vec_row(vc, C, mp, 0, i); v.offB = vec_off(vc, C, mp, 1, lcl[i], j);
         v.lbsum = __ldg(remLB + i) + (pd(j - 1) - pd(i + 1)) + __ldg(insLB + (size_t)lcl[i] * vc.LS + j);
         v.s0 = i - 1; v.sK = j;
      } else {       // layer j moves down, in front of layer i
         v.la = i; v.lb = j - 1; v.offA = vec_off(vc, C, mp, 0, lcl[j], i); v.offB = vec_row(vc, C, mp, 1, j);
         v.lbsum = __ldg(insLB + (size_t)lcl[j] * vc.LS + i) + (pd(j - 1) - pd(i)) + __ldg(remLB + j);
         v.s0 = i - 1; v.sK = j + 1;
      }
   } else if (op == VNSGPU_OP_EXCHANGE) {
      if (a >= T || a < 0 || b >= T || b < 0 || abs(a - b) < 2) return v;
      const int p = min(a, b) + 1, q = max(a, b) + 1;   // the two exchanged layers
      v.ok = 1; v.la = p + 1; v.lb = q - 1;
      v.offA = vec_off(vc, C, mp, 2, lcl[q], p); v.offB = vec_off(vc, C, mp, 3, lcl[p], q);
      v.lbsum = __ldg(repLB + (size_t)lcl[q] * vc.LS + p) + (pd(q - 1) - pd(p + 1)) + __ldg(repLB + (size_t)lcl[p] * vc.LS + q);
      v.s0 = p - 1; v.sK = q + 1;
   }
   return v;
}

template <int G, int NR>
__global__ void __launch_bounds__(VNS_EVAL_THREADS)
k_eval_vec(InstDev I, const int *__restrict__ Dv, PathDev P, SegCache seg, VecCache vc, long long n, long long idx0,
           const vnsgpu_cand *__restrict__ cands, const unsigned *__restrict__ perm, double budget, int skipflag,
           double *__restrict__ len, double *__restrict__ lb, uint8_t *__restrict__ flags, unsigned long long *d_first)
{
   extern __shared__ int smem_i[];
   constexpr unsigned FULL = 0xffffffffu;
   constexpr int INF = VNS_INT_INF;
   const int T = P.hdr[HDR_T], C = I.C, mp = I.mp;
   int *lcl = smem_i;
   int2 *cinfo = reinterpret_cast<int2 *>(smem_i + ((T + 2 + 1) & ~1));
   for (int l = threadIdx.x; l < T + 2; l += blockDim.x) lcl[l] = layer_cluster(I, P.tg, T, l);
   for (int c = threadIdx.x; c < C; c += blockDim.x) cinfo[c] = make_int2(I.off[c], I.m[c]);
   int *s_cum = reinterpret_cast<int *>(cinfo + C), *s_pd = s_cum + (T + 3);
   unsigned *s_rowoff = reinterpret_cast<unsigned *>(s_pd + (T + 2));
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
      if (slot < n) cd = __ldg(reinterpret_cast<const int4 *>(cands) + kidx);
      const VecCode vcd = decode_cand_vec(vc, T, C, mp, lcl, [&](int l) { return s_pd[l]; }, cd.x, cd.y, cd.z);
      unsigned p_rb = 0, p_pack;
      {
         const int mA = vcd.ok ? s_cum[vcd.la + 1] - s_cum[vcd.la] : 0, mB = vcd.ok ? s_cum[vcd.lb + 1] - s_cum[vcd.lb] : 0;
         const bool blk = vcd.ok && vcd.lb > vcd.la;
         if (blk) {   // block la -> lb: a table block when adjacent, else the segment product M(la, lb)
            if (vcd.lb == vcd.la + 1) {
               const int2 cia = cinfo[lcl[vcd.la]], cib = cinfo[lcl[vcd.lb]];
               p_rb = (unsigned)cia.x * (unsigned)I.N + (unsigned)cia.y * (unsigned)cib.x;
            } else
               p_rb = s_rowoff[vcd.la] + (unsigned)(mA * (s_cum[vcd.lb] - s_cum[vcd.la + 2]));
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
         const unsigned rb = __shfl_sync(FULL, p_rb, c), pk = __shfl_sync(FULL, p_pack, c);
         const int mA = pk & 63, mB = (pk >> 6) & 63, tmaxAll = (pk >> 12) & 63, nvmax = (pk >> 18) & 63;
         const bool blk = (pk >> 24) & 1;
         int vA[NR], vB[NR];
#pragma unroll
         for (int r = 0; r < NR; ++r) {
            const int j = gl + r * G;
            vA[r] = j < mA ? __ldg(vc.vec + offA + j) : INF;
            vB[r] = j < mB ? __ldg(vc.vec + offB + j) : INF;
         }
         if ((pk >> 25) & 1) {                                 // warp-uniform: some group of this iteration has a block
            const int trip = blk ? mA : 1, nvalid = blk ? mB : 1;
            int acc[NR];
#pragma unroll
            for (int rd = 0; rd < NR; ++rd) {
               acc[rd] = INF;
               if (rd == 0 || nvmax > rd * G) {                 // warp-uniform
                  const unsigned cbase = rb + (unsigned)min(gl + rd * G, nvalid - 1);
#pragma unroll
                  for (int rs = 0; rs < NR; ++rs) {
                     if (rs == 0 || tmaxAll > rs * G) {         // warp-uniform
                        const int tlim = max(min(trip - rs * G, G) - 1, 0);
                        const unsigned sb = cbase + (unsigned)(min(rs * G, trip - 1) * nvalid);
                        const int tm = min(tmaxAll - rs * G, G);
                        const int opnd = vA[rs];
#pragma unroll 2
                        for (int t = 0; t < tm; ++t) {
                           const int tt = min(t, tlim);
                           acc[rd] = __viaddmin_s32(__shfl_sync(FULL, opnd, tt, G), __ldg(Dv + (sb + (unsigned)(tt * nvalid))), acc[rd]);
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
         if (lb) lb[kidx] = bound;
         if (flags) flags[kidx] = (uint8_t)f;
         if ((f & VNSGPU_F_ACCEPT) && !(skipflag && (f & VNSGPU_F_LB_SKIP)) && (unsigned long long)(kidx + idx0) < best)
            best = (unsigned long long)(kidx + idx0);
      }
   }
   if (d_first && best != ULLONG_MAX) atomicMin(d_first, best);
}
