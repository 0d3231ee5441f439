// Distance-table kernels.
//
// k_table_pairs<KIND> is the whole table build in ONE pass: a warp owns one ordered cluster pair (c1, c2) =
// one contiguous m1 x m2 block of the block-major table (consecutive warps own consecutive blocks of the same
// row band, so the stores of a CTA form one contiguous run).  Per entry it computes the distance (Euclid /
// Dubins / gather from an explicit matrix), stores it as double AND - when the caller passed a buffer - as
// int32 for the exact integer path, and folds it into the pair's min / max (calcClusterDistancesBounds) and
// into the table's classification (all entries non-negative integers? largest entry).  HBM traffic: 8 N^2
// (+ 4 N^2) bytes written, nothing read back; the separate bounds / classification / conversion passes of the
// first version (8 N^2 read twice) are gone.
// k_table_gather (to_dense = 1) only serves vnsgpu_get_table_dense (tests).
#pragma once
#include "common.cuh"

struct BandPos {
   int c2, i, j, m2;
};

__device__ __forceinline__ BandPos band_decode(const InstDev &I, int m1, int q)
{
   BandPos p;
   p.c2 = __ldg(I.clof + q / m1);
   p.m2 = __ldg(I.m + p.c2);
   int r = q - m1 * __ldg(I.off + p.c2);
   p.i = r / p.m2;
   p.j = r - p.i * p.m2;
   return p;
}

#define VNS_TWO_PI 6.283185307179586476925286766559
// a CSC word whose straight segment has length 0 is feasible; rounding in sin/cos can push p^2 a few
// ulps below 0 there, so p^2 in [-eps, 0) is clamped to 0 instead of dropping the word
#define VNS_DUBINS_EPS 1e-9

__device__ __forceinline__ double mod2pi(double t) { return t - VNS_TWO_PI * floor(t / VNS_TWO_PI); }

// Shortest Dubins path (six words, normalised by rho).  No reference code exists;
// the known answers are the EXPLICIT matrices of datasets/dop_sop_dataset.
// Operation order mirrors oracle/vns_oracle.c:vo_dubins_length term by term.
static __device__ double dubins_length(double x0, double y0, double th0, double x1, double y1, double th1, double rho)
{
   double dx = x1 - x0, dy = y1 - y0;
   double d = sqrt(dx * dx + dy * dy) / rho;
   double phi = d > 0 ? atan2(dy, dx) : 0.0;
   double al = mod2pi(th0 - phi), be = mod2pi(th1 - phi);
   double sa, ca, sb, cb;
   sincos(al, &sa, &ca);
   sincos(be, &sb, &cb);
   double cab = cos(al - be);
   double best = CUDART_INF, psq, p, tau, u, t;
   psq = 2 + d * d - 2 * cab + 2 * d * (sa - sb);  // LSL
   if (psq >= -VNS_DUBINS_EPS) {
      psq = psq > 0 ? psq : 0;
      tau = atan2(cb - ca, d + sa - sb);
      t = mod2pi(tau - al) + sqrt(psq) + mod2pi(be - tau);
      best = t < best ? t : best;
   }
   psq = 2 + d * d - 2 * cab + 2 * d * (sb - sa);  // RSR
   if (psq >= -VNS_DUBINS_EPS) {
      psq = psq > 0 ? psq : 0;
      tau = atan2(ca - cb, d - sa + sb);
      t = mod2pi(al - tau) + sqrt(psq) + mod2pi(tau - be);
      best = t < best ? t : best;
   }
   psq = -2 + d * d + 2 * cab + 2 * d * (sa + sb);  // LSR
   if (psq >= -VNS_DUBINS_EPS) {
      psq = psq > 0 ? psq : 0;
      p = sqrt(psq);
      tau = atan2(-ca - cb, d + sa + sb) - atan2(-2.0, p);
      t = mod2pi(tau - al) + p + mod2pi(tau - mod2pi(be));
      best = t < best ? t : best;
   }
   psq = -2 + d * d + 2 * cab - 2 * d * (sa + sb);  // RSL
   if (psq >= -VNS_DUBINS_EPS) {
      psq = psq > 0 ? psq : 0;
      p = sqrt(psq);
      tau = atan2(ca + cb, d - sa - sb) - atan2(2.0, p);
      t = mod2pi(al - tau) + p + mod2pi(be - tau);
      best = t < best ? t : best;
   }
   u = (6 - d * d + 2 * cab + 2 * d * (sa - sb)) / 8;  // RLR
   if (fabs(u) <= 1) {
      p = mod2pi(VNS_TWO_PI - acos(u));
      t = mod2pi(al - atan2(ca - cb, d - sa + sb) + p / 2);
      t = t + p + mod2pi(al - be - t + p);
      best = t < best ? t : best;
   }
   u = (6 - d * d + 2 * cab + 2 * d * (sb - sa)) / 8;  // LRL
   if (fabs(u) <= 1) {
      p = mod2pi(VNS_TWO_PI - acos(u));
      t = mod2pi(-al - atan2(ca - cb, d + sa - sb) + p / 2);
      t = t + p + mod2pi(mod2pi(be) - al - t + mod2pi(p));
      best = t < best ? t : best;
   }
   return best * rho;
}

// a2: SOPLoader::getSOPFromWeightMatrix sop_loader.cpp:413-427 (W in sample order); to_dense = 1
// runs the inverse permutation (download for tests)
static __global__ void __launch_bounds__(256) k_table_gather(InstDev I, double *__restrict__ D, double *__restrict__ W, int to_dense)
{
   const int c1 = blockIdx.y;
   const int m1 = I.m[c1], o1 = I.off[c1];
   const int band = m1 * I.N;
   double *blk = D + (long long)o1 * I.N;
   for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < band; q += gridDim.x * blockDim.x) {
      BandPos p = band_decode(I, m1, q);
      long long w = (long long)(o1 + p.i) * I.N + I.off[p.c2] + p.j;
      if (to_dense) W[w] = blk[q];
      else blk[q] = W[w];
   }
}

struct TableJob {
   const double2 *pts;   // KIND 0: sample coordinates (sample order); KIND 1: cluster centres
   const double *W;      // KIND 2: explicit N x N matrix, row-major in sample order
   int rule;             // KIND 0: VNSGPU_RULE_*
   int S;                // KIND 1: heading samples per cluster
   double rho;           // KIND 1: turning radius
   int do_ceil;          // KIND 1
   double *D;            // block-major table (out)
   int *Di;              // its int32 copy (out) or null
   unsigned short *D16;  // its uint16 copy (out, saturated) or null: halves the bytes the state builds stream (vec.cuh)
   double *dmin, *dmax;  // C x C (out)
   unsigned long long *info;   // [0] += CTAs that saw a non-integer or negative entry, [1] = max over the bit patterns
                               // of the CTAs' largest entries (entries >= 0 order like their bit patterns)
};

// KIND 0 = a1/a3: SOPLoader::getSOPFromNodes sop_loader.cpp:457-491, GraphNode::distanceTo heuristic_types.h:67-75
//          (sqrt(dx*dx+dy*dy), no FMA contraction; rule: EUC_2D round :476, CEIL_2D ceil :480), getOP/getOPN :278,360 (raw)
// KIND 1 = a4: heading sample k of a cluster is the state (centre, 2 pi k / S)
// KIND 2 = a2: SOPLoader::getSOPFromWeightMatrix sop_loader.cpp:413-427
// all: a5 calcClusterDistancesBounds vns_sop_solver.cpp:262-297 in the epilogue
template <int KIND>
__global__ void __launch_bounds__(256) k_table_pairs(InstDev I, TableJob J)
{
   constexpr unsigned FULL = 0xffffffffu;
   const int lane = threadIdx.x & 31;
   const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
   const long long npairs = (long long)I.C * I.C;
   double hi_all = 0.0;
   int bad_all = 0;
   for (long long p = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; p < npairs; p += nwarps) {
      const int c1 = (int)(p / I.C), c2 = (int)(p - (long long)c1 * I.C);
      const int m1 = __ldg(I.m + c1), o1 = __ldg(I.off + c1), m2 = __ldg(I.m + c2), o2 = __ldg(I.off + c2);
      const long long base = (long long)o1 * I.N + (long long)m1 * o2;
      const int n = m1 * m2;
      double2 a, b;
      if (KIND == 1) { a = __ldg(J.pts + c1); b = __ldg(J.pts + c2); }
      // entry e = i * m2 + j; e advances by 32 per iteration: (i, j) follow without a division
      int i = lane / m2, j = lane - i * m2;
      const int di = 32 / m2, dj = 32 - di * m2;
      double lo = VNS_M, hi = -VNS_M;
      int bad = 0;
      for (int e = lane; e < n; e += 32) {
         double d;
         if (KIND == 0) {
            const double2 pa = __ldg(J.pts + o1 + i), pb = __ldg(J.pts + o2 + j);
            const double dx = __dsub_rn(pb.x, pa.x), dy = __dsub_rn(pb.y, pa.y);
            d = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
            if (J.rule == 1) d = round(d);
            else if (J.rule == 2) d = ceil(d);
         } else if (KIND == 1) {
            const double L = dubins_length(a.x, a.y, VNS_TWO_PI * i / J.S, b.x, b.y, VNS_TWO_PI * j / J.S, J.rho);
            d = J.do_ceil ? ceil(L) : L;
         } else {
            d = __ldg(J.W + ((long long)(o1 + i) * I.N + o2 + j));
         }
         J.D[base + e] = d;
         if (J.Di) J.Di[base + e] = (int)d;
         if (J.D16) J.D16[base + e] = (unsigned short)(d >= 0.0 && d < 65535.0 ? d : 65535.0);
         lo = d < lo ? d : lo;
         hi = d > hi ? d : hi;
         bad |= (d != floor(d)) || (d < 0);
         i += di; j += dj;
         if (j >= m2) { j -= m2; ++i; }
      }
#pragma unroll
      for (int s = 16; s > 0; s >>= 1) {
         double o = __shfl_xor_sync(FULL, lo, s);
         lo = o < lo ? o : lo;
         o = __shfl_xor_sync(FULL, hi, s);
         hi = o > hi ? o : hi;
      }
      bad_all |= bad;
      hi_all = hi > hi_all ? hi : hi_all;
      if (lane == 0) { J.dmin[p] = lo; J.dmax[p] = hi; }
   }
   // classification: one atomic per CTA
   __shared__ double s_hi[8];
   __shared__ int s_bad;
   if (threadIdx.x == 0) s_bad = 0;
   __syncthreads();
   if (__any_sync(FULL, bad_all) && lane == 0) s_bad = 1;
   if (lane == 0) s_hi[threadIdx.x >> 5] = hi_all;     // already the warp's maximum (hi was reduced per pair)
   __syncthreads();
   if (threadIdx.x == 0) {
      double h = 0.0;
      for (int w = 0; w < (int)(blockDim.x >> 5); ++w) h = s_hi[w] > h ? s_hi[w] : h;
      if (s_bad) atomicAdd(J.info, 1ull);
      else atomicMax(J.info + 1, (unsigned long long)__double_as_longlong(h));
   }
}
