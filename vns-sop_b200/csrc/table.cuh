// Distance-table kernels.  All write / read the block-major table with
// fully coalesced accesses: grid.y = source cluster c1 ("row band" of m[c1]*N contiguous entries),
// threads stride the band; the destination cluster of band position q is clof[q / m[c1]].
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

// a1/a3: SOPLoader::getSOPFromNodes sop_loader.cpp:457-491, GraphNode::distanceTo
// heuristic_types.h:67-75 (sqrt(dx*dx+dy*dy), no FMA contraction), getOP/getOPN :278,360 (raw)
__global__ void __launch_bounds__(256) k_table_euclid(InstDev I, const double2 *__restrict__ xy, int rule, double *__restrict__ D)
{
   const int c1 = blockIdx.y;
   const int m1 = I.m[c1], o1 = I.off[c1];
   const int band = m1 * I.N;
   double *out = D + (long long)o1 * I.N;
   for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < band; q += gridDim.x * blockDim.x) {
      BandPos p = band_decode(I, m1, q);
      double2 a = __ldg(xy + o1 + p.i), b = __ldg(xy + I.off[p.c2] + p.j);
      double dx = __dsub_rn(b.x, a.x), dy = __dsub_rn(b.y, a.y);
      double d = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
      if (rule == 1) d = round(d);       // EUC_2D  sop_loader.cpp:476
      else if (rule == 2) d = ceil(d);   // CEIL_2D sop_loader.cpp:480
      out[q] = d;
   }
}

#define VNS_TWO_PI 6.283185307179586476925286766559
// a CSC word whose straight segment has length 0 is feasible; rounding in sin/cos can push p^2 a few
// ulps below 0 there, so p^2 in [-eps, 0) is clamped to 0 instead of dropping the word
#define VNS_DUBINS_EPS 1e-9

__device__ __forceinline__ double mod2pi(double t) { return t - VNS_TWO_PI * floor(t / VNS_TWO_PI); }

// Shortest Dubins path (six words, normalised by rho).  No reference code exists;
// the known answers are the EXPLICIT matrices of datasets/dop_sop_dataset.
// Operation order mirrors oracle/vns_oracle.c:vo_dubins_length term by term.
__device__ double dubins_length(double x0, double y0, double th0, double x1, double y1, double th1, double rho)
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

// a4: heading sample k of a cluster is the state (centre, 2 pi k / S)
__global__ void __launch_bounds__(256) k_table_dubins(InstDev I, const double2 *__restrict__ ctr, int S, double rho, int do_ceil,
                                                       double *__restrict__ D)
{
   const int c1 = blockIdx.y;
   const int m1 = I.m[c1], o1 = I.off[c1];
   const int band = m1 * I.N;
   double *out = D + (long long)o1 * I.N;
   const double2 a = ctr[c1];
   for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < band; q += gridDim.x * blockDim.x) {
      BandPos p = band_decode(I, m1, q);
      double2 b = __ldg(ctr + p.c2);
      double L = dubins_length(a.x, a.y, VNS_TWO_PI * p.i / S, b.x, b.y, VNS_TWO_PI * p.j / S, rho);
      out[q] = do_ceil ? ceil(L) : L;
   }
}

// a2: SOPLoader::getSOPFromWeightMatrix sop_loader.cpp:413-427 (W in sample order); to_dense = 1
// runs the inverse permutation (download for tests)
__global__ void __launch_bounds__(256) k_table_gather(InstDev I, double *__restrict__ D, double *__restrict__ W, int to_dense)
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

// a5: calcClusterDistancesBounds vns_sop_solver.cpp:262-297, one warp per ordered cluster pair.
// Also classifies the table: info[0] = count of non-integer or negative entries,
// info[1] = bit pattern of the largest entry (entries >= 0 order like their bit patterns).
__global__ void __launch_bounds__(256) k_bounds(InstDev I, double *__restrict__ dmin, double *__restrict__ dmax, unsigned long long *info)
{
   const int lane = threadIdx.x & 31;
   const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
   const long long npairs = (long long)I.C * I.C;
   if (warp >= npairs) return;
   const int c1 = (int)(warp / I.C), c2 = (int)(warp - (long long)c1 * I.C);
   const double *blk = blk_ptr(I, c1, c2);
   const int n = I.m[c1] * I.m[c2];
   double lo = VNS_M, hi = -VNS_M;
   int bad = 0;
   for (int k = lane; k < n; k += 32) {
      double v = __ldg(blk + k);
      lo = v < lo ? v : lo;
      hi = v > hi ? v : hi;
      bad |= (v != floor(v)) || (v < 0);
   }
#pragma unroll
   for (int s = 16; s > 0; s >>= 1) {
      double o = __shfl_xor_sync(0xffffffffu, lo, s);
      lo = o < lo ? o : lo;
      o = __shfl_xor_sync(0xffffffffu, hi, s);
      hi = o > hi ? o : hi;
      bad |= __shfl_xor_sync(0xffffffffu, bad, s);
   }
   if (lane == 0) {
      dmin[warp] = lo;
      dmax[warp] = hi;
      if (bad) atomicAdd(info, 1ull);
      else atomicMax(info + 1, (unsigned long long)__double_as_longlong(hi));
   }
}
