// Jump-ahead for glibc rand() (random_r TYPE_3: x_i = x_{i-31} + x_{i-3} mod 2^32; math_common.cpp:21-28 draws from it).
// The 31-word state advances linearly: s' = M s with the companion matrix M (s'[t] = s[t+1], s'[30] = s[0] + s[28]), so
// the state d draws ahead is M^d s - one 31 x 31 matrix-vector product mod 2^32 per precomputed power.  Host only.
#pragma once
#include <cstdint>
#include <vector>

#include "common.cuh"

namespace rngjump {
typedef std::vector<unsigned> Mat;   // 31 x 31, row-major: out[t] = sum_j A[t][j] * s[j]

inline Mat mul(const Mat &A, const Mat &B)
{
   Mat Cm(31 * 31, 0u);
   for (int i = 0; i < 31; ++i)
      for (int k = 0; k < 31; ++k) {
         const unsigned a = A[i * 31 + k];
         if (!a) continue;
         for (int j = 0; j < 31; ++j) Cm[i * 31 + j] += a * B[k * 31 + j];
      }
   return Cm;
}
inline Mat identity()
{
   Mat P(31 * 31, 0u);
   for (int t = 0; t < 31; ++t) P[t * 31 + t] = 1u;
   return P;
}
inline Mat companion()
{
   Mat M(31 * 31, 0u);
   for (int t = 0; t < 30; ++t) M[t * 31 + t + 1] = 1u;
   M[30 * 31 + 0] = 1u; M[30 * 31 + 28] = 1u;
   return M;
}
inline Mat pow2(int log2_draws)        // M^(2^log2_draws)
{
   Mat S = companion();
   for (int d = 0; d < log2_draws; ++d) S = mul(S, S);
   return S;
}
// out[r][j][t] (t padded to 32) = coefficient of s[j] in state word t after r * 2^log2_draws draws, r = 0 .. count-1:
// the layout the device reads (one coalesced row per j)
inline void powers_for_device(int count, int log2_draws, std::vector<unsigned> &out)
{
   const Mat step = pow2(log2_draws);
   Mat P = identity();
   out.assign((size_t)count * 31 * 32, 0u);
   for (int r = 0; r < count; ++r) {
      for (int t = 0; t < 31; ++t)
         for (int j = 0; j < 31; ++j) out[((size_t)r * 31 + j) * 32 + t] = P[t * 31 + j];
      P = mul(step, P);
   }
}
// canonical form: r[0] oldest (x_{i-31}), f = 0, b = 28
inline void canonical(const RngState &g, unsigned s[31])
{
   for (int k = 0; k < 31; ++k) s[k] = (unsigned)g.r[(g.f + k) % 31];
}
// the state after `ndraws` further rand() calls
inline void advance(RngState &g, uint64_t ndraws)
{
   static std::vector<Mat> p2;         // M^(2^k)
   if (p2.empty()) {
      p2.push_back(companion());
      for (int k = 1; k < 48; ++k) p2.push_back(mul(p2.back(), p2.back()));
   }
   unsigned s[31], o[31];
   canonical(g, s);
   for (int k = 0; k < 48; ++k)
      if ((ndraws >> k) & 1) {
         const Mat &A = p2[k];
         for (int t = 0; t < 31; ++t) {
            unsigned acc = 0;
            for (int j = 0; j < 31; ++j) acc += A[t * 31 + j] * s[j];
            o[t] = acc;
         }
         for (int t = 0; t < 31; ++t) s[t] = o[t];
      }
   for (int k = 0; k < 31; ++k) g.r[k] = (int32_t)s[k];
   g.f = 0; g.b = 28; g.pad = 0;
}
}   // namespace rngjump
