// Device-side candidate generation (north star item 3): the proposals of insertRandom
// (vns_sop_solver.cpp:507-519: from in [0, |VNS|-1], to in [0, T], both redrawn while |from - to| <= 1) and exchangeRandom
// (:723-730: to in [0, T-1], redrawn while |from - to| <= 2) in DRAW ORDER, as device descriptors - the batched evaluation
// then needs no candidate bytes over PCIe at all.
//
// glibc's rand() stream is linear, so slice r of 256 draw pairs starts from the state M^(512 r) s: two 31 x 31
// matrix-vector products (powers of M^512 and of M^(512 * 64), rngjump.hpp).  Pass 1: every thread block generates its
// slice (warp 0: 17 x 31 values by shuffle prefix sums), maps the pairs to (from, to), counts the valid ones.  Pass 2:
// exclusive scan of the slice counts.  Pass 3: valid pair k of slice r becomes proposal offset[r] + k, written as a
// vnsgpu_cand (16 B) or a vnsgpu_pack_cand32 word together with the number of rand() calls consumed up to and including
// it - what rewinds the RNG to "just after the accepted proposal" (rngjump::advance on the host).
#include "ctx.hpp"
#include "rngjump.hpp"

#define GEN_PAIRS 256
#define GEN_LO 64
#define GEN_HI 1024

struct GenJob {
   const unsigned *lo, *hi;     // jump matrices [r][31][32]
   unsigned s0[31];             // canonical state at the start of the call
   int kind, nvns, T;
   long long nslices;
};

static __global__ void __launch_bounds__(GEN_PAIRS) k_gen_pairs(GenJob J, unsigned long long *__restrict__ pairs, unsigned *__restrict__ counts)
{
   __shared__ unsigned ring[31 + 31 * 17 + 2];
   __shared__ unsigned s1[32];
   __shared__ int wcnt[GEN_PAIRS / 32];
   constexpr unsigned FULL = 0xffffffffu;
   const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
   const long long r = blockIdx.x;
   if (warp == 0) {
      const unsigned *mh = J.hi + (size_t)(r / GEN_LO) * 31 * 32, *ml = J.lo + (size_t)(r % GEN_LO) * 31 * 32;
      unsigned acc = 0;
      if (lane < 31)
         for (int j = 0; j < 31; ++j) acc += __ldg(mh + j * 32 + lane) * J.s0[j];
      s1[lane] = acc;
      __syncwarp();
      acc = 0;
      if (lane < 31)
         for (int j = 0; j < 31; ++j) acc += __ldg(ml + j * 32 + lane) * s1[j];
      if (lane < 31) ring[lane] = acc;
      __syncwarp();
      for (int gen = 31; gen < 31 + 2 * GEN_PAIRS; gen += 31) {     // x_{gen+k} = x_{gen+k-31} + x_{gen+k-3}: scan with stride 3
         unsigned a = 0;
         if (lane < 31) a = ring[gen - 31 + lane];
         if (lane < 3) a += ring[gen - 3 + lane];
#pragma unroll
         for (int d = 3; d < 32; d <<= 1) {
            const unsigned t = __shfl_up_sync(FULL, a, d);
            if (lane >= d) a += t;
         }
         if (lane < 31) ring[gen + lane] = a;
         __syncwarp();
      }
   }
   __syncthreads();
   const int p = threadIdx.x;
   const unsigned r1 = ring[31 + 2 * p] >> 1, r2 = ring[31 + 2 * p + 1] >> 1;
   const int tohi = J.kind == 1 ? J.T : J.T - 1, gap = J.kind == 1 ? 1 : 2;
   const int from = (int)(((unsigned long long)r1 * (unsigned)J.nvns) >> 31);           // randIntMinMax, math_common.cpp:21-28
   const int to = (int)(((unsigned long long)r2 * (unsigned)(tohi + 1)) >> 31);
   const bool valid = abs(from - to) > gap;
   const unsigned bal = __ballot_sync(FULL, valid);
   if (lane == 0) wcnt[warp] = __popc(bal);
   __syncthreads();
   int rank = __popc(bal & ((1u << lane) - 1)), total = 0;
   for (int w = 0; w < GEN_PAIRS / 32; ++w) { if (w < warp) rank += wcnt[w]; total += wcnt[w]; }
   pairs[r * GEN_PAIRS + p] = valid ? (1ull << 63) | ((unsigned long long)rank << 40) | ((unsigned long long)to << 20) | (unsigned)from : 0ull;
   if (threadIdx.x == 0) counts[r] = (unsigned)total;
}

// exclusive scan of the slice counts (one block); total[0] = number of valid pairs, total[1] = draws consumed up to and
// including proposal n - 1 (0 when there are fewer than n)
static __global__ void __launch_bounds__(1024) k_gen_scan(long long nslices, const unsigned *__restrict__ counts, unsigned long long *__restrict__ offs,
                                                          unsigned long long *total)
{
   __shared__ unsigned long long part[1024];
   const long long chunk = (nslices + 1023) / 1024, lo = chunk * threadIdx.x, hi = min(nslices, lo + chunk);
   unsigned long long sum = 0;
   for (long long k = lo; k < hi; ++k) sum += counts[k];
   part[threadIdx.x] = sum;
   __syncthreads();
   for (int d = 1; d < 1024; d <<= 1) {
      const unsigned long long add = threadIdx.x >= d ? part[threadIdx.x - d] : 0;
      __syncthreads();
      part[threadIdx.x] += add;
      __syncthreads();
   }
   unsigned long long pos = part[threadIdx.x] - sum;
   for (long long k = lo; k < hi; ++k) { offs[k] = pos; pos += counts[k]; }
   if (threadIdx.x == 1023) total[0] = part[1023];
}

static __global__ void __launch_bounds__(GEN_PAIRS) k_gen_emit(int kind, int T, long long n, const int *__restrict__ vns, const unsigned long long *__restrict__ pairs,
                                                               const unsigned long long *__restrict__ offs, void *__restrict__ out, int format,
                                                               unsigned *__restrict__ draws, unsigned long long *total)
{
   const long long r = blockIdx.x, gp = r * GEN_PAIRS + threadIdx.x;
   const unsigned long long w = pairs[gp];
   if (!(w >> 63)) return;
   const long long idx = (long long)offs[r] + (long long)((w >> 40) & 0x1ff);
   if (idx >= n) return;
   const int from = (int)(w & 0xfffff), to = (int)((w >> 20) & 0xfffff);
   const bool in = from < T;
   const int c = in ? -1 : __ldg(vns + from);
   const int op = kind == 1 ? (in ? VNSGPU_OP_MOVE : VNSGPU_OP_ADD) : (in ? VNSGPU_OP_EXCHANGE : VNSGPU_OP_REPLACE);
   const int a = in ? from : c;
   if (format == 2)
      reinterpret_cast<unsigned *>(out)[idx] = (unsigned)(op & 7) | ((unsigned)(a & 0x3fff) << 3) | ((unsigned)(to & 0x3fff) << 17) | ((unsigned)(!in) << 31);
   else
      reinterpret_cast<int4 *>(out)[idx] = make_int4(op, a, to, c);
   const unsigned long long used = 2ull * (unsigned long long)(gp + 1);      // rand() calls up to and including this pair
   if (draws) draws[idx] = (unsigned)used;
   if (idx == n - 1) total[1] = used;
}

static vnsgpu_status gen_tables(vnsgpu_ctx *ctx)
{
   if (ctx->d_genjump) return VNSGPU_OK;
   std::vector<unsigned> lo, hi;
   rngjump::powers_for_device(GEN_LO, 9, lo);            // steps of 2 * GEN_PAIRS = 2^9 draws
   rngjump::powers_for_device(GEN_HI, 15, hi);           // steps of 2^9 * GEN_LO = 2^15 draws
   CK(cudaMalloc(&ctx->d_genjump, sizeof(unsigned) * (lo.size() + hi.size())));
   CK(cudaMemcpy(ctx->d_genjump, lo.data(), sizeof(unsigned) * lo.size(), cudaMemcpyHostToDevice));
   CK(cudaMemcpy(ctx->d_genjump + lo.size(), hi.data(), sizeof(unsigned) * hi.size(), cudaMemcpyHostToDevice));
   ctx->genjump_lo_words = lo.size();
   return VNSGPU_OK;
}

extern "C" {

void vnsgpu_rng_advance(vnsgpu_rng *rng, uint64_t ndraws) { rngjump::advance(*reinterpret_cast<RngState *>(rng), ndraws); }

vnsgpu_status vnsgpu_gen_proposals_dev(vnsgpu_ctx *ctx, int path, int kind, int64_t n, vnsgpu_rng *rng, void *d_out, int format,
                                       uint32_t *d_draws, void *stream)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   if (kind != 1 && kind != 2) return fail(ctx, VNSGPU_E_RANGE, "kind must be 1 (insert) or 2 (exchange)");
   if (format != 0 && format != 2) return fail(ctx, VNSGPU_E_RANGE, "format must be 0 (vnsgpu_cand) or 2 (vnsgpu_pack_cand32 words)");
   if (n <= 0) return VNSGPU_OK;
   cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
   PathHdr h;
   vnsgpu_status s = read_hdr(ctx, path, &h);
   if (s) return s;
   const int T = h.hdr[HDR_T], nvns = h.hdr[HDR_NVNS];
   if (kind == 1 ? nvns < 3 : (nvns < 4 || T < 3)) return fail(ctx, VNSGPU_E_STATE, "neighbourhood is empty for this state");
   if (nvns >= (1 << 20)) return fail(ctx, VNSGPU_E_RANGE, "more than 2^20 clusters");
   if (format == 2 && ctx->I.C >= 16384) return fail(ctx, VNSGPU_E_RANGE, "4-byte descriptors need indices below 16384");
   // exact probability that a draw pair is a proposal -> slices needed for n proposals, with 6 sigma + one slice of slack
   double pv = 0;
   {
      const int tohi = kind == 1 ? T : T - 1, gap = kind == 1 ? 1 : 2;
      long long ok = 0;
      for (int to = 0; to <= tohi; ++to) ok += nvns - (std::min(nvns - 1, to + gap) - std::max(0, to - gap) + 1);
      pv = (double)ok / ((double)nvns * (tohi + 1));
   }
   if (pv < 0.05) return fail(ctx, VNSGPU_E_STATE, "fewer than 5% of the draw pairs are proposals in this state: use vnsgpu_gen_proposals");
   const double need = (double)n / pv;
   const long long nslices = (long long)((need + 6.0 * std::sqrt(need * (1.0 - pv) / pv) + 2 * GEN_PAIRS) / GEN_PAIRS) + 1;
   if (nslices > (long long)GEN_LO * GEN_HI) return fail(ctx, VNSGPU_E_RANGE, "more than 16 Mi draw pairs in one call: split the batch");
   s = gen_tables(ctx);
   if (s) return s;
   const size_t bytes = sizeof(unsigned long long) * (size_t)nslices * GEN_PAIRS + sizeof(unsigned long long) * (size_t)nslices + sizeof(unsigned) * (size_t)nslices + 64;
   if (bytes > ctx->gen_bytes) {
      if (ctx->d_gen) { CK(cudaStreamSynchronize(st)); CK(cudaFree(ctx->d_gen)); }
      ctx->d_gen = nullptr; ctx->gen_bytes = 0;
      CK(cudaMalloc(&ctx->d_gen, bytes));
      ctx->gen_bytes = bytes;
   }
   unsigned long long *d_total = (unsigned long long *)ctx->d_gen, *d_pairs = d_total + 8, *d_offs = d_pairs + (size_t)nslices * GEN_PAIRS;
   unsigned *d_counts = (unsigned *)(d_offs + nslices);
   GenJob J;
   J.lo = ctx->d_genjump; J.hi = ctx->d_genjump + ctx->genjump_lo_words;
   rngjump::canonical(*reinterpret_cast<const RngState *>(rng), J.s0);
   J.kind = kind; J.nvns = nvns; J.T = T; J.nslices = nslices;
   CK(cudaMemsetAsync(d_total, 0, 64, st));
   k_gen_pairs<<<(unsigned)nslices, GEN_PAIRS, 0, st>>>(J, d_pairs, d_counts);
   LAUNCHED(ctx);
   k_gen_scan<<<1, 1024, 0, st>>>(nslices, d_counts, d_offs, d_total);
   LAUNCHED(ctx);
   k_gen_emit<<<(unsigned)nslices, GEN_PAIRS, 0, st>>>(kind, T, n, ctx->paths[path].d.vns, d_pairs, d_offs, d_out, format, d_draws, d_total);
   LAUNCHED(ctx);
   unsigned long long tot[2];
   CK(cudaMemcpyAsync(tot, d_total, sizeof(tot), cudaMemcpyDeviceToHost, st));
   CK(cudaStreamSynchronize(st));
   if ((int64_t)tot[0] < n || tot[1] == 0) return fail(ctx, VNSGPU_E_STATE, "vnsgpu_gen_proposals_dev: slice estimate fell short (6-sigma event): call again");
   rngjump::advance(*reinterpret_cast<RngState *>(rng), tot[1]);
   return VNSGPU_OK;
}

// One round of a host-driven search: n proposals of neighbourhood `kind` drawn on the device, evaluated against the
// path's state, first accepted one in draw order returned (index, descriptor) with *rng rewound to just after ITS draws
// (after all n proposals when none is accepted) - exactly what the reference's loop leaves behind (:507-646, :723-832).
// Moves 132 B in and ~30 B out, whatever n.
vnsgpu_status vnsgpu_propose_eval(vnsgpu_ctx *ctx, int path, int kind, int64_t n, vnsgpu_rng *rng, double budget, int skip,
                                  int64_t *first_accept, vnsgpu_cand *accepted)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   if (first_accept) *first_accept = -1;
   if (n <= 0) return VNSGPU_OK;
   const size_t bytes = sizeof(vnsgpu_cand) * (size_t)n + sizeof(unsigned) * (size_t)n + 64;
   if (bytes > ctx->gencands_bytes) {
      if (ctx->d_gencands) { CK(cudaStreamSynchronize(ctx->stream)); CK(cudaFree(ctx->d_gencands)); }
      ctx->d_gencands = nullptr; ctx->gencands_bytes = 0;
      CK(cudaMalloc(&ctx->d_gencands, bytes));
      ctx->gencands_bytes = bytes;
   }
   long long *d_first = (long long *)ctx->d_gencands;
   vnsgpu_cand *d_c = (vnsgpu_cand *)((char *)ctx->d_gencands + 64);
   unsigned *d_draws = (unsigned *)(d_c + n);
   const vnsgpu_rng rng0 = *rng;
   vnsgpu_status s = vnsgpu_gen_proposals_dev(ctx, path, kind, n, rng, d_c, 0, d_draws, ctx->stream);
   if (s) return s;
   s = vnsgpu_eval_batch_dev(ctx, path, n, d_c, budget, skip, nullptr, nullptr, nullptr, d_first, ctx->stream);
   if (s) return s;
   long long first = -1;
   CK(cudaMemcpyAsync(&first, d_first, sizeof(first), cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   if (first >= 0) {
      unsigned used = 0;
      vnsgpu_cand c;
      CK(cudaMemcpyAsync(&used, d_draws + first, sizeof(used), cudaMemcpyDeviceToHost, ctx->stream));
      CK(cudaMemcpyAsync(&c, d_c + first, sizeof(c), cudaMemcpyDeviceToHost, ctx->stream));
      CK(cudaStreamSynchronize(ctx->stream));
      *rng = rng0;
      vnsgpu_rng_advance(rng, used);
      if (accepted) *accepted = c;
   }
   if (first_accept) *first_accept = first;
   return VNSGPU_OK;
}

} // extern "C"
