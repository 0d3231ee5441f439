// Device-side candidate generation (north star item 3): the proposals of insertRandom
// (vns_sop_solver.cpp:507-519: from in [0, |VNS|-1], to in [0, T], both redrawn while |from - to| <= 1) and exchangeRandom
// (:723-730: to in [0, T-1], redrawn while |from - to| <= 2) in DRAW ORDER, as device descriptors - the batched evaluation
// then needs no candidate bytes over PCIe at all.
//
// glibc's rand() stream is linear, so warp w's 32 draw pairs start from the state M^(64 w) s: three 31 x 31 matrix-vector
// products (powers of M^64, M^4096 and M^262144, rngjump.hpp).  Pass 1: every warp generates its 64 draws (3 x 31 values
// by shuffle prefix sums), maps the pairs to (from, to); a thread block counts its 256 pairs' valid ones.  Pass 2:
// exclusive scan of the slice counts.  Pass 3: valid pair k of slice r becomes proposal offset[r] + k, written as a
// vnsgpu_cand (16 B) or a vnsgpu_pack_cand32 word together with the number of rand() calls consumed up to and including
// it - what rewinds the RNG to "just after the accepted proposal" (rngjump::advance on the host).
#include "ctx.hpp"
#include "rngjump.hpp"

#define GEN_PAIRS 256          // pairs per thread block (one slice of the scan)
#define GEN_WPAIRS 32          // pairs per warp: every warp jumps to its own 64 draws
#define GEN_L0 64              // jump tables: M^(64 a), a < 64;  M^(4096 b), b < 64;  M^(262144 c), c < 256  (64 Mi draws)
#define GEN_L1 64
#define GEN_L2 256

struct GenJob {
   const unsigned *j0, *j1, *j2;   // jump matrices [r][31][32]
   unsigned s0[31];                // canonical state at the start of the call
   int kind, nvns, T;
   long long nslices;
};

static __device__ __forceinline__ unsigned gen_matvec(const unsigned *__restrict__ m, const unsigned *v, int lane)
{
   unsigned acc = 0;
   if (lane < 31)
#pragma unroll 8
      for (int j = 0; j < 31; ++j) acc += __ldg(m + j * 32 + lane) * v[j];
   return acc;
}

static __global__ void __launch_bounds__(GEN_PAIRS) k_gen_pairs(GenJob J, unsigned long long *__restrict__ pairs, unsigned *__restrict__ counts)
{
   __shared__ unsigned ring[GEN_PAIRS / 32][31 + 31 * 3 + 3];
   __shared__ unsigned sv[GEN_PAIRS / 32][2][32];
   __shared__ int wcnt[GEN_PAIRS / 32];
   constexpr unsigned FULL = 0xffffffffu;
   const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
   const long long ws = (long long)blockIdx.x * (GEN_PAIRS / 32) + warp;       // this warp's slice of 64 draws
   unsigned *rg = ring[warp];
   {
      const unsigned a = gen_matvec(J.j2 + (size_t)(ws / (GEN_L0 * GEN_L1)) * 31 * 32, J.s0, lane);
      sv[warp][0][lane] = a;
      __syncwarp();
      const unsigned b = gen_matvec(J.j1 + (size_t)((ws / GEN_L0) % GEN_L1) * 31 * 32, sv[warp][0], lane);
      sv[warp][1][lane] = b;
      __syncwarp();
      const unsigned c = gen_matvec(J.j0 + (size_t)(ws % GEN_L0) * 31 * 32, sv[warp][1], lane);
      if (lane < 31) rg[lane] = c;
      __syncwarp();
      for (int gen = 31; gen < 31 + 2 * GEN_WPAIRS; gen += 31) {     // x_{gen+k} = x_{gen+k-31} + x_{gen+k-3}: scan with stride 3
         unsigned a2 = 0;
         if (lane < 31) a2 = rg[gen - 31 + lane];
         if (lane < 3) a2 += rg[gen - 3 + lane];
#pragma unroll
         for (int d = 3; d < 32; d <<= 1) {
            const unsigned t = __shfl_up_sync(FULL, a2, d);
            if (lane >= d) a2 += t;
         }
         if (lane < 31) rg[gen + lane] = a2;
         __syncwarp();
      }
   }
   const int p = threadIdx.x;
   const unsigned r1 = rg[31 + 2 * lane] >> 1, r2 = rg[31 + 2 * lane + 1] >> 1;
   const int tohi = J.kind == 1 ? J.T : J.T - 1, gap = J.kind == 1 ? 1 : 2;
   const int from = (int)(((unsigned long long)r1 * (unsigned)J.nvns) >> 31);           // randIntMinMax, math_common.cpp:21-28
   const int to = (int)(((unsigned long long)r2 * (unsigned)(tohi + 1)) >> 31);
   const bool valid = abs(from - to) > gap;
   const unsigned bal = __ballot_sync(FULL, valid);
   if (lane == 0) wcnt[warp] = __popc(bal);
   __syncthreads();
   int rank = __popc(bal & ((1u << lane) - 1)), total = 0;
   for (int w = 0; w < GEN_PAIRS / 32; ++w) { if (w < warp) rank += wcnt[w]; total += wcnt[w]; }
   pairs[(long long)blockIdx.x * GEN_PAIRS + p] = valid ? (1ull << 63) | ((unsigned long long)rank << 40) | ((unsigned long long)to << 20) | (unsigned)from : 0ull;
   if (threadIdx.x == 0) counts[blockIdx.x] = (unsigned)total;
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
   std::vector<unsigned> t0, t1, t2;
   rngjump::powers_for_device(GEN_L0, 6, t0);            // steps of 2 * GEN_WPAIRS = 2^6 draws
   rngjump::powers_for_device(GEN_L1, 12, t1);           // steps of 2^6 * GEN_L0 = 2^12 draws
   rngjump::powers_for_device(GEN_L2, 18, t2);           // steps of 2^12 * GEN_L1 = 2^18 draws
   CK(cudaMalloc(&ctx->d_genjump, sizeof(unsigned) * (t0.size() + t1.size() + t2.size())));
   CK(cudaMemcpy(ctx->d_genjump, t0.data(), sizeof(unsigned) * t0.size(), cudaMemcpyHostToDevice));
   CK(cudaMemcpy(ctx->d_genjump + t0.size(), t1.data(), sizeof(unsigned) * t1.size(), cudaMemcpyHostToDevice));
   CK(cudaMemcpy(ctx->d_genjump + t0.size() + t1.size(), t2.data(), sizeof(unsigned) * t2.size(), cudaMemcpyHostToDevice));
   ctx->genjump_lo_words = t0.size();
   ctx->genjump_mid_words = t1.size();
   return VNSGPU_OK;
}

// what one round of a host-driven search reads back: one 64-byte record, one copy, one synchronisation
struct GenPick {
   long long first;               // first accepted proposal or -1
   unsigned long long total_valid, total_draws;
   unsigned draws_first, pad;
   int4 cand;
};

static __global__ void k_gen_pick(const long long *d_first, const unsigned long long *total, const unsigned *draws, const void *cands, int format,
                                  GenPick *out)
{
   GenPick g;
   g.first = *d_first;
   g.total_valid = total[0]; g.total_draws = total[1];
   g.draws_first = 0; g.pad = 0; g.cand = make_int4(-1, 0, 0, -1);
   if (g.first >= 0) {
      g.draws_first = draws[g.first];
      if (format == 2) {
         const unsigned w = reinterpret_cast<const unsigned *>(cands)[g.first];
         g.cand.x = (int)(w & 7); g.cand.y = (int)((w >> 3) & 0x3fff); g.cand.z = (int)((w >> 17) & 0x3fff); g.cand.w = (w >> 31) ? g.cand.y : -1;
      } else
         g.cand = reinterpret_cast<const int4 *>(cands)[g.first];
   }
   *out = g;
}

// launches the three generation kernels on `st`; no synchronisation.  d_total: 2 x u64 (valid pairs, draws up to proposal n - 1)
static vnsgpu_status gen_launch(vnsgpu_ctx *ctx, int path, int kind, int64_t n, const vnsgpu_rng *rng, void *d_out, int format, uint32_t *d_draws,
                                cudaStream_t st, unsigned long long **d_total_out)
{
   PathHdr h;
   vnsgpu_status s = VNSGPU_OK;
   if (ctx->paths[path].hdr_ok) h = ctx->paths[path].hdr_cache;      // no round trip: every mutating call read the header back
   else s = read_hdr(ctx, path, &h);
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
   if (nslices * GEN_PAIRS * 2 > (long long)64 * GEN_L0 * GEN_L1 * GEN_L2) return fail(ctx, VNSGPU_E_RANGE, "more than 32 Mi draw pairs in one call: split the batch");
   s = gen_tables(ctx);
   if (s) return s;
   const size_t bytes = sizeof(unsigned long long) * (size_t)nslices * GEN_PAIRS + sizeof(unsigned long long) * (size_t)nslices + sizeof(unsigned) * (size_t)nslices + 256;
   if (bytes > ctx->gen_bytes) {
      if (ctx->d_gen) { CK(cudaStreamSynchronize(st)); CK(cudaFree(ctx->d_gen)); }
      ctx->d_gen = nullptr; ctx->gen_bytes = 0;
      CK(cudaMalloc(&ctx->d_gen, bytes));
      ctx->gen_bytes = bytes;
   }
   unsigned long long *d_total = (unsigned long long *)ctx->d_gen, *d_pairs = d_total + 32, *d_offs = d_pairs + (size_t)nslices * GEN_PAIRS;
   unsigned *d_counts = (unsigned *)(d_offs + nslices);
   GenJob J;
   J.j0 = ctx->d_genjump; J.j1 = J.j0 + ctx->genjump_lo_words; J.j2 = J.j1 + ctx->genjump_mid_words;
   rngjump::canonical(*reinterpret_cast<const RngState *>(rng), J.s0);
   J.kind = kind; J.nvns = nvns; J.T = T; J.nslices = nslices;
   CK(cudaMemsetAsync(d_total, 0, 64, st));
   k_gen_pairs<<<(unsigned)nslices, GEN_PAIRS, 0, st>>>(J, d_pairs, d_counts);
   LAUNCHED(ctx);
   k_gen_scan<<<1, 1024, 0, st>>>(nslices, d_counts, d_offs, d_total);
   LAUNCHED(ctx);
   k_gen_emit<<<(unsigned)nslices, GEN_PAIRS, 0, st>>>(kind, T, n, ctx->paths[path].d.vns, d_pairs, d_offs, d_out, format, d_draws, d_total);
   LAUNCHED(ctx);
   *d_total_out = d_total;
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
   unsigned long long *d_total = nullptr;
   vnsgpu_status s = gen_launch(ctx, path, kind, n, rng, d_out, format, d_draws, st, &d_total);
   if (s) return s;
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
   if (kind != 1 && kind != 2) return fail(ctx, VNSGPU_E_RANGE, "kind must be 1 (insert) or 2 (exchange)");
   if (first_accept) *first_accept = -1;
   if (n <= 0) return VNSGPU_OK;
   const size_t bytes = sizeof(vnsgpu_cand) * (size_t)n + sizeof(unsigned) * (size_t)n + 256;
   if (bytes > ctx->gencands_bytes) {
      if (ctx->d_gencands) { CK(cudaStreamSynchronize(ctx->stream)); CK(cudaFree(ctx->d_gencands)); }
      ctx->d_gencands = nullptr; ctx->gencands_bytes = 0;
      CK(cudaMalloc(&ctx->d_gencands, bytes));
      ctx->gencands_bytes = bytes;
   }
   if (!ctx->h_pick) CK(cudaMallocHost(&ctx->h_pick, sizeof(GenPick)));
   long long *d_first = (long long *)ctx->d_gencands;
   GenPick *d_pick = (GenPick *)((char *)ctx->d_gencands + 64);
   void *d_c = (char *)ctx->d_gencands + 256;
   unsigned *d_draws = (unsigned *)((char *)d_c + sizeof(vnsgpu_cand) * (size_t)n);
   cudaStream_t st = ctx->stream;
   // 4-byte descriptors that the endpoint-vector kernel reads in place, when the batch takes that form
   const int format = eval_dev_accepts_packed32(ctx, n) ? 2 : 0;
   unsigned long long *d_total = nullptr;
   vnsgpu_status s = gen_launch(ctx, path, kind, n, rng, d_c, format, d_draws, st, &d_total);
   if (s) return s;
   s = eval_dev_first(ctx, path, n, d_c, format, budget, skip, d_first, st);
   if (s) return s;
   k_gen_pick<<<1, 1, 0, st>>>(d_first, d_total, d_draws, d_c, format, d_pick);
   LAUNCHED(ctx);
   GenPick *hp = (GenPick *)ctx->h_pick;
   CK(cudaMemcpyAsync(hp, d_pick, sizeof(GenPick), cudaMemcpyDeviceToHost, st));
   CK(cudaStreamSynchronize(st));
   if ((int64_t)hp->total_valid < n || hp->total_draws == 0) return fail(ctx, VNSGPU_E_STATE, "vnsgpu_propose_eval: slice estimate fell short (6-sigma event): call again");
   if (hp->first >= 0) {
      vnsgpu_rng_advance(rng, hp->draws_first);
      if (accepted) { accepted->op = hp->cand.x; accepted->a = hp->cand.y; accepted->b = hp->cand.z; accepted->gain_cluster = hp->cand.w; }
   } else
      vnsgpu_rng_advance(rng, hp->total_draws);
   if (first_accept) *first_accept = hp->first;
   return VNSGPU_OK;
}

} // extern "C"
