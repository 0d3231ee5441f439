// Host side of the second search chain (chain2.cuh): fit test, per-context device tables (int32 cluster-pair minima,
// cluster offsets, RNG jump matrices, grid barrier + slots), cooperative launches.  Called from api_search.cu.
#include "ctx.hpp"
#include "path.cuh"
#include "search.cuh"
#include "eval.cuh"
#include "chain2.cuh"
#include "chain2_host.hpp"
#include "rngjump.hpp"

static __global__ void __launch_bounds__(256) k_dmin_to_int(int n, const double *__restrict__ dmin, int *__restrict__ out)
{
   const int k = blockIdx.x * blockDim.x + threadIdx.x;
   if (k < n) out[k] = (int)dmin[k];
}

// jump[r][j][t]: coefficient of s[j] in state word t after 2 * C2_PAIRS * r draws (rngjump.hpp)
static void build_jump(int ncta, std::vector<unsigned> &out)
{
   int lg = 0;
   while ((1 << lg) < 2 * C2_PAIRS) ++lg;
   rngjump::powers_for_device(ncta, lg, out);
}

// placement: bit 0 = cluster table in shared memory, bit 1 = F / Brev rows in shared memory (else per-block global slices)
bool c2_fits(const vnsgpu_ctx *ctx, size_t *smem, int *place)
{
   // automatic choice: below ~96 clusters a round is a handful of short chains and the first form's leaner round wins
   // (seeded krob200 solve 4.45 s against 4.89 s); from there on the lower-bound filter, the large window and the product
   // tree do (vm1084: 458 s -> 35.6 s)
   const bool want = ctx->chain_form == 0 || ((ctx->chain_form < 0 || ctx->chain_form == 3) && ctx->I.C >= 96);
   if (!want || !vnsgpu_int_path_active(ctx) || ctx->I.mp > 32 || !group_kernel_ok(ctx)) return false;
   // static shared memory of the kernels (control block, slot, RNG, the few words of search.cuh's helpers) + margin
   const size_t fixed = sizeof(C2Ctl) + sizeof(C2Slot) + sizeof(RngState) + sizeof(C2Tree) + 2048;
   const size_t greedy = sizeof(double) * 2 * ctx->I.mp * VNS_CHAIN_WARPS;   // the greedy start of vns_run borrows the buffer
   const int forced = getenv("VNSGPU_CHAIN_PLACE") ? atoi(getenv("VNSGPU_CHAIN_PLACE")) : -1;   // tests: force a placement
   for (int pl = 3; pl >= 0; --pl) {
      if (forced >= 0 && pl != (forced & 3)) continue;
      const size_t b = std::max(c2_smem_bytes(ctx->I.C, ctx->I.mp, (pl & 1) != 0, (pl & 2) != 0), greedy);
      if (b + fixed <= (size_t)ctx->max_smem) { *smem = b; *place = pl; return true; }
   }
   return false;
}

// blocks that cooperate on one chain: 1 up to ~256 clusters (with the product tree a round's evaluation is a few
// microseconds there: a grid barrier per round costs as much as it saves - vm1084, 219 clusters: 0.254 / 0.192 us per
// proposal on one block, 0.239 / 0.176 on eight), more as the neighbourhood grows (numIters = |VNS|^2, :343-362)
int c2_blocks(const vnsgpu_ctx *ctx)
{
   if (ctx->chain_blocks > 0) return std::min(ctx->chain_blocks, ctx->num_sms);
   const long long c2 = (long long)ctx->I.C * ctx->I.C;
   if (c2 < 65536) return 1;
   return (int)std::min<long long>(ctx->num_sms, std::max<long long>(2, c2 / 32768));
}

// product tree (chain2_tree.cuh): pays once a tour has a few dozen targets; one replica per thread block
static bool c2_tree_wanted(const vnsgpu_ctx *ctx)
{
   if (ctx->chain_tree >= 0) return ctx->chain_tree != 0;
   return ctx->I.C >= 96;
}

// int32 cluster-pair minima + {first sample, size} per cluster on the device (shared with the third form, chain3.cu)
vnsgpu_status c2_ensure_tables(vnsgpu_ctx *ctx)
{
   const InstDev &I = ctx->I;
   if (ctx->d_dmin32) return VNSGPU_OK;
   CK(cudaMalloc(&ctx->d_dmin32, sizeof(int) * (size_t)I.C * I.C));
   k_dmin_to_int<<<(I.C * I.C + 255) / 256, 256, 0, ctx->stream>>>(I.C * I.C, ctx->d_dmin, ctx->d_dmin32);
   LAUNCHED(ctx);
   std::vector<int2> ci(I.C);
   for (int c = 0; c < I.C; ++c) ci[c] = make_int2(ctx->off[c], ctx->m[c]);
   CK(cudaMalloc(&ctx->d_cinfo, sizeof(int2) * I.C));
   CK(cudaMemcpyAsync(ctx->d_cinfo, ci.data(), sizeof(int2) * I.C, cudaMemcpyHostToDevice, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   return VNSGPU_OK;
}

static vnsgpu_status c2_prepare(vnsgpu_ctx *ctx, int ncta, C2Env *E, int place, int nblocks)
{
   const int cinfo_in_smem = place & 1;
   const InstDev &I = ctx->I;
   {
      vnsgpu_status ts = c2_ensure_tables(ctx);
      if (ts) return ts;
   }
   if (ncta > 1 && ctx->c2_ncta < ncta) {
      if (ctx->d_c2grid) { CK(cudaStreamSynchronize(ctx->stream)); CK(cudaFree(ctx->d_c2grid)); }
      ctx->d_c2grid = nullptr;
      ctx->c2_ncta = 0;
      std::vector<unsigned> J;
      build_jump(ncta, J);
      const size_t o_slots = 256, o_jump = o_slots + sizeof(C2Slot) * 2 * ncta, total = o_jump + sizeof(unsigned) * J.size();
      CK(cudaMalloc(&ctx->d_c2grid, total));
      CK(cudaMemsetAsync(ctx->d_c2grid, 0, o_jump, ctx->stream));
      CK(cudaMemcpyAsync((char *)ctx->d_c2grid + o_jump, J.data(), sizeof(unsigned) * J.size(), cudaMemcpyHostToDevice, ctx->stream));
      CK(cudaStreamSynchronize(ctx->stream));
      ctx->c2_ncta = ncta;
      ctx->c2_slots_off = o_slots; ctx->c2_jump_off = o_jump;
   }
   E->I = I; E->Dv = ctx->d_D32; E->dmin = ctx->d_dmin32; E->g_cinfo = (const int2 *)ctx->d_cinfo; E->cinfo_in_smem = cinfo_in_smem;
   E->tree_base = nullptr; E->tree_stride = 0;
   if (c2_tree_wanted(ctx)) {
      int lvl_off[C2_MAXLEV + 2], nlev = 0;
      const long long stride = (long long)c2_tree_layout(I.C + 2, lvl_off, &nlev) * I.mp * I.mp;
      const size_t need = sizeof(int) * (size_t)stride * (size_t)nblocks;
      if (need > ctx->c2_tree_bytes) {
         if (ctx->d_c2tree) { CK(cudaStreamSynchronize(ctx->stream)); CK(cudaFree(ctx->d_c2tree)); }
         ctx->d_c2tree = nullptr; ctx->c2_tree_bytes = 0;
         CK(cudaMalloc(&ctx->d_c2tree, need));
         ctx->c2_tree_bytes = need;
      }
      E->tree_base = ctx->d_c2tree; E->tree_stride = stride;
   }
   E->rows_base = nullptr; E->rows_stride = 0;
   if (!(place & 2)) {
      const long long stride = ((long long)c2_rows_ints(I.C, I.mp) + 63) & ~63ll;
      const size_t need = sizeof(int) * (size_t)stride * (size_t)nblocks;
      if (need > ctx->c2_rows_bytes) {
         if (ctx->d_c2rows) { CK(cudaStreamSynchronize(ctx->stream)); CK(cudaFree(ctx->d_c2rows)); }
         ctx->d_c2rows = nullptr; ctx->c2_rows_bytes = 0;
         CK(cudaMalloc(&ctx->d_c2rows, need));
         ctx->c2_rows_bytes = need;
      }
      E->rows_base = ctx->d_c2rows; E->rows_stride = stride;
   }
   E->grid = C2Grid{nullptr, nullptr, nullptr, nullptr, 1};
   if (ncta > 1) {
      char *b = (char *)ctx->d_c2grid;
      CK(cudaMemsetAsync(b, 0, 256, ctx->stream));            // barrier counter + control words
      E->grid.bar = (unsigned long long *)b;
      E->grid.gctl = (int *)(b + 64);
      E->grid.slots = (C2Slot *)(b + ctx->c2_slots_off);
      E->grid.jump = (const unsigned *)(b + ctx->c2_jump_off);
      E->grid.ncta = ncta;
   }
   return VNSGPU_OK;
}

static int c2_variant(const vnsgpu_ctx *ctx)
{
   int G = ctx->group_lanes;
   if (const char *e = getenv("VNSGPU_CHAIN_G")) { const int g = atoi(e); if (g == 8 || g == 16 || g == 32) G = g; }
   while (G < 32 && (ctx->I.mp + G - 1) / G > (G == 8 ? 4 : 2)) G *= 2;
   const int nr = (ctx->I.mp + G - 1) / G;
   if (G == 8) return nr <= 1 ? 0 : (nr <= 2 ? 1 : 2);
   if (G == 16) return nr <= 1 ? 3 : 4;
   return 5;
}

// the kernels are instantiated in chain2_k8.cu / chain2_k16.cu (which: 0 k_chain2_ls, 1 k_chain2_vns_run, 2 k_chain2_vns_multi)
vnsgpu_status c2k_launch_g8(vnsgpu_ctx *ctx, int variant, int which, int nblocks, bool coop, size_t smem, void **args);
vnsgpu_status c2k_launch_g16(vnsgpu_ctx *ctx, int variant, int which, int nblocks, bool coop, size_t smem, void **args);

static vnsgpu_status c2_dispatch(vnsgpu_ctx *ctx, int which, int nblocks, bool coop, size_t smem, void **args)
{
   const int v = c2_variant(ctx);
   return v < 3 ? c2k_launch_g8(ctx, v, which, nblocks, coop, smem, args) : c2k_launch_g16(ctx, v, which, nblocks, coop, smem, args);
}

vnsgpu_status c2_launch_ls(vnsgpu_ctx *ctx, int path, int kind, int num_changes, double budget, int skip, size_t smem, int place)
{
   const int ncta = c2_blocks(ctx);
   C2Env E;
   vnsgpu_status s = c2_prepare(ctx, ncta, &E, place, ncta);
   if (s) return s;
   PathDev P = ctx->paths[path].d;
   RngState *rng = ctx->d_rng;
   LsStats *st = ctx->d_stats;
   long long *prof = ctx->d_prof;
   void *args[] = {&E, &P, &kind, &num_changes, &budget, &skip, &rng, &st, &prof};
   return c2_dispatch(ctx, 0, ncta, ncta > 1, smem, args);
}

vnsgpu_status c2_launch_vns(vnsgpu_ctx *ctx, int path, const VnsParamsDev &prm_, double *d_ir, double *d_il, double *d_it, size_t smem,
                            int place)
{
   const int ncta = c2_blocks(ctx);
   C2Env E;
   vnsgpu_status s = c2_prepare(ctx, ncta, &E, place, ncta);
   if (s) return s;
   PathDev P = ctx->paths[path].d;
   VnsParamsDev prm = prm_;
   RngState *rng = ctx->d_rng;
   VnsResultDev *res = ctx->d_res;
   long long *prof = ctx->d_prof;
   void *args[] = {&E, &P, &prm, &rng, &res, &d_ir, &d_il, &d_it, &prof};
   return c2_dispatch(ctx, 1, ncta, ncta > 1, smem, args);
}

vnsgpu_status c2_launch_multi(vnsgpu_ctx *ctx, int nruns, void *d_slots_, const VnsParamsDev &prm_, size_t smem, int place)
{
   C2Env E;
   vnsgpu_status s = c2_prepare(ctx, 1, &E, place, nruns);
   if (s) return s;
   C2RunSlot *d_slots = (C2RunSlot *)d_slots_;
   VnsParamsDev prm = prm_;
   void *args[] = {&E, &d_slots, &prm};
   return c2_dispatch(ctx, 2, nruns, false, smem, args);
}

size_t c2_run_slot_bytes() { return sizeof(C2RunSlot); }
void c2_run_slot_fill(void *slot, const PathDev &tour, const RngState &rng)
{
   C2RunSlot *s = (C2RunSlot *)slot;
   s->tour = tour; s->rng = rng;
   memset(&s->res, 0, sizeof(VnsResultDev));
}
const VnsResultDev &c2_run_slot_result(const void *slot) { return ((const C2RunSlot *)slot)->res; }
