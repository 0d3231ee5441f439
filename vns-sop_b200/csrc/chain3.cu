// Host side + kernel instantiations of the third search chain (chain3.cuh): the whole distance table in one SM's shared
// memory.  Called from api_search.cu.
#include "ctx.hpp"
#include "path.cuh"
#include "search.cuh"
#include "eval.cuh"
#include "chain3.cuh"
#include "chain2_host.hpp"

// fits when the int32 table (or its uint16 copy) plus the search state stay below the opt-in shared memory of one
// block.  chain_form: -1 (default) takes this form whenever it fits; 3 forces it; 0 / 1 / 2 name the other forms.
bool c3_fits(const vnsgpu_ctx *ctx, size_t *smem, int *u16)
{
   const bool want = ctx->chain_form == 3 || ctx->chain_form < 0;
   if (!want || !ctx->use_smem_chain || !vnsgpu_int_path_active(ctx) || ctx->I.mp > 32) return false;
   if ((long long)ctx->I.N * ctx->I.N >= (1ll << 19)) return false;      // block offsets are packed into 19 bits
   // static shared memory of the kernels (control block, generator state, the few words of search.cuh's helpers) + margin
   const size_t fixed = sizeof(C3Ctl) + sizeof(RngState) + 2048;
   const size_t greedy = sizeof(double) * 2 * ctx->I.mp * C3_WARPS;       // the greedy start borrows the table's place
   const int forced = getenv("VNSGPU_CHAIN3_U16") ? atoi(getenv("VNSGPU_CHAIN3_U16")) : -1;   // tests: force a copy
   for (int w = 0; w < 2; ++w) {
      if (forced >= 0 && w != (forced & 1)) continue;
      if (w == 1 && !(ctx->d_D16 && ctx->use_u16)) continue;
      const size_t b = std::max(c3_smem_bytes(ctx->I.C, ctx->I.N, ctx->I.mp, w ? 2 : 4), greedy);
      if (b + fixed <= (size_t)ctx->max_smem) { *smem = b; *u16 = w; return true; }
   }
   return false;
}

static vnsgpu_status c3_prepare(vnsgpu_ctx *ctx, C3Env *E, int u16)
{
   vnsgpu_status s = c2_ensure_tables(ctx);
   if (s) return s;
   E->I = ctx->I;
   E->tab_g = u16 ? (const void *)ctx->d_D16 : (const void *)ctx->d_D32;
   E->dmin_g = ctx->d_dmin32;
   E->cinfo_g = (const int2 *)ctx->d_cinfo;
   return VNSGPU_OK;
}

template <typename K> static vnsgpu_status c3k_launch(vnsgpu_ctx *ctx, K kernel, int nblocks, size_t smem, void **args)
{
   CK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
   CK(cudaLaunchKernel((const void *)kernel, dim3(nblocks), dim3(C3_THREADS), args, smem, ctx->stream));
   LAUNCHED(ctx);
   return VNSGPU_OK;
}

// which: 0 k_chain3_ls, 1 k_chain3_vns_run, 2 k_chain3_vns_multi
template <typename TT, int G> static vnsgpu_status c3k_pick(vnsgpu_ctx *ctx, int which, int nblocks, size_t smem, void **args)
{
   if (which == 0) return c3k_launch(ctx, k_chain3_ls<TT, G>, nblocks, smem, args);
   if (which == 1) return c3k_launch(ctx, k_chain3_vns_run<TT, G>, nblocks, smem, args);
   return c3k_launch(ctx, k_chain3_vns_multi<TT, G>, nblocks, smem, args);
}

static vnsgpu_status c3_dispatch(vnsgpu_ctx *ctx, int u16, int which, int nblocks, size_t smem, void **args)
{
   int G = ctx->I.mp <= 8 ? 8 : (ctx->I.mp <= 16 ? 16 : 32);              // one register per lane: the group holds the largest cluster
   if (const char *e = getenv("VNSGPU_CHAIN_G")) { const int g = atoi(e); if ((g == 8 || g == 16 || g == 32) && g >= G) G = g; }
   if (u16) {
      if (G == 8) return c3k_pick<unsigned short, 8>(ctx, which, nblocks, smem, args);
      if (G == 16) return c3k_pick<unsigned short, 16>(ctx, which, nblocks, smem, args);
      return c3k_pick<unsigned short, 32>(ctx, which, nblocks, smem, args);
   }
   if (G == 8) return c3k_pick<int, 8>(ctx, which, nblocks, smem, args);
   if (G == 16) return c3k_pick<int, 16>(ctx, which, nblocks, smem, args);
   return c3k_pick<int, 32>(ctx, which, nblocks, smem, args);
}

vnsgpu_status c3_launch_ls(vnsgpu_ctx *ctx, int path, int kind, int num_changes, double budget, int skip, size_t smem, int u16)
{
   C3Env E;
   vnsgpu_status s = c3_prepare(ctx, &E, u16);
   if (s) return s;
   PathDev P = ctx->paths[path].d;
   RngState *rng = ctx->d_rng;
   LsStats *st = ctx->d_stats;
   long long *prof = ctx->d_prof;
   void *args[] = {&E, &P, &kind, &num_changes, &budget, &skip, &rng, &st, &prof};
   return c3_dispatch(ctx, u16, 0, 1, smem, args);
}

vnsgpu_status c3_launch_vns(vnsgpu_ctx *ctx, int path, const VnsParamsDev &prm_, double *d_ir, double *d_il, double *d_it, size_t smem,
                            int u16)
{
   C3Env E;
   vnsgpu_status s = c3_prepare(ctx, &E, u16);
   if (s) return s;
   PathDev P = ctx->paths[path].d;
   VnsParamsDev prm = prm_;
   RngState *rng = ctx->d_rng;
   VnsResultDev *res = ctx->d_res;
   long long *prof = ctx->d_prof;
   void *args[] = {&E, &P, &prm, &rng, &res, &d_ir, &d_il, &d_it, &prof};
   return c3_dispatch(ctx, u16, 1, 1, smem, args);
}

// d_slots: the second form's slot layout (c2_run_slot_fill / c2_run_slot_result of chain2.cu)
vnsgpu_status c3_launch_multi(vnsgpu_ctx *ctx, int nruns, void *d_slots_, const VnsParamsDev &prm_, size_t smem, int u16)
{
   static_assert(sizeof(C3RunSlot) == sizeof(PathDev) + sizeof(RngState) + sizeof(VnsResultDev), "slot layout");
   if (c2_run_slot_bytes() != sizeof(C3RunSlot)) return fail(ctx, VNSGPU_E_STATE, "run slot layouts differ");
   C3Env E;
   vnsgpu_status s = c3_prepare(ctx, &E, u16);
   if (s) return s;
   C3RunSlot *d_slots = (C3RunSlot *)d_slots_;
   VnsParamsDev prm = prm_;
   void *args[] = {&E, &d_slots, &prm};
   return c3_dispatch(ctx, u16, 2, nruns, smem, args);
}
