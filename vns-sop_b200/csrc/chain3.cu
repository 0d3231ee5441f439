// Host side + kernel instantiations of the third search chain (chain3.cuh): the whole distance table in one SM's shared
// memory.  Called from api_search.cu.
#include "chain3_kernels.inl"
#include "chain2_host.hpp"

// variant: bit 0 = the uint16 copy of the table, bit 1 = the table stays in global memory (read-only path).
// Shared-memory table: when the int32 table (or its uint16 copy) plus the search state stay below the opt-in shared memory of
// one block.  Global table: instances below 96 clusters whose table does not fit (from 96 clusters on the second form's
// product tree wins).  chain_form: -1 (default) takes this form whenever one of the two applies; 3 forces it; 0 / 1 / 2 name
// the other forms.
bool c3_fits(const vnsgpu_ctx *ctx, size_t *smem, int *variant, bool dense)
{
   const bool want = ctx->chain_form == 3 || ctx->chain_form < 0;
   if (!want || !ctx->use_smem_chain || !vnsgpu_int_path_active(ctx) || ctx->I.mp > 32) return false;
   if ((long long)ctx->I.N * ctx->I.N >= (1ll << 19)) return false;      // block offsets are packed into 19 bits
   // static shared memory of the kernels (control block, generator state, the few words of search.cuh's helpers) + margin
   const size_t fixed = sizeof(C3Ctl) + sizeof(RngState) + 2048;
   const size_t greedy = sizeof(double) * 2 * ctx->I.mp * C3_WARPS;       // the greedy start borrows the table's place
   const int forced = getenv("VNSGPU_CHAIN3_U16") ? atoi(getenv("VNSGPU_CHAIN3_U16")) : -1;   // tests: force a variant (0 .. 3)
   // dense (more seeded runs than SMs in one launch): the table stays in global memory when the state is small enough for
   // several thread blocks per SM - they share the table in L1 instead of holding a copy each
   for (int k = 0; k < 4; ++k) {
      const int v = dense ? (k + 2) & 3 : k;
      if (forced >= 0 && v != (forced & 3)) continue;
      const bool u16 = v & 1, tg = (v & 2) != 0;
      if (u16 && !(ctx->d_D16 && ctx->use_u16)) continue;
      if (tg && forced < 0 && (ctx->I.C >= 96 || (!u16 && ctx->d_D16 && ctx->use_u16))) continue;   // global: the smaller copy when there is one
      const size_t b = std::max(c3_smem_bytes(ctx->I.C, ctx->I.N, ctx->I.mp, u16 ? 2 : 4, tg), greedy);
      if (dense && tg && 2 * (b + fixed) > (size_t)ctx->max_smem) continue;                          // not even two blocks per SM: no point
      if (b + fixed <= (size_t)ctx->max_smem) { *smem = b; *variant = v; return true; }
   }
   return false;
}

static vnsgpu_status c3_prepare(vnsgpu_ctx *ctx, C3Env *E, int variant)
{
   const int u16 = variant & 1;
   vnsgpu_status s = c2_ensure_tables(ctx);
   if (s) return s;
   E->I = ctx->I;
   E->tab_g = u16 ? (const void *)ctx->d_D16 : (const void *)ctx->d_D32;
   E->dmin_g = ctx->d_dmin32;
   E->cinfo_g = (const int2 *)ctx->d_cinfo;
   return VNSGPU_OK;
}

// kernels with the table in global memory: chain3_g.cu
vnsgpu_status c3k_pick_global(vnsgpu_ctx *ctx, int u16, int G, int which, int nblocks, size_t smem, void **args);

static vnsgpu_status c3_dispatch(vnsgpu_ctx *ctx, int variant, int which, int nblocks, size_t smem, void **args)
{
   int G = ctx->I.mp <= 8 ? 8 : (ctx->I.mp <= 16 ? 16 : 32);              // one register per lane: the group holds the largest cluster
   if (const char *e = getenv("VNSGPU_CHAIN_G")) { const int g = atoi(e); if ((g == 8 || g == 16 || g == 32) && g >= G) G = g; }
   if (variant == 0) C3_PICK(int);
   if (variant == 1) C3_PICK(unsigned short);
   return c3k_pick_global(ctx, variant & 1, G, which, nblocks, smem, args);
#undef C3_PICK
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
