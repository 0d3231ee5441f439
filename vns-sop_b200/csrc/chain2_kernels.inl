// Kernel instantiations of the second search chain for one group width (included by chain2_k8.cu with C2K_G8 and by
// chain2_k16.cu without: the 18 instantiations compile in two units side by side instead of one after the other).
#include "ctx.hpp"
#include "path.cuh"
#include "search.cuh"
#include "eval.cuh"
#include "chain2.cuh"
#include "chain2_host.hpp"

template <typename K> static vnsgpu_status c2k_launch(vnsgpu_ctx *ctx, K kernel, int nblocks, bool coop, size_t smem, void **args)
{
   CK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
   if (coop) CK(cudaLaunchCooperativeKernel((const void *)kernel, dim3(nblocks), dim3(C2_THREADS), args, smem, ctx->stream));
   else CK(cudaLaunchKernel((const void *)kernel, dim3(nblocks), dim3(C2_THREADS), args, smem, ctx->stream));
   LAUNCHED(ctx);
   return VNSGPU_OK;
}

// which: 0 k_chain2_ls, 1 k_chain2_vns_run, 2 k_chain2_vns_multi
template <int G, int NR> static vnsgpu_status c2k_pick(vnsgpu_ctx *ctx, int which, int nblocks, bool coop, size_t smem, void **args)
{
   if (which == 0) return c2k_launch(ctx, k_chain2_ls<G, NR>, nblocks, coop, smem, args);
   if (which == 1) return c2k_launch(ctx, k_chain2_vns_run<G, NR>, nblocks, coop, smem, args);
   return c2k_launch(ctx, k_chain2_vns_multi<G, NR>, nblocks, coop, smem, args);
}

#ifdef C2K_G8
vnsgpu_status c2k_launch_g8(vnsgpu_ctx *ctx, int variant, int which, int nblocks, bool coop, size_t smem, void **args)
{
   if (variant == 0) return c2k_pick<8, 1>(ctx, which, nblocks, coop, smem, args);
   if (variant == 1) return c2k_pick<8, 2>(ctx, which, nblocks, coop, smem, args);
   return c2k_pick<8, 4>(ctx, which, nblocks, coop, smem, args);
}
#else
vnsgpu_status c2k_launch_g16(vnsgpu_ctx *ctx, int variant, int which, int nblocks, bool coop, size_t smem, void **args)
{
   if (variant == 3) return c2k_pick<16, 1>(ctx, which, nblocks, coop, smem, args);
   if (variant == 4) return c2k_pick<16, 2>(ctx, which, nblocks, coop, smem, args);
   return c2k_pick<32, 1>(ctx, which, nblocks, coop, smem, args);
}
#endif
