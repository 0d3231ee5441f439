// Launch helpers of the third search chain's kernels (included by chain3.cu: shared-memory table, and by chain3_g.cu: table in
// global memory - the 36 instantiations compile in two units side by side).
#include "ctx.hpp"
#include "path.cuh"
#include "search.cuh"
#include "eval.cuh"
#include "chain3.cuh"

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

#define C3_PICK(TT)                                                                                                   \
   do {                                                                                                               \
      if (G == 8) return c3k_pick<TT, 8>(ctx, which, nblocks, smem, args);                                            \
      if (G == 16) return c3k_pick<TT, 16>(ctx, which, nblocks, smem, args);                                          \
      return c3k_pick<TT, 32>(ctx, which, nblocks, smem, args);                                                       \
   } while (0)
