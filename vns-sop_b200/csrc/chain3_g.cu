// third search chain, kernels that read the distance table from global memory (see chain3_kernels.inl)
#include "chain3_kernels.inl"

vnsgpu_status c3k_pick_global(vnsgpu_ctx *ctx, int u16, int G, int which, int nblocks, size_t smem, void **args)
{
   if (u16) C3_PICK(C3G<unsigned short>);
   C3_PICK(C3G<int>);
}
