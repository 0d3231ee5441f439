// Host entry points of the second search chain (chain2.cu), used by api_search.cu.
#pragma once
#include "ctx.hpp"

vnsgpu_status c2_ensure_tables(vnsgpu_ctx *ctx);
bool c2_fits(const vnsgpu_ctx *ctx, size_t *smem, int *place);
int c2_blocks(const vnsgpu_ctx *ctx);
vnsgpu_status c2_launch_ls(vnsgpu_ctx *ctx, int path, int kind, int num_changes, double budget, int skip, size_t smem, int place);
vnsgpu_status c2_launch_vns(vnsgpu_ctx *ctx, int path, const VnsParamsDev &prm, double *d_ir, double *d_il, double *d_it, size_t smem,
                            int place);
vnsgpu_status c2_launch_multi(vnsgpu_ctx *ctx, int nruns, void *d_slots, const VnsParamsDev &prm, size_t smem, int place);
size_t c2_run_slot_bytes();
void c2_run_slot_fill(void *slot, const PathDev &tour, const RngState &rng);
const VnsResultDev &c2_run_slot_result(const void *slot);

// third form (chain3.cu): the whole search state (and the distance table when it fits) in shared memory.  u16 = variant:
// bit 0 the uint16 copy of the table, bit 1 the table stays in global memory
bool c3_fits(const vnsgpu_ctx *ctx, size_t *smem, int *u16, bool dense = false);
vnsgpu_status c3_launch_ls(vnsgpu_ctx *ctx, int path, int kind, int num_changes, double budget, int skip, size_t smem, int u16);
vnsgpu_status c3_launch_vns(vnsgpu_ctx *ctx, int path, const VnsParamsDev &prm, double *d_ir, double *d_il, double *d_it, size_t smem,
                            int u16);
vnsgpu_status c3_launch_multi(vnsgpu_ctx *ctx, int nruns, void *d_slots, const VnsParamsDev &prm, size_t smem, int u16);
