// libvnsgpu.so: C ABI (include/vnsgpu.h) over the sm_100a kernels.  No CPU fallback exists: every
// entry point either runs its CUDA kernels or returns an error.  This unit: context + distance table.
#include "ctx.hpp"
#include "table.cuh"

extern "C" {

const char *vnsgpu_version(void) { return "vnsgpu 0.1 (sm_100a)"; }

vnsgpu_status vnsgpu_create(vnsgpu_ctx **out, int device)
{
   *out = nullptr;
   int n = 0;
   if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0 || device >= n) return VNSGPU_E_CUDA;
   vnsgpu_ctx *ctx = new vnsgpu_ctx();
   ctx->device = device;
   *out = ctx;
   CK(cudaSetDevice(device));
   CK(cudaStreamCreateWithFlags(&ctx->own, cudaStreamNonBlocking));
   ctx->stream = ctx->own;
   CK(cudaDeviceGetAttribute(&ctx->num_sms, cudaDevAttrMultiProcessorCount, device));
   CK(cudaDeviceGetAttribute(&ctx->max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
   CK(cudaMalloc(&ctx->d_rng, sizeof(RngState)));
   CK(cudaMalloc(&ctx->d_stats, sizeof(LsStats)));
   CK(cudaMalloc(&ctx->d_first, sizeof(long long)));
   CK(cudaMalloc(&ctx->d_res, sizeof(VnsResultDev)));
   CK(cudaMalloc(&ctx->d_prof, 4 * sizeof(long long)));
   CK(cudaMemset(ctx->d_prof, 0, 4 * sizeof(long long)));
   CK(cudaMalloc(&ctx->d_info, 2 * sizeof(unsigned long long)));
   CK(cudaMalloc(&ctx->d_stop, sizeof(int)));
   if (const char *e = getenv("VNSGPU_VEC_STAGED")) ctx->use_vec_staged = atoi(e) != 0;     // experiments / tests
   return VNSGPU_OK;
}

static void free_instance(vnsgpu_ctx *ctx)
{
   for (PathHost &p : ctx->paths) cudaFree(p.block);
   ctx->paths.clear();
   cudaFree(ctx->d_m); cudaFree(ctx->d_off); cudaFree(ctx->d_clof); cudaFree(ctx->d_reward);
   cudaFree(ctx->d_D); cudaFree(ctx->d_dmin); cudaFree(ctx->d_dmax); cudaFree(ctx->d_seg); cudaFree(ctx->d_D32); cudaFree(ctx->d_D16);
   cudaFree(ctx->d_hist); cudaFree(ctx->d_perm); cudaFree(ctx->d_segmeta); cudaFree(ctx->d_vec); cudaFree(ctx->d_veclb);
   cudaFree(ctx->d_dmin32); cudaFree(ctx->d_cinfo);
   ctx->d_dmin32 = nullptr; ctx->d_cinfo = nullptr;
   ctx->d_segmeta = nullptr; ctx->seg_ok = false; ctx->d_vec = nullptr; ctx->d_veclb = nullptr;
   ctx->d_D32 = nullptr; ctx->d_D16 = nullptr; ctx->d_hist = ctx->d_perm = nullptr; ctx->perm_cap = 0;
   ctx->d_m = ctx->d_off = ctx->d_clof = ctx->d_seg = nullptr;
   ctx->d_reward = ctx->d_D = ctx->d_dmin = ctx->d_dmax = nullptr;
   ctx->have_clusters = ctx->have_table = false;
}

void vnsgpu_destroy(vnsgpu_ctx *ctx)
{
   if (!ctx) return;
   cudaSetDevice(ctx->device);
   cudaStreamSynchronize(ctx->stream);
   vnsgpu_comm_destroy(ctx);
   free_instance(ctx);
   cudaFree(ctx->d_scratch); cudaFree(ctx->d_rng); cudaFree(ctx->d_stats); cudaFree(ctx->d_first);
   cudaFree(ctx->d_res); cudaFree(ctx->d_info); cudaFree(ctx->d_prof); cudaFree(ctx->d_stop); cudaFree(ctx->d_c2grid); cudaFree(ctx->d_c2tree); cudaFree(ctx->d_c2rows);
   cudaFree(ctx->d_genjump); cudaFree(ctx->d_gen); cudaFree(ctx->d_gencands); cudaFree(ctx->d_best);
   if (ctx->s_in) {
      cudaStreamDestroy(ctx->s_in); cudaStreamDestroy(ctx->s_out); cudaEventDestroy(ctx->ev_start); cudaEventDestroy(ctx->ev_out);
      for (int k = 0; k < VNS_PIPE_EVENTS; ++k) { cudaEventDestroy(ctx->ev_in[k]); cudaEventDestroy(ctx->ev_k[k]); }
   }
   if (ctx->own) cudaStreamDestroy(ctx->own);
   if (ctx->hg_exec) cudaGraphExecDestroy(ctx->hg_exec);
   if (ctx->h_first) cudaFreeHost(ctx->h_first);
   if (ctx->h_pick) cudaFreeHost(ctx->h_pick);
   delete ctx;
}

const char *vnsgpu_last_error(const vnsgpu_ctx *ctx) { return ctx ? ctx->err.c_str() : "no context"; }
int64_t vnsgpu_launch_count(const vnsgpu_ctx *ctx) { return ctx->launches; }
int64_t vnsgpu_table_bytes(const vnsgpu_ctx *ctx) { return 8ll * ctx->I.N * ctx->I.N; }

vnsgpu_status vnsgpu_set_stream(vnsgpu_ctx *ctx, void *stream)
{
   ctx->stream = stream ? (cudaStream_t)stream : ctx->own;
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_set_target_reward(vnsgpu_ctx *ctx, double target)
{
   ctx->target_reward = target;
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_set_smem_chain(vnsgpu_ctx *ctx, int enable)
{
   ctx->use_smem_chain = enable != 0;
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_set_chain_form(vnsgpu_ctx *ctx, int form)
{
   ctx->chain_form = form < 0 ? -1 : (form == 3 ? 3 : (form == 2 ? 0 : (form != 0)));
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_set_chain_tree(vnsgpu_ctx *ctx, int mode)
{
   ctx->chain_tree = mode < 0 ? -1 : (mode != 0);
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_set_chain_blocks(vnsgpu_ctx *ctx, int blocks)
{
   ctx->chain_blocks = blocks < 0 ? 0 : blocks;
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_set_segment_cache(vnsgpu_ctx *ctx, int enable)
{
   ctx->use_seg = enable != 0;
   ctx->use_seg_tiles = enable != 2;      // 2: products on, but built by the one-level chain kernel (tests compare both)
   ctx->use_seg_tiles_always = enable == 3;   // 3: two-level build whatever the size
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_set_vector_cache(vnsgpu_ctx *ctx, int enable)
{
   ctx->use_vec_pack = enable != 3;        // 3: vectors always, one cluster per warp in the build (tests compare both)
   ctx->use_vec = enable < 0 ? 0 : (enable > 2 ? 2 : enable);
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_set_u16_table(vnsgpu_ctx *ctx, int enable)
{
   ctx->use_u16 = enable != 0;
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_set_batch_sort(vnsgpu_ctx *ctx, int enable)
{
   ctx->sort_batches = enable != 0;
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_set_int_path(vnsgpu_ctx *ctx, int enable)
{
   ctx->use_int = enable != 0;
   return VNSGPU_OK;
}

int vnsgpu_int_path_active(const vnsgpu_ctx *ctx) { return ctx->use_int && ctx->d_D32 != nullptr && group_kernel_ok(ctx); }

vnsgpu_status vnsgpu_sync(vnsgpu_ctx *ctx)
{
   CK(cudaStreamSynchronize(ctx->stream));
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_set_clusters(vnsgpu_ctx *ctx, int C, const int *m, const double *reward, int start, int goal)
{
   CK(cudaSetDevice(ctx->device));
   if (C < 2 || start < 0 || start >= C || goal < 0 || goal >= C) return fail(ctx, VNSGPU_E_RANGE, "bad cluster description");
   CK(cudaStreamSynchronize(ctx->stream));
   free_instance(ctx);
   ctx->m.assign(m, m + C);
   ctx->reward.assign(reward, reward + C);
   ctx->off.assign(C + 1, 0);
   int mp = 0;
   for (int c = 0; c < C; ++c) {
      if (m[c] < 1) return fail(ctx, VNSGPU_E_RANGE, "empty cluster");
      ctx->off[c + 1] = ctx->off[c] + m[c];
      mp = std::max(mp, m[c]);
   }
   const int N = ctx->off[C];
   if ((long long)mp * N >= (1ll << 31)) return fail(ctx, VNSGPU_E_RANGE, "row band exceeds 2^31 entries");
   std::vector<int> clof(N);
   for (int c = 0; c < C; ++c)
      for (int s = ctx->off[c]; s < ctx->off[c + 1]; ++s) clof[s] = c;
   CK(cudaMalloc(&ctx->d_m, sizeof(int) * C));
   CK(cudaMalloc(&ctx->d_off, sizeof(int) * (C + 1)));
   CK(cudaMalloc(&ctx->d_clof, sizeof(int) * N));
   CK(cudaMalloc(&ctx->d_reward, sizeof(double) * C));
   CK(cudaMalloc(&ctx->d_D, sizeof(double) * (size_t)N * N + 64));
   CK(cudaMalloc(&ctx->d_dmin, sizeof(double) * (size_t)C * C));
   CK(cudaMalloc(&ctx->d_dmax, sizeof(double) * (size_t)C * C));
   CK(cudaMemcpyAsync(ctx->d_m, m, sizeof(int) * C, cudaMemcpyHostToDevice, ctx->stream));
   CK(cudaMemcpyAsync(ctx->d_off, ctx->off.data(), sizeof(int) * (C + 1), cudaMemcpyHostToDevice, ctx->stream));
   CK(cudaMemcpyAsync(ctx->d_clof, clof.data(), sizeof(int) * N, cudaMemcpyHostToDevice, ctx->stream));
   CK(cudaMemcpyAsync(ctx->d_reward, reward, sizeof(double) * C, cudaMemcpyHostToDevice, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   InstDev &I = ctx->I;
   I.C = C; I.N = N; I.mp = mp; I.start = start; I.goal = goal; I.rows = 2 * C + 2; I.exact_int = 0; I.symmetric = 0;
   I.m = ctx->d_m; I.off = ctx->d_off; I.clof = ctx->d_clof; I.reward = ctx->d_reward;
   I.D = ctx->d_D; I.dmin = ctx->d_dmin;
   CK(cudaMalloc(&ctx->d_seg, sizeof(int) * 2 * I.rows));
   // lanes per candidate: the smallest power of two >= the mean cluster size (>= 8); larger clusters
   // take up to 4 registers per lane
   {
      const double mean = (double)N / C;
      int G = mean <= 8 ? 8 : (mean <= 16 ? 16 : 32);
      while (G < 32 && (mp + G - 1) / G > 4) G *= 2;
      ctx->group_lanes = G;
   }
   ctx->have_clusters = true;
   return VNSGPU_OK;
}

static dim3 band_grid(const vnsgpu_ctx *ctx)
{
   const long long band = (long long)ctx->I.mp * ctx->I.N;
   const int gx = (int)std::min<long long>((band + 255) / 256, 4096);
   return dim3(gx, ctx->I.C);
}

} // extern "C"

// The whole table build: ONE fused pass (k_table_pairs: entries as double and as int32, cluster-pair bounds,
// classification), then the decision whether the exact integer path applies.
template <int KIND>
static vnsgpu_status build_table(vnsgpu_ctx *ctx, TableJob J, bool maybe_int)
{
   const InstDev &I = ctx->I;
   const long long nn = (long long)I.N * I.N;
   // room behind the int32 table for the segment products of any path: <= (N + mp)^2 / 2 entries
   const long long segcap = ((long long)I.N + I.mp + 1) * ((long long)I.N + I.mp + 1) / 2;
   if (ctx->d_D32) { CK(cudaFree(ctx->d_D32)); ctx->d_D32 = nullptr; }
   if (ctx->d_dmin32) { CK(cudaFree(ctx->d_dmin32)); CK(cudaFree(ctx->d_cinfo)); ctx->d_dmin32 = nullptr; ctx->d_cinfo = nullptr; }
   if (maybe_int) {   // optimistic: the copy is dropped again if the classification says no
      ctx->seg_ok = true;                                  // block offsets are 64-bit
      // + 8 KB: k_eval_vec lets a row pointer run up to 31 rows past a block.  Everything behind the table is zeroed
      // once, so that such a read always finds a number in [0, 2^29) (zero, or a product of an earlier path state)
      const size_t extra = sizeof(int) * (size_t)(ctx->seg_ok ? segcap : 0) + 8192;
      if (cudaMalloc(&ctx->d_D32, sizeof(int) * (size_t)nn + extra) != cudaSuccess) {
         cudaGetLastError();
         ctx->d_D32 = nullptr;
      } else
         CK(cudaMemsetAsync(ctx->d_D32 + nn, 0, extra, ctx->stream));
   }
   if (ctx->d_D16) { CK(cudaFree(ctx->d_D16)); ctx->d_D16 = nullptr; }
   // + 4 KB of zeros: the third search chain reads whole groups of rows without clamping (chain3.cuh)
   if (ctx->d_D32 && cudaMalloc(&ctx->d_D16, sizeof(unsigned short) * (size_t)nn + 4096) != cudaSuccess) {   // optimistic as well
      cudaGetLastError();
      ctx->d_D16 = nullptr;
   }
   if (ctx->d_D16) CK(cudaMemsetAsync(ctx->d_D16 + nn, 0, 4096, ctx->stream));
   J.D16 = ctx->d_D16;
   CK(cudaMemsetAsync(ctx->d_info, 0, 2 * sizeof(unsigned long long), ctx->stream));
   J.D = ctx->d_D; J.Di = ctx->d_D32; J.dmin = ctx->d_dmin; J.dmax = ctx->d_dmax; J.info = ctx->d_info;
   const long long npairs = (long long)I.C * I.C;
   const int blocks = (int)std::min<long long>((npairs + 7) / 8, (long long)ctx->num_sms * 64);
   cudaEvent_t e0, e1;
   CK(cudaEventCreate(&e0));
   CK(cudaEventCreate(&e1));
   CK(cudaEventRecord(e0, ctx->stream));
   k_table_pairs<KIND><<<blocks, 256, 0, ctx->stream>>>(I, J);
   LAUNCHED(ctx);
   CK(cudaEventRecord(e1, ctx->stream));
   unsigned long long info[2];
   CK(cudaMemcpyAsync(info, ctx->d_info, sizeof(info), cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   CK(cudaEventElapsedTime(&ctx->table_kernel_ms, e0, e1));
   cudaEventDestroy(e0); cudaEventDestroy(e1);
   double mx;
   memcpy(&mx, &info[1], sizeof(double));
   // integers below 2^53 / (layers+1): every partial path sum is exact in double
   ctx->I.exact_int = (info[0] == 0 && mx * (double)(I.rows + 1) < 9007199254740992.0) ? 1 : 0;
   // the int32 copy feeds the DPX path: INF + INF and every partial path sum must stay below 2^31
   if (ctx->d_D32 && !(ctx->I.exact_int && mx * (double)(I.rows + 1) < 536870912.0)) {
      CK(cudaFree(ctx->d_D32));
      ctx->d_D32 = nullptr;
   }
   if (ctx->d_D16 && !(ctx->d_D32 && mx < 65535.0)) {      // the uint16 copy needs every entry below its saturation value
      CK(cudaFree(ctx->d_D16));
      ctx->d_D16 = nullptr;
   }
   ctx->I.symmetric = KIND == 0 ? 1 : 0;   // sqrt(dx*dx+dy*dy): the same bits in both directions
   ctx->have_table = true;
   return VNSGPU_OK;
}

extern "C" {

vnsgpu_status vnsgpu_build_table_euclid(vnsgpu_ctx *ctx, const double *xy, int rule)
{
   CK(cudaSetDevice(ctx->device));
   if (!ctx->have_clusters) return fail(ctx, VNSGPU_E_STATE, "set_clusters first");
   if (rule < 0 || rule > 2) return fail(ctx, VNSGPU_E_RANGE, "unknown edge weight rule");
   const size_t bytes = sizeof(double) * 2 * ctx->I.N;
   vnsgpu_status s = need_scratch(ctx, bytes);
   if (s) return s;
   CK(cudaMemcpyAsync(ctx->d_scratch, xy, bytes, cudaMemcpyHostToDevice, ctx->stream));
   TableJob J{};
   J.pts = (const double2 *)ctx->d_scratch; J.rule = rule;
   return build_table<0>(ctx, J, rule != VNSGPU_RULE_RAW);
}

vnsgpu_status vnsgpu_build_table_dubins(vnsgpu_ctx *ctx, const double *centers, int S, double rho, int do_ceil)
{
   CK(cudaSetDevice(ctx->device));
   if (!ctx->have_clusters) return fail(ctx, VNSGPU_E_STATE, "set_clusters first");
   if (S < 1 || !(rho > 0)) return fail(ctx, VNSGPU_E_RANGE, "bad Dubins parameters");
   const size_t bytes = sizeof(double) * 2 * ctx->I.C;
   vnsgpu_status s = need_scratch(ctx, bytes);
   if (s) return s;
   CK(cudaMemcpyAsync(ctx->d_scratch, centers, bytes, cudaMemcpyHostToDevice, ctx->stream));
   TableJob J{};
   J.pts = (const double2 *)ctx->d_scratch; J.S = S; J.rho = rho; J.do_ceil = do_ceil;
   return build_table<1>(ctx, J, do_ceil != 0);
}

vnsgpu_status vnsgpu_set_table_dense(vnsgpu_ctx *ctx, const double *W)
{
   CK(cudaSetDevice(ctx->device));
   if (!ctx->have_clusters) return fail(ctx, VNSGPU_E_STATE, "set_clusters first");
   const size_t bytes = sizeof(double) * (size_t)ctx->I.N * ctx->I.N;
   vnsgpu_status s = need_scratch(ctx, bytes);
   if (s) return s;
   CK(cudaMemcpyAsync(ctx->d_scratch, W, bytes, cudaMemcpyHostToDevice, ctx->stream));
   TableJob J{};
   J.W = (const double *)ctx->d_scratch;
   return build_table<2>(ctx, J, true);
}

double vnsgpu_table_kernel_ms(const vnsgpu_ctx *ctx) { return (double)ctx->table_kernel_ms; }

vnsgpu_status vnsgpu_get_table_dense(vnsgpu_ctx *ctx, double *W)
{
   CK(cudaSetDevice(ctx->device));
   if (!ctx->have_table) return fail(ctx, VNSGPU_E_STATE, "no table");
   const size_t bytes = sizeof(double) * (size_t)ctx->I.N * ctx->I.N;
   vnsgpu_status s = need_scratch(ctx, bytes);
   if (s) return s;
   k_table_gather<<<band_grid(ctx), 256, 0, ctx->stream>>>(ctx->I, ctx->d_D, (double *)ctx->d_scratch, 1);
   LAUNCHED(ctx);
   CK(cudaMemcpyAsync(W, ctx->d_scratch, bytes, cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_get_bounds(vnsgpu_ctx *ctx, double *dmin, double *dmax)
{
   CK(cudaSetDevice(ctx->device));
   if (!ctx->have_table) return fail(ctx, VNSGPU_E_STATE, "no table");
   const size_t bytes = sizeof(double) * (size_t)ctx->I.C * ctx->I.C;
   if (dmin) CK(cudaMemcpyAsync(dmin, ctx->d_dmin, bytes, cudaMemcpyDeviceToHost, ctx->stream));
   if (dmax) CK(cudaMemcpyAsync(dmax, ctx->d_dmax, bytes, cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   return VNSGPU_OK;
}

} // extern "C"

