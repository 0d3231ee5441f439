// C ABI: path handles + batched evaluation
#include "ctx.hpp"
#include "path.cuh"
#include "search.cuh"
#include "eval.cuh"
#include "vec.cuh"
#include "kernels.cuh"
#include "chain2_host.hpp"

extern "C" {

vnsgpu_status vnsgpu_path_create(vnsgpu_ctx *ctx, int *path)
{
   CK(cudaSetDevice(ctx->device));
   if (!ctx->have_table) return fail(ctx, VNSGPU_E_STATE, "build or set the distance table first");
   const InstDev &I = ctx->I;
   const size_t rows = I.rows, mp = I.mp;
   size_t o_hdr = 0;
   size_t o_hd = o_hdr + sizeof(int) * HDR_INTS;
   size_t o_tg = al16(o_hd + sizeof(double) * HD_DBLS);
   size_t o_vns = al16(o_tg + sizeof(int) * rows);
   size_t o_F = al16(o_vns + sizeof(int) * rows);
   size_t o_B = al16(o_F + sizeof(double) * rows * mp);
   size_t o_minF = al16(o_B + sizeof(double) * rows * mp);
   size_t o_minB = al16(o_minF + sizeof(double) * rows);
   size_t total = al16(o_minB + sizeof(double) * rows);
   PathHost ph;
   CK(cudaMalloc(&ph.block, total));
   ph.bytes = total;
   CK(cudaMemsetAsync(ph.block, 0, total, ctx->stream));
   char *b = (char *)ph.block;
   ph.d.hdr = (int *)(b + o_hdr); ph.d.hd = (double *)(b + o_hd);
   ph.d.tg = (int *)(b + o_tg); ph.d.vns = (int *)(b + o_vns);
   ph.d.F = (double *)(b + o_F); ph.d.B = (double *)(b + o_B);
   ph.d.minF = (double *)(b + o_minF); ph.d.minB = (double *)(b + o_minB);
   int slot = -1;   // reuse a destroyed handle's slot
   for (size_t k = 0; k < ctx->paths.size() && slot < 0; ++k)
      if (!ctx->paths[k].block) slot = (int)k;
   if (slot < 0) {
      ctx->paths.push_back(ph);
      slot = (int)ctx->paths.size() - 1;
   } else
      ctx->paths[slot] = ph;
   *path = slot;
   k_path_update<<<1, 64, 0, ctx->stream>>>(ctx->I, ph.d);   // empty path start -> goal
   LAUNCHED(ctx);
   CK(cudaStreamSynchronize(ctx->stream));
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_path_destroy(vnsgpu_ctx *ctx, int path)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   CK(cudaStreamSynchronize(ctx->stream));
   CK(cudaFree(ctx->paths[path].block));
   ctx->paths[path] = PathHost();
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_path_copy(vnsgpu_ctx *ctx, int dst, int src)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, dst) || !path_ok(ctx, src)) return VNSGPU_E_STATE;
   k_path_copy<<<1, 256, 0, ctx->stream>>>(ctx->I, ctx->paths[dst].d, ctx->paths[src].d);
   LAUNCHED(ctx);
   ctx->paths[dst].hdr_ok = false;
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_path_set(vnsgpu_ctx *ctx, int path, const int *vns, int nvns, int T)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   const InstDev &I = ctx->I;
   if (nvns < 0 || nvns > I.C || T < 0 || T > nvns) return fail(ctx, VNSGPU_E_RANGE, "bad T / nvns");
   for (int k = 0; k < nvns; ++k)
      if (vns[k] < 0 || vns[k] >= I.C) return fail(ctx, VNSGPU_E_RANGE, "cluster id out of range");
   const PathDev &d = ctx->paths[path].d;
   int hdr[HDR_INTS] = {T, nvns, 0, 0, 0, 0, 0, 0};
   CK(cudaMemcpyAsync(d.hdr, hdr, sizeof(hdr), cudaMemcpyHostToDevice, ctx->stream));
   if (nvns) CK(cudaMemcpyAsync(d.vns, vns, sizeof(int) * nvns, cudaMemcpyHostToDevice, ctx->stream));
   if (T) CK(cudaMemcpyAsync(d.tg, vns, sizeof(int) * T, cudaMemcpyHostToDevice, ctx->stream));
   // update() (vns_sop_path.cpp:1321-1404) walks T layers one after the other; on long tours over an integer table every
   // layer's rows come from the segment products instead, all layers in parallel (VNSGPU_APPLY_PRODUCTS: tests force either)
   bool par = path_products_available(ctx) && T >= 128;
   if (const char *e = getenv("VNSGPU_APPLY_PRODUCTS")) par = path_products_available(ctx) && atoi(e) != 0 && T >= 1;
   if (par) {
      vnsgpu_status ps = path_rows_from_products(ctx, path, ctx->stream, true);
      if (ps) return ps;
   } else {
      k_path_update<<<1, 64, 0, ctx->stream>>>(ctx->I, d);
      LAUNCHED(ctx);
   }
   return check_path_err(ctx, path);
}

vnsgpu_status vnsgpu_path_add_point(vnsgpu_ctx *ctx, int path, int cluster, int idx)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   PathHdr h;
   vnsgpu_status s = read_hdr(ctx, path, &h);
   if (s) return s;
   const int T = h.hdr[HDR_T];
   if (idx > T || idx < 0) return fail(ctx, VNSGPU_E_RANGE, "addPoint idx exceeds maximum");   // vns_sop_path.cpp:69-72
   if (cluster < 0 || cluster >= ctx->I.C || T + 3 > ctx->I.rows) return fail(ctx, VNSGPU_E_RANGE, "addPoint cluster / capacity");
   // the batched evaluation (layer maps, endpoint vectors, segment products, search chains) is sized for tours of at most C
   // targets - what the solver can build (every cluster at most once); the reference's addPoint itself has no such limit
   if (T + 1 > ctx->I.C) return fail(ctx, VNSGPU_E_RANGE, "addPoint: more targets than clusters");
   k_path_edit<<<1, 256, 0, ctx->stream>>>(ctx->I, ctx->paths[path].d, 0, cluster, idx);
   LAUNCHED(ctx);
   return check_path_err(ctx, path);
}

vnsgpu_status vnsgpu_path_remove_point(vnsgpu_ctx *ctx, int path, int idx)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   PathHdr h;
   vnsgpu_status s = read_hdr(ctx, path, &h);
   if (s) return s;
   if (idx >= h.hdr[HDR_T] || idx < 0) return fail(ctx, VNSGPU_E_RANGE, "removePoint idx exceeds maximum");   // :158-161
   k_path_edit<<<1, 256, 0, ctx->stream>>>(ctx->I, ctx->paths[path].d, 1, 0, idx);
   LAUNCHED(ctx);
   return check_path_err(ctx, path);
}

vnsgpu_status vnsgpu_path_get(vnsgpu_ctx *ctx, int path, int *T, int *targets, int *nvns, int *vns, double *length,
                              double *length_rear, double *reward)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   PathHdr h;
   vnsgpu_status s = read_hdr(ctx, path, &h);
   if (s) return s;
   const PathDev &d = ctx->paths[path].d;
   if (T) *T = h.hdr[HDR_T];
   if (nvns) *nvns = h.hdr[HDR_NVNS];
   if (length) *length = h.hd[HD_LEN];
   if (length_rear) *length_rear = h.hd[HD_LEN_REAR];
   if (reward) *reward = h.hd[HD_REWARD];
   if (targets && h.hdr[HDR_T]) CK(cudaMemcpyAsync(targets, d.tg, sizeof(int) * h.hdr[HDR_T], cudaMemcpyDeviceToHost, ctx->stream));
   if (vns && h.hdr[HDR_NVNS]) CK(cudaMemcpyAsync(vns, d.vns, sizeof(int) * h.hdr[HDR_NVNS], cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_path_get_solution(vnsgpu_ctx *ctx, int path, int *cluster, int *node, double *length_calc)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   PathHdr h;
   vnsgpu_status s = read_hdr(ctx, path, &h);
   if (s) return s;
   const int n = h.hdr[HDR_T] + 2;
   s = need_scratch(ctx, sizeof(int) * 2 * n + 16);
   if (s) return s;
   double *d_calc = (double *)ctx->d_scratch;
   int *d_cl = (int *)((char *)ctx->d_scratch + 16), *d_nd = d_cl + n;
   k_path_solution<<<1, 32, 0, ctx->stream>>>(ctx->I, ctx->paths[path].d, d_cl, d_nd, d_calc);
   LAUNCHED(ctx);
   if (cluster) CK(cudaMemcpyAsync(cluster, d_cl, sizeof(int) * n, cudaMemcpyDeviceToHost, ctx->stream));
   if (node) CK(cudaMemcpyAsync(node, d_nd, sizeof(int) * n, cudaMemcpyDeviceToHost, ctx->stream));
   if (length_calc) CK(cudaMemcpyAsync(length_calc, d_calc, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_path_get_dp(vnsgpu_ctx *ctx, int path, double *F, double *B, int stride)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   if (stride < ctx->I.mp) return fail(ctx, VNSGPU_E_RANGE, "stride < max cluster size");
   PathHdr h;
   vnsgpu_status s = read_hdr(ctx, path, &h);
   if (s) return s;
   const PathDev &d = ctx->paths[path].d;
   const size_t rows = h.hdr[HDR_T] + 2, w = sizeof(double) * ctx->I.mp;
   if (F) CK(cudaMemcpy2DAsync(F, sizeof(double) * stride, d.F, w, w, rows, cudaMemcpyDeviceToHost, ctx->stream));
   if (B) CK(cudaMemcpy2DAsync(B, sizeof(double) * stride, d.B, w, w, rows, cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   return VNSGPU_OK;
}

static int eval_grid(const vnsgpu_ctx *ctx, int64_t n)
{
   const int64_t wpb = VNS_EVAL_THREADS / 32;
   const int64_t need = (n + wpb - 1) / wpb;
   const int64_t cap = (int64_t)ctx->num_sms * 8;   // 8 resident CTAs of 256 threads per SM
   return (int)std::max<int64_t>(1, std::min(need, cap));
}

// one evaluation of n candidates on `st`; idx0 = index of cands[0] in the caller's batch (first-accept
// key); slot = which of the VNS_PIPE_EVENTS sort-scratch regions to use (chunks in flight concurrently)
#define VNS_SORT_MIN 8192
#define VNS_SEG_MIN 16384
// integer tables: build the segment products of the path's current state (exact, see eval.cuh SegCache)
static bool seg_wanted(const vnsgpu_ctx *ctx, int64_t n)
{
   // block by block a candidate costs ~T/3 block steps, the products ~T^2/2 once per state: they pay from a batch of a
   // few T on.  Fixed floor for small instances (a handful of launches costs more than it saves there), n C >= 2^20 for
   // large ones (64 k-sample DOP state, 1024 candidates: 3.3 ms block by block, 1.4 ms with the products)
   return ctx->use_seg && vnsgpu_int_path_active(ctx) && ctx->seg_ok && (n >= VNS_SEG_MIN || n * (int64_t)ctx->I.C >= (1 << 20));
}

// evaluation form of a batch: 0 block by block, 1 segment products, 2 endpoint vectors + segment products.
// The vectors cost 2 C (T+2) + 4 T (T+2) block steps per path state (T is only known on the device; typically
// T ~ C / 3) and save ~4 block steps per candidate.  Measured break-even on the HBM-sized instances (batch sweeps
// r1t, r2d): 0.6 M candidates at C = 2002, m = 10 and 0.2 M at C = 2002, m = 32 - hence batch >= C^2 / 4, C^2 / 16 for
// clusters of more than 16 samples (the 4 KB block gathers of the product form cost more than their bytes).
static int eval_mode(const vnsgpu_ctx *ctx, int64_t n)
{
   if (!seg_wanted(ctx, n)) return 0;
   const int64_t C = ctx->I.C;
   if (ctx->use_vec && ctx->I.mp <= 32 && ((ctx->I.mp > 16 ? 16 : 4) * n >= C * C || ctx->use_vec == 2) && vec_ints((int)C, (int)C + 2, ctx->I.mp) < (1ull << 31)) return 2;
   return 1;
}

// A batch of tryToAdd candidates only (the greedy grid of best.cu) on a large instance: endpoint vectors whatever the batch
// size - ADD is a dot-min of U and W, so the build stops after k_vec_uw (no pair vectors) and nothing is sorted
static int eval_mode_hinted(vnsgpu_ctx *ctx, int64_t n)
{
   const bool add_only = ctx->batch_add_only;
   ctx->batch_add_only = false;
   int mode = eval_mode(ctx, n);
   const int64_t C = ctx->I.C;
   ctx->vec_skip_pairs = add_only && C >= 512 && seg_wanted(ctx, n) && ctx->use_vec && ctx->I.mp <= 32 &&
                         vec_ints((int)C, (int)C + 2, ctx->I.mp) < (1ull << 31);
   if (ctx->vec_skip_pairs) mode = 2;
   return mode;
}

// T + 2 as the host last read it - when it did and the launch is not being captured into a graph (a graph is replayed on
// later states) - else `bound`.  Only for grids of kernels whose result does not depend on the grid.
static int known_layers(vnsgpu_ctx *ctx, int path, cudaStream_t st, int bound)
{
   cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
   if (cudaStreamIsCapturing(st, &cap) != cudaSuccess) { cudaGetLastError(); return bound; }
   const PathHost &ph = ctx->paths[path];
   return (ph.hdr_ok && cap == cudaStreamCaptureStatusNone) ? std::min(bound, ph.hdr_cache.hdr[HDR_T] + 2) : bound;
}

static vnsgpu_status prepare_seg(vnsgpu_ctx *ctx, int path, cudaStream_t st)
{
   const InstDev &I = ctx->I;
   const int R = I.rows + 3;
   if (!ctx->d_segmeta) CK(cudaMalloc(&ctx->d_segmeta, sizeof(int) * 6 * R));
   PhaseTimer pt(ctx, VNSGPU_PHASE_SEG, st);
   k_seg_offsets<<<1, 256, sizeof(int) * 2 * R, st>>>(I, ctx->paths[path].d, ctx->d_segmeta);
   LAUNCHED(ctx);
   const int nst = I.mp <= 22 ? 16 : 8;                      // ring depth: <= 32 KB of staged table blocks
   const size_t smem = seg_build_smem(I.rows, I.mp, nst);
   if (smem <= (size_t)ctx->max_smem && I.mp <= 32 && I.N < 65536) {
      const int grid = std::min(I.rows, I.C + 2);
      // T is only known on the device: the two-level build is chosen from the bound of T (a tour visits every cluster at
      // most once); on a short tour its extra launches find nothing to do
      // (below ~256 layers the one-level chains are as fast: vm1084, 219 clusters: 0.12 ms)
      // and with clusters of ~32 samples both forms are bound by the m^3 add-mins of the products themselves: 64 k-sample DOP
      // state, T = 600: 0.99 ms one-level, 1.24 ms two-level; 20 k-node state, m = 10, T = 793: 0.78 -> 0.33 ms)
      const int tile = (ctx->use_seg_tiles && ((I.C + 2 > 8 * SEG_TILE && I.mp <= 16) || ctx->use_seg_tiles_always)) ? SEG_TILE : 0;
      if (nst == 16) {
         if (smem > 48 * 1024) CK(cudaFuncSetAttribute(k_seg_build<16, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
         k_seg_build<16, 4><<<grid, VNS_SEG_THREADS, smem, st>>>(I, ctx->paths[path].d, ctx->d_D32, ctx->d_segmeta, tile);
      } else {
         if (smem > 48 * 1024) CK(cudaFuncSetAttribute(k_seg_build<8, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
         k_seg_build<8, 1><<<grid, VNS_SEG_THREADS, smem, st>>>(I, ctx->paths[path].d, ctx->d_D32, ctx->d_segmeta, tile);
      }
      LAUNCHED(ctx);
      if (tile) {
         const int ntiles = (grid + SEG_TILE - 1) / SEG_TILE + 1;
         const size_t sm1 = sizeof(int) * (size_t)(2 + SEG_TILES_RING) * I.mp * I.mp;      // mp <= 16 here: 8 KB
         k_seg_tiles<<<ntiles, 256, sm1, st>>>(I, ctx->paths[path].d, ctx->d_D32, ctx->d_segmeta);
         LAUNCHED(ctx);
         // the cross products stride over their tasks: grids from the header as last read (see prepare_eval), else the bound
         const int lk = known_layers(ctx, path, st, grid);
         const int force_g = getenv("VNSGPU_SEG_CROSS_GRID") ? atoi(getenv("VNSGPU_SEG_CROSS_GRID")) : 0;     // tests: a grid below the tasks
         const unsigned g3 = force_g > 0 ? force_g : (unsigned)((lk + 1 + 7) / 8);
         const long long ntk = (lk + SEG_TILE - 1) / SEG_TILE + 1;                                            // >= the kernel's nt
         const unsigned g4 = force_g > 0 ? force_g : (unsigned)std::max(1ll, (SEG_TILE / 2 * ntk * (ntk - 1) + 7) / 8);
#define VNS_SEG_CROSS(GG, KU)                                                                                         \
   do {                                                                                                               \
      k_seg_cross<0, GG, KU><<<g3, 256, 0, st>>>(I, ctx->paths[path].d, ctx->d_D32, ctx->d_segmeta);                  \
      LAUNCHED(ctx);                                                                                                  \
      k_seg_cross<1, GG, KU><<<g4, 256, 0, st>>>(I, ctx->paths[path].d, ctx->d_D32, ctx->d_segmeta);                  \
   } while (0)
         if (I.mp <= 8) VNS_SEG_CROSS(8, 8);
         else if (I.mp <= 10 && ctx->use_vec_pack) VNS_SEG_CROSS(10, 12);
         else if (I.mp <= 12) VNS_SEG_CROSS(16, 12);
         else if (I.mp <= 16) VNS_SEG_CROSS(16, 16);
         else VNS_SEG_CROSS(32, 32);
#undef VNS_SEG_CROSS
      }
   } else
      k_seg_build_g<<<I.rows, 128, 0, st>>>(I, ctx->paths[path].d, ctx->d_D32, ctx->d_segmeta);
   LAUNCHED(ctx);
   return VNSGPU_OK;
}

// ---- every layer's F / B row from the segment products, all layers in parallel (integer tables) ------------------------
// update() (vns_sop_path.cpp:1321-1404) and updateAfterInsert / updateAfterRemove (:1406-1540) walk the layers one after the
// other: T dependent steps (k_path_update: 0.95 ms on the 20 000-node state).  With the exact products M(a,b) of the path's
// CURRENT targets in place, F[l] = F[0] (x) M(0,l) and B[l] = M(l,S) (x) B[S] are one block step per layer, independent of
// each other: one warp per (layer, direction).  Exact because integer min-plus products are associative.
static __global__ void __launch_bounds__(256) k_rows_from_seg(InstDev I, PathDev P, const int *__restrict__ Dv, const int *__restrict__ meta)
{
   const int T = P.hdr[HDR_T], S = T + 1, mp = I.mp;
   const SegMeta sm = seg_meta(I, meta);
   const int lane = threadIdx.x & 31;
   const int nw = (int)((gridDim.x * blockDim.x) >> 5);
   for (int job = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5); job < 2 * (S + 1); job += nw) {
      const bool fwd = job <= S;
      const int l = fwd ? job : job - (S + 1);
      const int ml = sm.m(l);
      double best = VNS_M;
      if (fwd) {
         if (l == 0) {
            for (int j = lane; j < ml; j += 32) best = fmin(best, P.F[j]);
         } else {
            const int m0 = sm.m(0);
            const int *M = Dv + sm.off(0, l);
            for (int j = lane; j < ml; j += 32) {
               int acc = VNS_INT_INF;
               for (int i = 0; i < m0; ++i) acc = __viaddmin_s32((int)P.F[i], __ldg(M + (size_t)i * ml + j), acc);
               P.F[(size_t)l * mp + j] = (double)acc;
               best = fmin(best, (double)acc);
            }
         }
         best = warp_min(best);
         if (lane == 0) { P.minF[l] = best; if (l == S) P.hd[HD_LEN] = best; }
      } else {
         const double *BS = P.B + (size_t)S * mp;
         if (l == S) {
            for (int j = lane; j < ml; j += 32) best = fmin(best, BS[j]);
         } else {
            const int mS = sm.m(S);
            const int *M = Dv + sm.off(l, S);
            for (int i = lane; i < ml; i += 32) {
               int acc = VNS_INT_INF;
               for (int j = 0; j < mS; ++j) acc = __viaddmin_s32((int)BS[j], __ldg(M + (size_t)i * mS + j), acc);
               P.B[(size_t)l * mp + i] = (double)acc;
               best = fmin(best, (double)acc);
            }
         }
         best = warp_min(best);
         if (lane == 0) { P.minB[l] = best; if (l == 0) P.hd[HD_LEN_REAR] = best; }
      }
   }
}

// the rows that do not come from a product: F[0] = 0 (heuristic_types.h:184-187), B[S] = 0, and getReward
static __global__ void __launch_bounds__(64) k_rows_ends(InstDev I, PathDev P)
{
   const int T = P.hdr[HDR_T], S = T + 1;
   const int m0 = I.m[I.start], mS = I.m[I.goal];
   for (int j = threadIdx.x; j < m0; j += blockDim.x) P.F[j] = 0.0;
   for (int j = threadIdx.x; j < mS; j += blockDim.x) P.B[(size_t)S * I.mp + j] = 0.0;
   if (threadIdx.x == 0) P.hd[HD_REWARD] = path_reward(I, P.tg, T);
}

// addPoint's consistency check (vns_sop_path.cpp:80-88) behind k_rows_from_seg
static __global__ void k_rows_check(PathDev P)
{
   if (fabs(P.hd[HD_LEN] - P.hd[HD_LEN_REAR]) > 0.01 && !P.hdr[HDR_ERR]) P.hdr[HDR_ERR] = 4;
}

static vnsgpu_status prepare_eval(vnsgpu_ctx *ctx, int path, int mode, cudaStream_t st)
{
   if (mode == 0) return VNSGPU_OK;
   // the caller built the products of exactly this state a moment ago (path_rows_from_products): not again
   const bool fresh = ctx->seg_fresh_path == path;
   ctx->seg_fresh_path = -1;
   vnsgpu_status s = fresh ? VNSGPU_OK : prepare_seg(ctx, path, st);
   if (s || mode == 1) return s;
   const InstDev &I = ctx->I;
   const int LS = I.C + 2;
   if (!ctx->d_vec) {
      CK(cudaMalloc(&ctx->d_vec, sizeof(int) * vec_ints(I.C, LS, I.mp)));
      CK(cudaMalloc(&ctx->d_veclb, sizeof(int) * (vec_lb_ints(I.C, LS) + (size_t)2 * LS * I.mp)));   // + F and B rows as int32
      CK(cudaMemsetAsync(ctx->d_vec, 0, sizeof(int) * vec_ints(I.C, LS, I.mp), st));
      CK(cudaMemsetAsync(ctx->d_veclb, 0, sizeof(int) * vec_lb_ints(I.C, LS), st));
   }
   // T is only known on the device: grids sized for its bound (T + 2 <= C), surplus warps return at once
   const int Smax = std::min(I.rows, I.C);
   // the builds stream table blocks and nothing else: from the uint16 copy when there is one (half the bytes)
   const bool u16 = ctx->d_D16 && ctx->use_u16;
   const size_t slot = vec_slot_bytes(I.mp, u16 ? sizeof(unsigned short) : sizeof(int));
   const int thr = 2 * slot * 8 <= 40 * 1024 ? 256 : 128;      // 2 staged blocks per warp: <= 40 KB per CTA
   const int wpc = thr / 32;
   const size_t smem = 2 * slot * wpc;
   // grid.y = layer lanes (the kernels stride over the layers): about eight waves of resident blocks.  The warps' work is
   // uneven (cluster sizes, cache hits), so the block scheduler's dynamic assignment is worth more than a grid that is
   // resident at once: measured on the 20 k-node / 64 k-sample states, 1 wave 0.557 / 1.98 ms, the old 1.35 waves 0.483 /
   // 1.67 ms, 4 waves 0.477 / 1.60 ms, 8 waves 0.464 / 1.56 ms, 16 waves the same (profiles/r4y_waves.txt).
   // k_vec_pairs' grid.x covers the layers: from the header as last read back when there is one and the launch is not being
   // captured into a graph (a graph is replayed on later states), else from the bound - the surplus warps return at once
   const int Sknown = known_layers(ctx, path, st, Smax + 1) - 1;
   auto layer_lanes = [&](auto kern, size_t sm, int gx) {
      int nb = 0;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, thr, sm) != cudaSuccess || nb < 1) { cudaGetLastError(); nb = 1; }
      return std::max(1, std::min(Smax, 8 * nb * ctx->num_sms / gx));
   };
   const int force_gxp = getenv("VNSGPU_VEC_PAIRS_GX") ? atoi(getenv("VNSGPU_VEC_PAIRS_GX")) : 0;   // tests: a grid.x below the layers
   const int gxu = (I.C + wpc - 1) / wpc, gxp = force_gxp > 0 ? force_gxp : (Sknown + 1 + wpc - 1) / wpc;
   const int gyu = u16 ? layer_lanes(k_vec_uw<unsigned short>, smem, gxu) : layer_lanes(k_vec_uw<int>, smem, gxu);
   const size_t smem_p = 2 * vec_pairs_slot_bytes(I.mp, u16 ? sizeof(unsigned short) : sizeof(int)) * wpc;     // blocks + their four vectors
   const int gyp = u16 ? layer_lanes(k_vec_pairs<unsigned short>, smem_p, gxp) : layer_lanes(k_vec_pairs<int>, smem_p, gxp);
   // cluster-pair minima as int32 (shared with the chain forms; the call synchronises ctx->stream when it builds them)
   if (!ctx->d_dmin32) { vnsgpu_status ds = c2_ensure_tables(ctx); if (ds) return ds; }
   // F / B rows as int32 for the build kernels: from 512 clusters on (below that the conversions inside the kernels cost
   // less than one more launch: krob200's vec phase 14.4 us without, 17.0 us with it)
   int *fbi = I.C >= 512 ? ctx->d_veclb + vec_lb_ints(I.C, LS) : nullptr;
   PhaseTimer pt(ctx, VNSGPU_PHASE_VEC, st);
   if (fbi) {
      k_rows_to_int<<<std::min(64, (Smax + 2) * I.mp / 256 + 1), 256, 0, st>>>(I, ctx->paths[path].d, fbi, LS);
      LAUNCHED(ctx);
      k_vec_lbs<<<8 * ctx->num_sms, 256, 0, st>>>(I, ctx->paths[path].d, ctx->d_segmeta, ctx->d_dmin32, ctx->d_veclb, LS);
      LAUNCHED(ctx);
   }
   // small clusters on a symmetric table: 2 - 4 clusters per warp (k_vec_uw_packed)
   const int pslot = (I.symmetric && ctx->use_vec_pack) ? (I.mp <= 8 ? 8 : (I.mp <= 10 ? 10 : (I.mp <= 16 ? 16 : 0))) : 0;
   if (pslot) {
      const int pk = 32 / pslot, gxk = ((I.C + pk - 1) / pk + wpc - 1) / wpc;
      const size_t smk = (size_t)pk * slot * wpc;
      int gyk = 1;
#define VNS_VEC_PACKED(TT, SL, TAB)                                                                                   \
   gyk = layer_lanes(k_vec_uw_packed<TT, SL>, smk, gxk),                                                                 \
   k_vec_uw_packed<TT, SL><<<dim3(gxk, gyk), thr, smk, st>>>(I, ctx->paths[path].d, TAB, ctx->d_segmeta, ctx->d_vec, ctx->d_veclb, LS, fbi, ctx->d_dmin32)
      if (u16) { if (pslot == 8) VNS_VEC_PACKED(unsigned short, 8, ctx->d_D16); else if (pslot == 10) VNS_VEC_PACKED(unsigned short, 10, ctx->d_D16); else VNS_VEC_PACKED(unsigned short, 16, ctx->d_D16); }
      else { if (pslot == 8) VNS_VEC_PACKED(int, 8, ctx->d_D32); else if (pslot == 10) VNS_VEC_PACKED(int, 10, ctx->d_D32); else VNS_VEC_PACKED(int, 16, ctx->d_D32); }
#undef VNS_VEC_PACKED
   } else if (u16) k_vec_uw<unsigned short><<<dim3(gxu, gyu), thr, smem, st>>>(I, ctx->paths[path].d, ctx->d_D16, ctx->d_segmeta, ctx->d_vec, ctx->d_veclb, LS, fbi, ctx->d_dmin32);
   else k_vec_uw<int><<<dim3(gxu, gyu), thr, smem, st>>>(I, ctx->paths[path].d, ctx->d_D32, ctx->d_segmeta, ctx->d_vec, ctx->d_veclb, LS, fbi, ctx->d_dmin32);
   LAUNCHED(ctx);
   if (ctx->vec_skip_pairs) return VNSGPU_OK;      // tryToAdd candidates only: U and W are all they read
   if (u16) k_vec_pairs<unsigned short><<<dim3(gxp, gyp), thr, smem_p, st>>>(I, ctx->paths[path].d, ctx->d_D16, ctx->d_segmeta, ctx->d_vec, LS);
   else k_vec_pairs<int><<<dim3(gxp, gyp), thr, smem_p, st>>>(I, ctx->paths[path].d, ctx->d_D32, ctx->d_segmeta, ctx->d_vec, LS);
   LAUNCHED(ctx);
   return VNSGPU_OK;
}

// tables beyond the caches (synthetic scaled instances): latency form of k_eval_vec's block step, batch sorted by product row
static bool vec_lat(const vnsgpu_ctx *ctx) { return (size_t)ctx->I.N * ctx->I.N * sizeof(int) > ((size_t)48 << 20); }

// Is the batch evaluated in sorted order?  The endpoint-vector kernel on a cache-resident table mostly is not: every
// candidate is at most one block step there, so all the sort can do is group the candidates without one (a warp
// iteration skips the block product only if all its 32/G groups do).  With clusters of at most 2 G samples (one pass of
// the block product) that is worth less than the three sort passes and the scattered descriptor reads / result writes
// of a permuted batch (krob200, 4 Mi candidates: 11.5 -> 12.5 G evals/s without the sort); with larger clusters (up to
// 4 x 4 passes per product) the grouping still pays (vm1084: 5.8 G sorted, 5.2 G unsorted).
static bool eval_sorts(const vnsgpu_ctx *ctx, int64_t n, int useseg)
{
   const bool light = useseg == 2 && !vec_lat(ctx) && ctx->I.mp <= 2 * ctx->group_lanes;
   return group_kernel_ok(ctx) && ctx->sort_batches && n >= VNS_SORT_MIN && n < (1ll << 32) && !light && !ctx->vec_skip_pairs;
}

// fmt / d_leni: only with useseg == 2 and !eval_sorts(): d_cands are vnsgpu_pack_cand32 words read in place (fmt 2),
// lengths are written as int32 as well
static vnsgpu_status launch_eval(vnsgpu_ctx *ctx, int path, int64_t n, int64_t idx0, const vnsgpu_cand *d_cands, double budget,
                                 int skip, double *d_len, double *d_lb, uint8_t *d_flags, unsigned long long *d_first, cudaStream_t st,
                                 int slot, int64_t perm_off, int useseg, int fmt = 0, int *d_leni = nullptr)
{
   const InstDev &I = ctx->I;
   const PathDev &P = ctx->paths[path].d;
   if (group_kernel_ok(ctx)) {
      const unsigned *d_perm = nullptr;
      if (eval_sorts(ctx, n, useseg)) {
         const int nb = I.rows + 1;
         if (!ctx->d_hist) CK(cudaMalloc(&ctx->d_hist, sizeof(unsigned) * 2 * nb * VNS_PIPE_EVENTS));
         if ((int64_t)ctx->perm_cap < perm_off + n) {
            if (ctx->d_perm) { CK(cudaDeviceSynchronize()); CK(cudaFree(ctx->d_perm)); }
            ctx->d_perm = nullptr; ctx->perm_cap = 0;
            CK(cudaMalloc(&ctx->d_perm, sizeof(unsigned) * (size_t)(perm_off + n)));
            ctx->perm_cap = (size_t)(perm_off + n);
         }
         unsigned *hist = ctx->d_hist + (size_t)slot * 2 * nb, *cursor = hist + nb, *perm = ctx->d_perm + perm_off;
         PhaseTimer pt(ctx, VNSGPU_PHASE_SORT, st);
         const int sgrid = (int)std::min<int64_t>((n + 2047) / 2048, (int64_t)ctx->num_sms * 8);
         CK(cudaMemsetAsync(hist, 0, sizeof(unsigned) * nb, st));
         // endpoint vectors on a table beyond the caches: order the block candidates by product row (vec_sort_key)
         const int sortmode = (useseg == 2 && vec_lat(ctx)) ? 3 : useseg;
         k_sort_hist<<<sgrid, 256, sizeof(unsigned) * nb, st>>>(I, P, n, d_cands, nb, hist, sortmode);
         LAUNCHED(ctx);
         k_sort_scan<<<1, 1024, 0, st>>>(nb, hist, cursor);
         LAUNCHED(ctx);
         k_sort_scatter<<<sgrid, 256, sizeof(unsigned) * 2 * nb, st>>>(I, P, n, d_cands, nb, cursor, perm, sortmode);
         LAUNCHED(ctx);
         d_perm = perm;
      }
      PhaseTimer pt(ctx, VNSGPU_PHASE_EVAL, st);
      // T is only known on the device; size the layer map for its bound (a tour visits every cluster at most once)
      const size_t smem = eval_smem_bytes(std::min(I.rows, I.C), I.C);
      const int G = ctx->group_lanes, NR = (I.mp + G - 1) / G <= 1 ? 1 : ((I.mp + G - 1) / G <= 2 ? 2 : 4);
      const int64_t per_cta = (VNS_EVAL_THREADS / 32) * (32 / G);
      const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((n + per_cta - 1) / per_cta, (int64_t)ctx->num_sms * 8));
      SegCache seg{};
      if (useseg) {
         const int R = I.rows + 3;
         seg.cum = ctx->d_segmeta; seg.pd = ctx->d_segmeta + R;
         seg.rowoff = reinterpret_cast<const long long *>(ctx->d_segmeta + 2 * R);
      }
#define VNS_LAUNCH_EVAL(V, GG, RR, SS, TAB)                                                                          \
   do {                                                                                                              \
      if (smem > 48 * 1024)                                                                                          \
         CK(cudaFuncSetAttribute(k_eval_groups<V, GG, RR, SS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
      k_eval_groups<V, GG, RR, SS><<<grid, VNS_EVAL_THREADS, smem, st>>>(I, TAB, P, seg, n, idx0, d_cands, d_perm, budget, skip, \
                                                                          d_len, d_lb, d_flags, d_first);              \
   } while (0)
#define VNS_LAUNCH_EVAL_V(V, SS, TAB)                                                                                \
   do {                                                                                                              \
      if (G == 8 && NR == 1) VNS_LAUNCH_EVAL(V, 8, 1, SS, TAB);                                                       \
      else if (G == 8 && NR == 2) VNS_LAUNCH_EVAL(V, 8, 2, SS, TAB);                                                  \
      else if (G == 8) VNS_LAUNCH_EVAL(V, 8, 4, SS, TAB);                                                             \
      else if (G == 16 && NR == 1) VNS_LAUNCH_EVAL(V, 16, 1, SS, TAB);                                                \
      else if (G == 16) VNS_LAUNCH_EVAL(V, 16, 2, SS, TAB);                                                           \
      else VNS_LAUNCH_EVAL(V, 32, 1, SS, TAB);                                                                        \
   } while (0)
      if (useseg == 2) {
         VecCache vc{ctx->d_vec, ctx->d_veclb, I.C + 2};
#define VNS_LAUNCH_VEC_L(GG, RR, LL)                                                                                 \
   do {                                                                                                              \
      if (smem > 48 * 1024) CK(cudaFuncSetAttribute(k_eval_vec<GG, RR, LL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
      k_eval_vec<GG, RR, LL><<<grid, VNS_EVAL_THREADS, smem, st>>>(I, ctx->d_D32, P, seg, vc, n, idx0, d_cands, d_perm, budget, skip, d_len, \
                                                                   d_lb, d_flags, d_first, fmt, d_leni);             \
   } while (0)
#define VNS_LAUNCH_VEC(GG, RR)                                                                                       \
   do {                                                                                                              \
      if (lat) VNS_LAUNCH_VEC_L(GG, RR, true);                                                                        \
      else VNS_LAUNCH_VEC_L(GG, RR, false);                                                                           \
   } while (0)
         // tables beyond the caches: the same kernel with the NEXT candidate's block and vectors prefetched into L1 - or,
         // with clusters of at most 16 samples, the staged form (every lane copies its candidate's block and vectors into
         // shared memory with cp.async: k_eval_vec_st)
         const bool lat = vec_lat(ctx);
         if (lat && I.mp <= 16 && ctx->use_vec_staged) {
            const size_t meta_b = (smem + 15) & ~(size_t)15, per_warp = sizeof(int) * 32 * (size_t)vecst_slot_ints(I.mp);
            const int wpc = (int)std::min<size_t>(8, ((size_t)ctx->max_smem - 2048 - meta_b) / per_warp);
            if (wpc >= 2) {
               const size_t sm_st = meta_b + per_warp * wpc;
               const int gst = (int)std::max<int64_t>(1, std::min<int64_t>((n + 32 * wpc - 1) / (32 * wpc), ctx->num_sms));
#define VNS_LAUNCH_ST(GG, RR)                                                                                        \
   do {                                                                                                              \
      CK(cudaFuncSetAttribute(k_eval_vec_st<GG, RR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_st));        \
      k_eval_vec_st<GG, RR><<<gst, 32 * wpc, sm_st, st>>>(I, ctx->d_D32, P, seg, vc, n, idx0, d_cands, d_perm, budget, skip, d_len, d_lb, \
                                                          d_flags, d_first, fmt, d_leni, (int)(meta_b / 4));          \
   } while (0)
               if (G == 8 && NR == 1) VNS_LAUNCH_ST(8, 1);
               else if (G == 8) VNS_LAUNCH_ST(8, 2);
               else VNS_LAUNCH_ST(16, 1);
#undef VNS_LAUNCH_ST
               LAUNCHED(ctx);
               return VNSGPU_OK;
            }
         }
         if (G == 8 && NR == 1) VNS_LAUNCH_VEC(8, 1);
         else if (G == 8 && NR == 2) VNS_LAUNCH_VEC(8, 2);
         else if (G == 8) VNS_LAUNCH_VEC(8, 4);
         else if (G == 16 && NR == 1) VNS_LAUNCH_VEC(16, 1);
         else if (G == 16) VNS_LAUNCH_VEC(16, 2);
         else VNS_LAUNCH_VEC(32, 1);
#undef VNS_LAUNCH_VEC_L
#undef VNS_LAUNCH_VEC
      } else if (ctx->use_int && ctx->d_D32) {
         if (useseg) VNS_LAUNCH_EVAL_V(int, true, ctx->d_D32);
         else VNS_LAUNCH_EVAL_V(int, false, ctx->d_D32);
      } else VNS_LAUNCH_EVAL_V(double, false, ctx->d_D);
#undef VNS_LAUNCH_EVAL_V
#undef VNS_LAUNCH_EVAL
   } else {
      const size_t smem = sizeof(double) * 2 * I.mp * (VNS_EVAL_THREADS / 32);
      k_eval_batch<<<eval_grid(ctx, n), VNS_EVAL_THREADS, smem, st>>>(I, P, n, idx0, d_cands, budget, skip, d_len, d_lb, d_flags, d_first);
   }
   LAUNCHED(ctx);
   return VNSGPU_OK;
}

} // extern "C"

bool path_products_available(const vnsgpu_ctx *ctx)
{
   return ctx->use_seg && vnsgpu_int_path_active(ctx) && ctx->seg_ok && ctx->I.mp <= 32 && ctx->I.N < 65536;
}

// targets (and T, F[0], B[S]) of `path` are in place: products of this state, then all rows from them.  A caller that
// evaluates a batch on the same state next sets ctx->seg_fresh_path = path right before that call: prepare_eval then
// keeps these products (the flag lives for that one call)
vnsgpu_status path_rows_from_products(vnsgpu_ctx *ctx, int path, cudaStream_t st, bool init_ends)
{
   if (!path_products_available(ctx)) return fail(ctx, VNSGPU_E_STATE, "segment products not available for this table");
   if (init_ends) {
      k_rows_ends<<<1, 64, 0, st>>>(ctx->I, ctx->paths[path].d);
      LAUNCHED(ctx);
   }
   vnsgpu_status s = prepare_seg(ctx, path, st);
   if (s) return s;
   const int jobs = 2 * (ctx->I.rows + 2);
   k_rows_from_seg<<<std::min((jobs + 7) / 8, 8 * ctx->num_sms), 256, 0, st>>>(ctx->I, ctx->paths[path].d, ctx->d_D32, ctx->d_segmeta);
   LAUNCHED(ctx);
   k_rows_check<<<1, 1, 0, st>>>(ctx->paths[path].d);
   LAUNCHED(ctx);
   ctx->paths[path].hdr_ok = false;
   return VNSGPU_OK;
}

bool eval_dev_accepts_packed32(const vnsgpu_ctx *ctx, int64_t n)
{
   const int mode = eval_mode(ctx, n);
   return mode == 2 && !eval_sorts(ctx, n, mode) && ctx->I.C < 16384 && group_kernel_ok(ctx);
}

vnsgpu_status eval_dev_first(vnsgpu_ctx *ctx, int path, int64_t n, const void *d_cands, int format, double budget, int skip, long long *d_first,
                             cudaStream_t st)
{
   CK(cudaMemsetAsync(d_first, 0xFF, sizeof(long long), st));   // -1 = none
   const int useseg = eval_mode(ctx, n);
   vnsgpu_status s = prepare_eval(ctx, path, useseg, st);
   if (s) return s;
   return launch_eval(ctx, path, n, 0, reinterpret_cast<const vnsgpu_cand *>(d_cands), budget, skip, nullptr, nullptr, nullptr,
                      (unsigned long long *)d_first, st, 0, 0, useseg, format == 2 ? 2 : 0, nullptr);
}

extern "C" {

int vnsgpu_eval_mode(const vnsgpu_ctx *ctx, int64_t n)
{
   const int mode = eval_mode(ctx, n);
   return mode | (eval_sorts(ctx, n, mode) ? 16 : 0) | (vec_lat(ctx) ? 32 : 0) | (ctx->group_lanes << 8);
}

vnsgpu_status vnsgpu_eval_batch_dev(vnsgpu_ctx *ctx, int path, int64_t n, const vnsgpu_cand *d_cands, double budget, int skip,
                                    double *d_len, double *d_lb, uint8_t *d_flags, long long *d_first_accept, void *stream)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   if (n <= 0) return VNSGPU_OK;
   cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
   if (d_first_accept) CK(cudaMemsetAsync(d_first_accept, 0xFF, sizeof(long long), st));   // -1 = none
   const int useseg = eval_mode_hinted(ctx, n);
   vnsgpu_status s = prepare_eval(ctx, path, useseg, st);
   if (!s) s = launch_eval(ctx, path, n, 0, d_cands, budget, skip, d_len, d_lb, d_flags, (unsigned long long *)d_first_accept, st, 0, 0, useseg);
   ctx->vec_skip_pairs = false;
   return s;
}

// One rank's shard of a batch that several GPUs evaluate: the key of the first accepted candidate comes
// out as index_base + local index, VNSGPU_NO_KEY when the shard has none - the ranks all-reduce(MIN) the 8 bytes as they
// are (no conversion kernels between the evaluation and the collective).
vnsgpu_status vnsgpu_eval_batch_dev_shard(vnsgpu_ctx *ctx, int path, int64_t n, const vnsgpu_cand *d_cands, double budget, int skip,
                                          double *d_len, double *d_lb, uint8_t *d_flags, long long *d_first_key, int64_t index_base,
                                          void *stream)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   if (!d_first_key || index_base < 0) return VNSGPU_E_RANGE;
   cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
   CK(cudaMemsetAsync(d_first_key, 0x7F, sizeof(long long), st));     // VNSGPU_NO_KEY: above every index, positive as int64
   if (n <= 0) return VNSGPU_OK;
   const int useseg = eval_mode(ctx, n);
   {
      vnsgpu_status s = prepare_eval(ctx, path, useseg, st);
      if (s) return s;
   }
   return launch_eval(ctx, path, n, index_base, d_cands, budget, skip, d_len, d_lb, d_flags, (unsigned long long *)d_first_key, st, 0, 0,
                      useseg);
}

// Host buffers.  The batch is cut into chunks that flow through three streams - H2D of chunk c+1, the
// kernel on chunk c and D2H of chunk c-1 overlap - so the call runs at PCIe speed when the caller's
// buffers are pinned (pageable buffers work, the copies then serialise inside the driver).
__global__ void __launch_bounds__(256) k_unpack_cands(long long n, const unsigned long long *__restrict__ src, vnsgpu_cand *__restrict__ dst)
{
   for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (long long)gridDim.x * blockDim.x) {
      const unsigned long long w = src[q];
      int4 c;
      c.x = (int)(w & 7); c.y = (int)((w >> 3) & 0xfffff); c.z = (int)((w >> 23) & 0xfffff); c.w = (int)(w >> 43) - 1;
      reinterpret_cast<int4 *>(dst)[q] = c;
   }
}

static bool host_pinned(const void *p)
{
   if (!p) return true;
   cudaPointerAttributes at;
   if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
   return at.type == cudaMemoryTypeHost;
}

// 4-byte descriptors (vnsgpu_pack_cand32): op 3 bits, a 14 bits, b 14 bits, bit 31 = the gain cluster is a
__global__ void __launch_bounds__(256) k_unpack_cands32(long long n, const unsigned *__restrict__ src, vnsgpu_cand *__restrict__ dst)
{
   for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (long long)gridDim.x * blockDim.x) {
      const unsigned w = src[q];
      int4 c;
      c.x = (int)(w & 7); c.y = (int)((w >> 3) & 0x3fff); c.z = (int)((w >> 17) & 0x3fff); c.w = (w >> 31) ? c.y : -1;
      reinterpret_cast<int4 *>(dst)[q] = c;
   }
}

// lengths of an all-integer table as int32 (exact: every length is an integer below 2^31); NaN (bad indices) -> INT_MIN
__global__ void __launch_bounds__(256) k_len_to_i32(long long n, const double *__restrict__ len, int *__restrict__ out)
{
   for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (long long)gridDim.x * blockDim.x) {
      const double l = len[q];
      out[q] = l == l ? (int)l : INT_MIN;
   }
}

// result words of vnsgpu_eval_batch_packed32r: length | flags << 29 (lengths of the exact int32 path are below 2^29), all
// ones for bad indices
__global__ void __launch_bounds__(256) k_len_to_word(long long n, const double *__restrict__ len, const uint8_t *__restrict__ flags,
                                                     unsigned *__restrict__ out)
{
   for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (long long)gridDim.x * blockDim.x) {
      const double l = len[q];
      const unsigned f = flags[q];
      out[q] = (l == l && !(f & VNSGPU_F_BAD)) ? ((unsigned)(int)l | (f << 29)) : 0xffffffffu;
   }
}

// src: vnsgpu_cand[n] (packed == 0), uint64_t[n] in the vnsgpu_pack_cand format (packed == 1) or uint32_t[n] in the
// vnsgpu_pack_cand32 format (packed == 2); len_i32 (integer tables only): lengths as int32 instead of len
static vnsgpu_status eval_batch_host(vnsgpu_ctx *ctx, int path, int64_t n, const void *src, int packed, double budget, int skip,
                                     double *len, double *lb, uint8_t *flags, int64_t *first_accept, int32_t *len_i32 = nullptr,
                                     bool resw = false)
{
   // resw: len_i32 receives result words (length | flags << 29) and `flags` is not copied out
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   if (first_accept) *first_accept = -1;
   if (n <= 0) return VNSGPU_OK;
   const size_t o_len = al16(sizeof(vnsgpu_cand) * n), o_lb = o_len + al16(sizeof(double) * n);
   const size_t o_fl = o_lb + al16(sizeof(double) * n);
   const size_t o_pk = o_fl + al16(n);
   const size_t o_li = o_pk + (packed ? al16(sizeof(unsigned long long) * n) : 0);
   vnsgpu_status s = need_scratch(ctx, o_li + (len_i32 ? al16(sizeof(int) * n) : 0));
   if (s) return s;
   char *b = (char *)ctx->d_scratch;
   unsigned long long *d_pk = (unsigned long long *)(b + o_pk);
   unsigned *d_pk32 = (unsigned *)(b + o_pk);
   int *d_li = (int *)(b + o_li);
   vnsgpu_cand *d_c = (vnsgpu_cand *)b;
   double *d_len = (double *)(b + o_len), *d_lb = (double *)(b + o_lb);
   uint8_t *d_fl = (uint8_t *)(b + o_fl);
   if (!ctx->s_in) {
      CK(cudaStreamCreateWithFlags(&ctx->s_in, cudaStreamNonBlocking));
      CK(cudaStreamCreateWithFlags(&ctx->s_out, cudaStreamNonBlocking));
      for (int k = 0; k < VNS_PIPE_EVENTS; ++k) {
         CK(cudaEventCreateWithFlags(&ctx->ev_in[k], cudaEventDisableTiming));
         CK(cudaEventCreateWithFlags(&ctx->ev_k[k], cudaEventDisableTiming));
      }
      CK(cudaEventCreateWithFlags(&ctx->ev_start, cudaEventDisableTiming));
      CK(cudaEventCreateWithFlags(&ctx->ev_out, cudaEventDisableTiming));
      if (const char *e = getenv("VNSGPU_HOST_GRAPH")) ctx->use_graph = atoi(e) != 0;
   }
   // Chunk schedule.  The kernel is the slower stage (PCIe moves a candidate's 16 B in and 9-17 B out faster
   // than it is evaluated), and it is most efficient on large chunks (one sort + launch ramp per chunk): a
   // small first chunk starts the kernel early, a small last chunk shortens the exposed D2H tail, the
   // middle is cut large.  VNSGPU_PIPE_CHUNKS=k forces k equal chunks (experiments).
   std::vector<int64_t> cut;   // chunk sizes
   if (const char *e = getenv("VNSGPU_PIPE_CHUNKS")) {
      const int k = std::min(VNS_PIPE_EVENTS, std::max(1, atoi(e)));
      const int64_t c = (((n + k - 1) / k) + 255) & ~(int64_t)255;
      for (int64_t lo = 0; lo < n; lo += c) cut.push_back(std::min(c, n - lo));
   } else if (n < 262144) {
      cut.push_back(n);
   } else {
      static const int w[] = {1, 4, 4, 4, 2, 1};
      const int64_t unit = ((n / 16) + 255) & ~(int64_t)255;
      int64_t lo = 0;
      for (int k = 0; k < 6 && lo < n; ++k) {
         const int64_t c = k == 5 ? n - lo : std::min<int64_t>(unit * w[k], n - lo);
         cut.push_back(c);
         lo += c;
      }
   }
   cudaStream_t sk = ctx->stream;
   const int useseg = eval_mode(ctx, n);
   if (!ctx->h_first) CK(cudaMallocHost(&ctx->h_first, sizeof(long long)));

   // The call is ~100 stream operations (6 chunks x copies, events, 8 kernels).  A caller that repeats it with the
   // same buffers and batch size - a solver evaluating one neighbourhood batch per search state - gets it as ONE
   // CUDA graph launch from the third identical call on (the second one records it): the key is every pointer and
   // size the operations bake in.  The path's state (targets, T, F / B) is read by the kernels on the device, so the
   // graph stays valid while the search moves on.  Pageable host buffers and VNSGPU_HOST_GRAPH=0 keep the eager form.
   HostGraphKey key;
   memset(&key, 0, sizeof(key));
   key.I = ctx->I; key.P = ctx->paths[path].d;
   key.src = src; key.len = len; key.lb = lb; key.flags = flags; key.leni = len_i32;
   key.scratch = ctx->d_scratch; key.D32 = ctx->d_D32; key.D16 = (ctx->d_D16 && ctx->use_u16) ? ctx->d_D16 : nullptr; key.vec = ctx->d_vec; key.veclb = ctx->d_veclb;
   key.segmeta = ctx->d_segmeta; key.hist = ctx->d_hist; key.perm = ctx->d_perm; key.sk = sk;
   key.n = n; key.budget = budget; key.skip = skip; key.packed = packed; key.mode = useseg;
   key.sort = ctx->sort_batches + 2 * (int)resw; key.use_int = ctx->use_int; key.G = ctx->group_lanes; key.perm_cap = ctx->perm_cap;
   auto finish = [&]() -> vnsgpu_status {
      CK(cudaStreamSynchronize(sk));
      if (first_accept) *first_accept = (int64_t)*ctx->h_first;
      return VNSGPU_OK;
   };
   if (ctx->hg_exec && memcmp(&key, &ctx->hg_key, sizeof(key)) == 0) {
      CK(cudaGraphLaunch(ctx->hg_exec, sk));
      ctx->launches += ctx->hg_launches;
      return finish();
   }
   if (ctx->hg_exec) { cudaGraphExecDestroy(ctx->hg_exec); ctx->hg_exec = nullptr; }
   bool capture = ctx->use_graph && ctx->hg_seen && memcmp(&key, &ctx->hg_last, sizeof(key)) == 0 && n >= 262144 &&
                  host_pinned(src) && host_pinned(len) && host_pinned(lb) && host_pinned(flags) && host_pinned(len_i32);
   ctx->hg_last = key; ctx->hg_seen = true;
   if (capture && cudaStreamBeginCapture(sk, cudaStreamCaptureModeRelaxed) != cudaSuccess) { cudaGetLastError(); capture = false; }
   const int64_t launches0 = ctx->launches;

   auto issue = [&]() -> vnsgpu_status {
      vnsgpu_status s = prepare_eval(ctx, path, useseg, sk);
      if (s) return s;
      CK(cudaMemsetAsync(ctx->d_first, 0xFF, sizeof(long long), sk));
      CK(cudaEventRecord(ctx->ev_start, sk));          // order behind whatever the caller queued before
      CK(cudaStreamWaitEvent(ctx->s_in, ctx->ev_start, 0));
      CK(cudaStreamWaitEvent(ctx->s_out, ctx->ev_start, 0));
      int64_t lo = 0;
      for (int c = 0; c < (int)cut.size(); lo += cut[c], ++c) {
         const int64_t cn = cut[c];
         if (packed == 2)
            CK(cudaMemcpyAsync(d_pk32 + lo, (const unsigned *)src + lo, sizeof(unsigned) * cn, cudaMemcpyHostToDevice, ctx->s_in));
         else if (packed)
            CK(cudaMemcpyAsync(d_pk + lo, (const unsigned long long *)src + lo, sizeof(unsigned long long) * cn, cudaMemcpyHostToDevice, ctx->s_in));
         else
            CK(cudaMemcpyAsync(d_c + lo, (const vnsgpu_cand *)src + lo, sizeof(vnsgpu_cand) * cn, cudaMemcpyHostToDevice, ctx->s_in));
         CK(cudaEventRecord(ctx->ev_in[c], ctx->s_in));
         CK(cudaStreamWaitEvent(sk, ctx->ev_in[c], 0));
         // endpoint-vector kernel on an unsorted chunk: it reads the 4-byte descriptors in place and writes int32 lengths
         // itself (two kernels and 36 B / candidate of HBM traffic less per chunk)
         const bool direct = packed == 2 && useseg == 2 && !eval_sorts(ctx, cn, useseg);
         if (packed && !direct) {
            const int ug = (int)std::min<int64_t>((cn + 255) / 256, (int64_t)ctx->num_sms * 8);
            if (packed == 2) k_unpack_cands32<<<ug, 256, 0, sk>>>(cn, d_pk32 + lo, d_c + lo);
            else k_unpack_cands<<<ug, 256, 0, sk>>>(cn, d_pk + lo, d_c + lo);
            LAUNCHED(ctx);
         }
         if (direct)
            s = launch_eval(ctx, path, cn, lo, reinterpret_cast<const vnsgpu_cand *>(d_pk32 + lo), budget, skip, len ? d_len + lo : nullptr,
                            lb ? d_lb + lo : nullptr, flags ? d_fl + lo : nullptr, (unsigned long long *)ctx->d_first, sk, c, lo, useseg,
                            resw ? 6 : 2, len_i32 ? d_li + lo : nullptr);
         else
            s = launch_eval(ctx, path, cn, lo, d_c + lo, budget, skip, (len || len_i32) ? d_len + lo : nullptr, lb ? d_lb + lo : nullptr,
                            (flags || resw) ? d_fl + lo : nullptr, (unsigned long long *)ctx->d_first, sk, c, lo, useseg);
         if (s) return s;
         if (len_i32 && !direct) {
            const int cg = (int)std::min<int64_t>((cn + 255) / 256, (int64_t)ctx->num_sms * 8);
            if (resw) k_len_to_word<<<cg, 256, 0, sk>>>(cn, d_len + lo, d_fl + lo, (unsigned *)d_li + lo);
            else k_len_to_i32<<<cg, 256, 0, sk>>>(cn, d_len + lo, d_li + lo);
            LAUNCHED(ctx);
         }
         CK(cudaEventRecord(ctx->ev_k[c], sk));
         CK(cudaStreamWaitEvent(ctx->s_out, ctx->ev_k[c], 0));
         if (len_i32) CK(cudaMemcpyAsync(len_i32 + lo, d_li + lo, sizeof(int) * cn, cudaMemcpyDeviceToHost, ctx->s_out));
         if (len) CK(cudaMemcpyAsync(len + lo, d_len + lo, sizeof(double) * cn, cudaMemcpyDeviceToHost, ctx->s_out));
         if (lb) CK(cudaMemcpyAsync(lb + lo, d_lb + lo, sizeof(double) * cn, cudaMemcpyDeviceToHost, ctx->s_out));
         if (flags) CK(cudaMemcpyAsync(flags + lo, d_fl + lo, cn, cudaMemcpyDeviceToHost, ctx->s_out));
      }
      CK(cudaMemcpyAsync(ctx->h_first, ctx->d_first, sizeof(long long), cudaMemcpyDeviceToHost, sk));
      CK(cudaEventRecord(ctx->ev_out, ctx->s_out));         // join the D2H stream (the H2D stream is joined by ev_in)
      CK(cudaStreamWaitEvent(sk, ctx->ev_out, 0));
      return VNSGPU_OK;
   };
   s = issue();
   if (capture) {
      cudaGraph_t g = nullptr;
      const cudaError_t e = cudaStreamEndCapture(sk, &g);
      if (s || e != cudaSuccess || !g) {
         cudaGetLastError();
         if (g) cudaGraphDestroy(g);
         if (!s) { ctx->err = std::string("cudaStreamEndCapture: ") + cudaGetErrorString(e); s = VNSGPU_E_CUDA; }
         return s;
      }
      const cudaError_t ei = cudaGraphInstantiate(&ctx->hg_exec, g, 0);
      cudaGraphDestroy(g);
      if (ei != cudaSuccess) { ctx->hg_exec = nullptr; ctx->err = std::string("cudaGraphInstantiate: ") + cudaGetErrorString(ei); return VNSGPU_E_CUDA; }
      ctx->hg_key = key;
      ctx->hg_launches = ctx->launches - launches0;
      CK(cudaGraphLaunch(ctx->hg_exec, sk));
   } else if (s)
      return s;
   return finish();
}

vnsgpu_status vnsgpu_eval_batch(vnsgpu_ctx *ctx, int path, int64_t n, const vnsgpu_cand *cands, double budget, int skip, double *len,
                                double *lb, uint8_t *flags, int64_t *first_accept)
{
   return eval_batch_host(ctx, path, n, cands, 0, budget, skip, len, lb, flags, first_accept);
}

vnsgpu_status vnsgpu_eval_batch_packed(vnsgpu_ctx *ctx, int path, int64_t n, const uint64_t *packed, double budget, int skip, double *len,
                                       double *lb, uint8_t *flags, int64_t *first_accept)
{
   return eval_batch_host(ctx, path, n, packed, 1, budget, skip, len, lb, flags, first_accept);
}

vnsgpu_status vnsgpu_eval_batch_packed32(vnsgpu_ctx *ctx, int path, int64_t n, const uint32_t *packed, double budget, int skip,
                                         int32_t *len, uint8_t *flags, int64_t *first_accept)
{
   if (!vnsgpu_int_path_active(ctx)) {
      ctx->err = "vnsgpu_eval_batch_packed32: int32 lengths need an all-integer table (exact int32 path not active)";
      return VNSGPU_E_STATE;
   }
   return eval_batch_host(ctx, path, n, packed, 2, budget, skip, nullptr, nullptr, flags, first_accept, len);
}

vnsgpu_status vnsgpu_eval_batch_packed32r(vnsgpu_ctx *ctx, int path, int64_t n, const uint32_t *packed, double budget, int skip,
                                          uint32_t *result, int64_t *first_accept)
{
   if (!vnsgpu_int_path_active(ctx)) {
      ctx->err = "vnsgpu_eval_batch_packed32r: result words need an all-integer table (exact int32 path not active)";
      return VNSGPU_E_STATE;
   }
   return eval_batch_host(ctx, path, n, packed, 2, budget, skip, nullptr, nullptr, nullptr, first_accept, (int32_t *)result, true);
}

vnsgpu_status vnsgpu_cand_bytes(vnsgpu_ctx *ctx, int path, int64_t n, const vnsgpu_cand *cands, int64_t *bytes)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   *bytes = 0;
   if (n <= 0) return VNSGPU_OK;
   vnsgpu_status s = need_scratch(ctx, sizeof(vnsgpu_cand) * n + 16);
   if (s) return s;
   unsigned long long *d_tot = (unsigned long long *)ctx->d_scratch;
   vnsgpu_cand *d_c = (vnsgpu_cand *)((char *)ctx->d_scratch + 16);
   CK(cudaMemsetAsync(d_tot, 0, 8, ctx->stream));
   CK(cudaMemcpyAsync(d_c, cands, sizeof(vnsgpu_cand) * n, cudaMemcpyHostToDevice, ctx->stream));
   const int grid = (int)std::min<int64_t>((n + 255) / 256, 1184);
   k_cand_bytes<<<grid, 256, 0, ctx->stream>>>(ctx->I, ctx->paths[path].d, n, d_c, d_tot, vnsgpu_int_path_active(ctx) ? 4 : 8);
   LAUNCHED(ctx);
   unsigned long long tot = 0;
   CK(cudaMemcpyAsync(&tot, d_tot, 8, cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   *bytes = (int64_t)tot;
   return VNSGPU_OK;
}

} // extern "C"
