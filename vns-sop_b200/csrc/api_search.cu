// C ABI: RNG replay + neighbourhood drivers
#include "ctx.hpp"
#include "path.cuh"
#include "search.cuh"
#include "eval.cuh"
#include "chain.cuh"
#include "kernels_search.cuh"
#include "chain2_host.hpp"

// insertRandom / exchangeRandom: the shared-memory chain (chain.cuh) when the state fits one SM's shared
// memory, else the global-memory chain of search.cuh
template <typename V>
static bool chain_fits(const vnsgpu_ctx *ctx, size_t *bytes)
{
   *bytes = chain_smem_bytes<V>(ctx->I.C, ctx->I.mp);
   return ctx->use_smem_chain && group_kernel_ok(ctx) && ctx->I.mp <= 32 && ctx->I.C + 2 <= VNS_CH_THREADS &&
          *bytes + sizeof(ChainCtl) + 1024 <= (size_t)ctx->max_smem;
}

template <typename V, int G, int NR>
static vnsgpu_status launch_chain(vnsgpu_ctx *ctx, const V *tab, int path, int kind, int num_changes, double budget, int skip, size_t smem)
{
   CK(cudaFuncSetAttribute(k_chain_ls<V, G, NR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
   k_chain_ls<V, G, NR><<<1, VNS_CH_THREADS, smem, ctx->stream>>>(ctx->I, tab, ctx->paths[path].d, kind, num_changes, budget, skip,
                                                                    ctx->d_rng, ctx->d_stats, ctx->d_prof);
   LAUNCHED(ctx);
   return VNSGPU_OK;
}

template <typename V>
static vnsgpu_status launch_chain_v(vnsgpu_ctx *ctx, const V *tab, int path, int kind, int num_changes, double budget, int skip, size_t smem)
{
   const int G = ctx->group_lanes, nr = (ctx->I.mp + G - 1) / G;
   if (G == 8 && nr <= 1) return launch_chain<V, 8, 1>(ctx, tab, path, kind, num_changes, budget, skip, smem);
   if (G == 8 && nr <= 2) return launch_chain<V, 8, 2>(ctx, tab, path, kind, num_changes, budget, skip, smem);
   if (G == 8) return launch_chain<V, 8, 4>(ctx, tab, path, kind, num_changes, budget, skip, smem);
   if (G == 16 && nr <= 1) return launch_chain<V, 16, 1>(ctx, tab, path, kind, num_changes, budget, skip, smem);
   if (G == 16) return launch_chain<V, 16, 2>(ctx, tab, path, kind, num_changes, budget, skip, smem);
   return launch_chain<V, 32, 1>(ctx, tab, path, kind, num_changes, budget, skip, smem);
}

static vnsgpu_status launch_local_search(vnsgpu_ctx *ctx, int path, int kind, int num_changes, double budget, int skip)
{
   size_t smem = 0;
   int cis = 0;
   if (c3_fits(ctx, &smem, &cis)) return c3_launch_ls(ctx, path, kind, num_changes, budget, skip, smem, cis);
   if (ctx->use_smem_chain && c2_fits(ctx, &smem, &cis)) return c2_launch_ls(ctx, path, kind, num_changes, budget, skip, smem, cis);
   if (vnsgpu_int_path_active(ctx) && chain_fits<int>(ctx, &smem))
      return launch_chain_v<int>(ctx, ctx->d_D32, path, kind, num_changes, budget, skip, smem);
   if (chain_fits<double>(ctx, &smem)) return launch_chain_v<double>(ctx, ctx->d_D, path, kind, num_changes, budget, skip, smem);
   k_local_search<<<1, VNS_CHAIN_THREADS, chain_smem(ctx), ctx->stream>>>(ctx->I, ctx->paths[path].d, kind, num_changes, budget, skip,
                                                                          ctx->d_rng, ctx->d_stats);
   LAUNCHED(ctx);
   return VNSGPU_OK;
}

template <typename V, int G, int NR>
static vnsgpu_status launch_chain_vns(vnsgpu_ctx *ctx, const V *tab, int path, int work, const VnsParamsDev &prm, double *d_ir, double *d_il,
                                      double *d_it, size_t smem)
{
   CK(cudaFuncSetAttribute(k_chain_vns_run<V, G, NR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
   k_chain_vns_run<V, G, NR><<<1, VNS_CH_THREADS, smem, ctx->stream>>>(ctx->I, tab, ctx->paths[path].d, ctx->paths[work].d, prm, ctx->d_rng,
                                                                         ctx->d_res, d_ir, d_il, d_it, ctx->d_seg, ctx->d_prof);
   LAUNCHED(ctx);
   return VNSGPU_OK;
}

template <typename V>
static vnsgpu_status launch_chain_vns_v(vnsgpu_ctx *ctx, const V *tab, int path, int work, const VnsParamsDev &prm, double *d_ir,
                                        double *d_il, double *d_it, size_t smem)
{
   const int G = ctx->group_lanes, nr = (ctx->I.mp + G - 1) / G;
   if (G == 8 && nr <= 1) return launch_chain_vns<V, 8, 1>(ctx, tab, path, work, prm, d_ir, d_il, d_it, smem);
   if (G == 8 && nr <= 2) return launch_chain_vns<V, 8, 2>(ctx, tab, path, work, prm, d_ir, d_il, d_it, smem);
   if (G == 8) return launch_chain_vns<V, 8, 4>(ctx, tab, path, work, prm, d_ir, d_il, d_it, smem);
   if (G == 16 && nr <= 1) return launch_chain_vns<V, 16, 1>(ctx, tab, path, work, prm, d_ir, d_il, d_it, smem);
   if (G == 16) return launch_chain_vns<V, 16, 2>(ctx, tab, path, work, prm, d_ir, d_il, d_it, smem);
   return launch_chain_vns<V, 32, 1>(ctx, tab, path, work, prm, d_ir, d_il, d_it, smem);
}

template <typename V, int G, int NR>
static vnsgpu_status launch_chain_multi(vnsgpu_ctx *ctx, const V *tab, int nruns, VnsRunSlot *d_slots, const VnsParamsDev &prm, size_t smem)
{
   CK(cudaFuncSetAttribute(k_chain_vns_multi<V, G, NR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
   k_chain_vns_multi<V, G, NR><<<nruns, VNS_CH_THREADS, smem, ctx->stream>>>(ctx->I, tab, d_slots, prm);
   LAUNCHED(ctx);
   return VNSGPU_OK;
}

template <typename V>
static vnsgpu_status launch_chain_multi_v(vnsgpu_ctx *ctx, const V *tab, int nruns, VnsRunSlot *d_slots, const VnsParamsDev &prm, size_t smem)
{
   const int G = ctx->group_lanes, nr = (ctx->I.mp + G - 1) / G;
   if (G == 8 && nr <= 1) return launch_chain_multi<V, 8, 1>(ctx, tab, nruns, d_slots, prm, smem);
   if (G == 8 && nr <= 2) return launch_chain_multi<V, 8, 2>(ctx, tab, nruns, d_slots, prm, smem);
   if (G == 8) return launch_chain_multi<V, 8, 4>(ctx, tab, nruns, d_slots, prm, smem);
   if (G == 16 && nr <= 1) return launch_chain_multi<V, 16, 1>(ctx, tab, nruns, d_slots, prm, smem);
   if (G == 16) return launch_chain_multi<V, 16, 2>(ctx, tab, nruns, d_slots, prm, smem);
   return launch_chain_multi<V, 32, 1>(ctx, tab, nruns, d_slots, prm, smem);
}

extern "C" {

vnsgpu_status vnsgpu_chain_cycles(vnsgpu_ctx *ctx, int64_t *cycles4, int reset)
{
   CK(cudaSetDevice(ctx->device));
   CK(cudaMemcpyAsync(cycles4, ctx->d_prof, 4 * sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
   if (reset) CK(cudaMemsetAsync(ctx->d_prof, 0, 4 * sizeof(long long), ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   return VNSGPU_OK;
}

void vnsgpu_rng_seed(vnsgpu_rng *rng, uint32_t seed) { rng_seed_host(*reinterpret_cast<RngState *>(rng), seed); }
int32_t vnsgpu_rng_next(vnsgpu_rng *rng) { return rng_next(*reinterpret_cast<RngState *>(rng)); }

vnsgpu_status vnsgpu_gen_proposals(vnsgpu_ctx *ctx, int path, int kind, int64_t n, vnsgpu_rng *rng, vnsgpu_cand *cands)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   if (kind != 1 && kind != 2) return fail(ctx, VNSGPU_E_RANGE, "kind must be 1 (insert) or 2 (exchange)");
   PathHdr h;
   vnsgpu_status s = read_hdr(ctx, path, &h);
   if (s) return s;
   const int T = h.hdr[HDR_T], nvns = h.hdr[HDR_NVNS];
   if (kind == 1 ? nvns < 3 : (nvns < 4 || T < 3)) return fail(ctx, VNSGPU_E_STATE, "neighbourhood is empty for this state");
   std::vector<int> vns(nvns);
   CK(cudaMemcpyAsync(vns.data(), ctx->paths[path].d.vns, sizeof(int) * nvns, cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   RngState &g = *reinterpret_cast<RngState *>(rng);
   const int gap = kind == 1 ? 1 : 2, tohi = kind == 1 ? T : T - 1;
   for (int64_t k = 0; k < n; ++k) {
      int from = rng_int(g, 0, nvns - 1), to = rng_int(g, 0, tohi);        // vns_sop_solver.cpp:507-508 / 723-724
      while (std::abs(from - to) <= gap) {                                 // :510-519 / :726-730
         from = rng_int(g, 0, nvns - 1);
         to = rng_int(g, 0, tohi);
      }
      vnsgpu_cand &c = cands[k];
      if (from < T) { c.op = kind == 1 ? VNSGPU_OP_MOVE : VNSGPU_OP_EXCHANGE; c.a = from; c.b = to; c.gain_cluster = -1; }
      else { c.op = kind == 1 ? VNSGPU_OP_ADD : VNSGPU_OP_REPLACE; c.a = vns[from]; c.b = to; c.gain_cluster = vns[from]; }
   }
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_local_search(vnsgpu_ctx *ctx, int path, int kind, int num_changes, double budget, int skip, vnsgpu_rng *rng,
                                  vnsgpu_ls_stats *stats)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   if (kind != 1 && kind != 2) return fail(ctx, VNSGPU_E_RANGE, "kind must be 1 (insert) or 2 (exchange)");
   CK(cudaMemcpyAsync(ctx->d_rng, rng, sizeof(RngState), cudaMemcpyHostToDevice, ctx->stream));
   vnsgpu_status ls = launch_local_search(ctx, path, kind, num_changes, budget, skip);
   if (ls) return ls;
   LsStats st;
   CK(cudaMemcpyAsync(rng, ctx->d_rng, sizeof(RngState), cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaMemcpyAsync(&st, ctx->d_stats, sizeof(st), cudaMemcpyDeviceToHost, ctx->stream));
   vnsgpu_status s = check_path_err(ctx, path);
   if (stats) {
      stats->proposals = st.proposals; stats->lb_skips = st.lb_skips; stats->evals = st.evals;
      stats->accepts = st.accepts; stats->rounds = st.rounds;
   }
   return s;
}

vnsgpu_status vnsgpu_shake(vnsgpu_ctx *ctx, int path, int kind, double budget, vnsgpu_rng *rng)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   if (kind != 1 && kind != 2) return fail(ctx, VNSGPU_E_RANGE, "kind must be 1 (pathInsert) or 2 (pathExchange)");
   CK(cudaMemcpyAsync(ctx->d_rng, rng, sizeof(RngState), cudaMemcpyHostToDevice, ctx->stream));
   k_shake<<<1, VNS_CHAIN_THREADS, chain_smem(ctx), ctx->stream>>>(ctx->I, ctx->paths[path].d, kind, budget, ctx->d_rng, ctx->d_seg);
   LAUNCHED(ctx);
   CK(cudaMemcpyAsync(rng, ctx->d_rng, sizeof(RngState), cudaMemcpyDeviceToHost, ctx->stream));
   return check_path_err(ctx, path);
}

vnsgpu_status vnsgpu_fit_to_budget(vnsgpu_ctx *ctx, int path, double budget)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   k_fit<<<1, VNS_CHAIN_THREADS, chain_smem(ctx), ctx->stream>>>(ctx->I, ctx->paths[path].d, budget);
   LAUNCHED(ctx);
   return check_path_err(ctx, path);
}

vnsgpu_status vnsgpu_greedy(vnsgpu_ctx *ctx, int path, double budget)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   k_greedy<<<1, VNS_CHAIN_THREADS, chain_smem(ctx), ctx->stream>>>(ctx->I, ctx->paths[path].d, budget);
   LAUNCHED(ctx);
   return check_path_err(ctx, path);
}

vnsgpu_status vnsgpu_vns_run(vnsgpu_ctx *ctx, int path, const vnsgpu_vns_params *p, vnsgpu_vns_result *result, double *impr_reward,
                             double *impr_length, double *impr_time_ms)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   const int cap = std::max(1, p->max_improvements);
   vnsgpu_status s = need_scratch(ctx, sizeof(double) * 3 * cap);
   if (s) return s;
   size_t smem2 = 0;
   int cis = 0;
   const bool c3 = c3_fits(ctx, &smem2, &cis);
   const bool c2 = !c3 && ctx->use_smem_chain && c2_fits(ctx, &smem2, &cis);
   int work = -1;
   if (!c2 && !c3) {
      s = vnsgpu_path_create(ctx, &work);   // actualVNSGOPPath (vns_sop_solver.cpp:195); the second chain keeps it in shared memory
      if (s) return s;
   }
   double *d_ir = (double *)ctx->d_scratch, *d_il = d_ir + cap, *d_it = d_il + cap;
   RngState g;
   rng_seed_host(g, p->seed);
   CK(cudaMemcpyAsync(ctx->d_rng, &g, sizeof(g), cudaMemcpyHostToDevice, ctx->stream));
   VnsParamsDev prm;
   prm.num_iterations = p->num_iterations; prm.num_iterations_unimproved = p->num_iterations_unimproved;
   prm.skip = p->lower_bound_skip; prm.greedy = p->greedy_initial; prm.budget = p->budget;
   prm.max_time_ms = p->max_time_ms; prm.max_improvements = cap;
   prm.target_reward = 0.0; prm.stop_flag = nullptr;
   if (c3)
      s = c3_launch_vns(ctx, path, prm, d_ir, d_il, d_it, smem2, cis);
   else if (c2)
      s = c2_launch_vns(ctx, path, prm, d_ir, d_il, d_it, smem2, cis);
   else {
      size_t smem = 0;
      const size_t extra = chain_smem(ctx) + sizeof(RngState) + 64;   // shake / fit / greedy scratch + statics
      if (vnsgpu_int_path_active(ctx) && chain_fits<int>(ctx, &smem) && smem + extra + sizeof(ChainCtl) + 1024 <= (size_t)ctx->max_smem)
         s = launch_chain_vns_v<int>(ctx, ctx->d_D32, path, work, prm, d_ir, d_il, d_it, smem + chain_smem(ctx));
      else if (chain_fits<double>(ctx, &smem) && smem + extra + sizeof(ChainCtl) + 1024 <= (size_t)ctx->max_smem)
         s = launch_chain_vns_v<double>(ctx, ctx->d_D, path, work, prm, d_ir, d_il, d_it, smem + chain_smem(ctx));
      else {
         k_vns_run<<<1, VNS_CHAIN_THREADS, chain_smem(ctx), ctx->stream>>>(ctx->I, ctx->paths[path].d, ctx->paths[work].d, prm, ctx->d_rng,
                                                                           ctx->d_res, d_ir, d_il, d_it, ctx->d_seg);
         LAUNCHED(ctx);
      }
   }
   if (s) { if (work >= 0) vnsgpu_path_destroy(ctx, work); return s; }
   ctx->paths[path].hdr_ok = false;
   VnsResultDev r;
   CK(cudaMemcpyAsync(&r, ctx->d_res, sizeof(r), cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   const int n = std::min(r.n_improvements, cap);
   if (impr_reward) CK(cudaMemcpy(impr_reward, d_ir, sizeof(double) * n, cudaMemcpyDeviceToHost));
   if (impr_length) CK(cudaMemcpy(impr_length, d_il, sizeof(double) * n, cudaMemcpyDeviceToHost));
   if (impr_time_ms) CK(cudaMemcpy(impr_time_ms, d_it, sizeof(double) * n, cudaMemcpyDeviceToHost));
   if (result) {
      result->num_iters = r.num_iters; result->num_iters_last_impr = r.num_iters_last_impr;
      result->n_improvements = r.n_improvements; result->status = r.status;
      result->reward = r.reward; result->length = r.length; result->time_ms = r.time_ms;
      result->time_last_impr_ms = r.time_last_impr_ms;
      result->proposals = r.proposals; result->evals = r.evals; result->accepts = r.accepts;
   }
   if (work >= 0) vnsgpu_path_destroy(ctx, work);
   if (r.status == 4) return fail(ctx, VNSGPU_E_INCONSISTENT, "inconsistent path length in improved tour");
   if (r.status) return fail(ctx, VNSGPU_E_STATE, "device-side error flag in vns_run");
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_vns_run_batch(vnsgpu_ctx *ctx, int path, int nruns, const uint32_t *seeds, const vnsgpu_vns_params *p,
                                   vnsgpu_vns_result *results, int *best_run, int *tour_paths)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   if (nruns < 1) return fail(ctx, VNSGPU_E_RANGE, "nruns < 1");
   size_t smem = 0;
   int cis = 0;
   const bool c3 = c3_fits(ctx, &smem, &cis, nruns > ctx->num_sms);
   const bool c2 = c3 || (ctx->use_smem_chain && c2_fits(ctx, &smem, &cis));   // c2: the slot layout both forms share
   const size_t extra = chain_smem(ctx) + sizeof(RngState) + 64;
   bool as_int = false;
   if (!c2) {
      as_int = vnsgpu_int_path_active(ctx) && chain_fits<int>(ctx, &smem) && smem + extra + sizeof(ChainCtl) + 1024 <= (size_t)ctx->max_smem;
      if (!as_int && !(chain_fits<double>(ctx, &smem) && smem + extra + sizeof(ChainCtl) + 1024 <= (size_t)ctx->max_smem))
         return fail(ctx, VNSGPU_E_TOO_LARGE, "instance too large for the shared-memory chain: use vnsgpu_vns_run per seed");
   }
   // first form: two paths per run (tourVNSGOPPath, actualVNSGOPPath); second form: the tour only.  Every tour a copy of `path`
   std::vector<int> tours(nruns, -1), acts(nruns, -1);
   const size_t slot_bytes = c2 ? c2_run_slot_bytes() : sizeof(VnsRunSlot);
   std::vector<char> slots(slot_bytes * (size_t)nruns);
   int *d_segs = nullptr;
   void *d_slots = nullptr;
   vnsgpu_status s = VNSGPU_OK;
   auto cuda_ok = [&](cudaError_t e, const char *what) {
      if (e == cudaSuccess || s) return;
      ctx->err = std::string(what) + ": " + cudaGetErrorString(e);
      s = VNSGPU_E_CUDA;
   };
   if (!c2) cuda_ok(cudaMalloc(&d_segs, sizeof(int) * 2 * ctx->I.rows * (size_t)nruns), "cudaMalloc(segs)");
   cuda_ok(cudaMalloc(&d_slots, slot_bytes * (size_t)nruns), "cudaMalloc(slots)");
   for (int r = 0; r < nruns && !s; ++r) {
      int a = -1, b = -1;
      s = vnsgpu_path_create(ctx, &a);
      tours[r] = a;
      if (!s && !c2) { s = vnsgpu_path_create(ctx, &b); acts[r] = b; }
      if (!s) s = vnsgpu_path_copy(ctx, a, path);
      if (s) break;
      RngState g;
      rng_seed_host(g, seeds[r]);
      if (c2)
         c2_run_slot_fill(slots.data() + slot_bytes * r, ctx->paths[a].d, g);
      else {
         VnsRunSlot &sl = *reinterpret_cast<VnsRunSlot *>(slots.data() + slot_bytes * r);
         sl.tour = ctx->paths[a].d; sl.act = ctx->paths[b].d; sl.rng = g;
         memset(&sl.res, 0, sizeof(VnsResultDev));
         sl.seg = d_segs + (size_t)r * 2 * ctx->I.rows;
      }
   }
   if (!s) {
      cuda_ok(cudaMemcpyAsync(d_slots, slots.data(), slots.size(), cudaMemcpyHostToDevice, ctx->stream), "cudaMemcpyAsync(slots)");
      VnsParamsDev prm;
      prm.num_iterations = p->num_iterations; prm.num_iterations_unimproved = p->num_iterations_unimproved;
      prm.skip = p->lower_bound_skip; prm.greedy = p->greedy_initial; prm.budget = p->budget;
      prm.max_time_ms = p->max_time_ms; prm.max_improvements = 0;
      prm.target_reward = ctx->target_reward; prm.stop_flag = nullptr;
      if (ctx->target_reward > 0) {
         cuda_ok(cudaMemsetAsync(ctx->d_stop, 0, sizeof(int), ctx->stream), "cudaMemsetAsync(stop)");
         prm.stop_flag = ctx->d_stop;
      }
      if (!s) {
         if (c3) s = c3_launch_multi(ctx, nruns, d_slots, prm, smem, cis);
         else if (c2) s = c2_launch_multi(ctx, nruns, d_slots, prm, smem, cis);
         else
            s = as_int ? launch_chain_multi_v<int>(ctx, ctx->d_D32, nruns, (VnsRunSlot *)d_slots, prm, smem + chain_smem(ctx))
                       : launch_chain_multi_v<double>(ctx, ctx->d_D, nruns, (VnsRunSlot *)d_slots, prm, smem + chain_smem(ctx));
      }
   }
   if (!s) {
      cuda_ok(cudaMemcpyAsync(slots.data(), d_slots, slots.size(), cudaMemcpyDeviceToHost, ctx->stream), "cudaMemcpyAsync(results)");
      cuda_ok(cudaStreamSynchronize(ctx->stream), "cudaStreamSynchronize");
   }
   if (!s) {
      auto result_of = [&](int r) -> const VnsResultDev & {
         return c2 ? c2_run_slot_result(slots.data() + slot_bytes * r) : reinterpret_cast<const VnsRunSlot *>(slots.data() + slot_bytes * r)->res;
      };
      int best = 0, bad = 0;
      for (int r = 0; r < nruns; ++r) {
         const VnsResultDev &q = result_of(r);
         if (q.status) bad = q.status;
         if (results) {
            vnsgpu_vns_result &o = results[r];
            o.num_iters = q.num_iters; o.num_iters_last_impr = q.num_iters_last_impr; o.n_improvements = q.n_improvements;
            o.status = q.status; o.reward = q.reward; o.length = q.length; o.time_ms = q.time_ms;
            o.time_last_impr_ms = q.time_last_impr_ms; o.proposals = q.proposals; o.evals = q.evals; o.accepts = q.accepts;
         }
         const VnsResultDev &bq = result_of(best);
         if (q.reward > bq.reward || (q.reward == bq.reward && q.length < bq.length)) best = r;
      }
      if (best_run) *best_run = best;
      s = vnsgpu_path_copy(ctx, path, tours[best]);      // the best tour replaces `path`
      if (!s) cuda_ok(cudaStreamSynchronize(ctx->stream), "cudaStreamSynchronize");
      if (!s && bad) s = fail(ctx, bad == 4 ? VNSGPU_E_INCONSISTENT : VNSGPU_E_STATE, "device-side error flag in vns_run_batch");
   }
   // every exit path releases the temporaries
   cudaStreamSynchronize(ctx->stream);
   const std::string keep = ctx->err;
   for (int r = 0; r < nruns; ++r) {
      if (acts[r] >= 0) vnsgpu_path_destroy(ctx, acts[r]);
      if (tours[r] >= 0 && (s || !tour_paths)) vnsgpu_path_destroy(ctx, tours[r]);
      if (tour_paths) tour_paths[r] = s ? -1 : tours[r];
   }
   cudaFree(d_segs); cudaFree(d_slots);
   if (s) ctx->err = keep;
   return s;
}

} // extern "C"
