// Full-neighbourhood modes on top of the batched evaluation: every candidate of a neighbourhood is
// evaluated by ONE vnsgpu_eval_batch_dev call (endpoint vectors + segment products on integer tables), a device
// reduction picks the move, a one-block kernel applies it.
//   * vnsgpu_greedy_batched: greedy_insertion (vns_sop_solver.cpp:1091-1155) - per round the (|VNS| - T) x (T + 1) grid
//     of tryToAdd, strict minimum of (d - L) / reward among d <= budget, first grid index on ties (:1119).  Same tour as
//     vnsgpu_greedy and as the reference (trajectory-exact: the reference's loop is already an enumeration).
//   * vnsgpu_best_move / vnsgpu_apply_move / vnsgpu_best_improvement: the opt-in, NOT trajectory-exact local search -
//     the whole insert (MOVE / ADD) or exchange (EXCHANGE / REPLACE) neighbourhood in the reference's index order
//     (from-major), the best improving candidate instead of the first drawn one.
#include "ctx.hpp"
#include "path.cuh"
#include "search.cuh"

struct BestRec {            // device result of the reduction
   long long idx;           // enumeration index of the chosen candidate, -1: none
   double a, b;             // greedy: (d - L) / reward, d;  best-improvement: added reward, length after the move
   int from, to;            // VNS-vector index that moves / target index it goes to
   int kind, pad;
};

struct BestKey {
   double a, b;
   long long idx;
};

// total orders (smaller = better).  MODE 0: a ascending, first index.  MODE 1: a (added reward) descending, b (length)
// ascending, first index
template <int MODE> __device__ __forceinline__ bool key_better(const BestKey &x, const BestKey &y)
{
   if (x.idx < 0) return false;
   if (y.idx < 0) return true;
   if (MODE == 0) return x.a < y.a || (x.a == y.a && x.idx < y.idx);
   return x.a > y.a || (x.a == y.a && (x.b < y.b || (x.b == y.b && x.idx < y.idx)));
}

// ---- enumeration of a neighbourhood (device descriptors, index order = the reference's loop order) ----------------------
// kind 0: greedy grid  g = (av - T) * (T + 1) + pos  -> tryToAdd(VNS[av], pos)                      (:1109-1127)
// kind 1: insert       k = from * (T + 1) + to        -> tryToMove(from, to) / tryToAdd(VNS[from], to)    (:507-519 pairs)
// kind 2: exchange     k = from * T + to              -> tryToExchange(from, to) / tryToReplace(VNS[from], to)  (:723-730)
// pairs the reference redraws (|from - to| <= gap) get op = -1: the evaluation flags them VNSGPU_F_BAD, never chosen
static __global__ void __launch_bounds__(256) k_nbh_cands(PathDev P, int kind, long long k0, long long n, vnsgpu_cand *__restrict__ out)
{
   const int T = P.hdr[HDR_T];
   const int ncol = kind == 2 ? T : T + 1;
   // this rank's shard: enumeration indices k0 .. k0 + n - 1
   for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (long long)gridDim.x * blockDim.x) {
      const long long k = k0 + q;
      const int row = (int)(k / ncol), to = (int)(k - (long long)row * ncol);
      vnsgpu_cand c;
      if (kind == 0) {
         const int cl = P.vns[T + row];
         c.op = VNSGPU_OP_ADD; c.a = cl; c.b = to; c.gain_cluster = cl;
      } else {
         const int from = row, gap = kind == 1 ? 1 : 2;
         if (abs(from - to) <= gap) { c.op = -1; c.a = 0; c.b = 0; c.gain_cluster = -1; }
         else if (from < T) { c.op = kind == 1 ? VNSGPU_OP_MOVE : VNSGPU_OP_EXCHANGE; c.a = from; c.b = to; c.gain_cluster = -1; }
         else {
            const int cl = P.vns[from];
            c.op = kind == 1 ? VNSGPU_OP_ADD : VNSGPU_OP_REPLACE; c.a = cl; c.b = to; c.gain_cluster = cl;
         }
      }
      out[q] = c;
   }
}

// ---- arg-best over the evaluated neighbourhood ---------------------------------------------------------------------------
// partial: one BestKey per block; the last block to finish (ticket) reduces the partials and writes *rec
template <int MODE>
static __global__ void __launch_bounds__(256) k_best_reduce(InstDev I, PathDev P, int kind, long long n, const vnsgpu_cand *__restrict__ cands,
                                                            const double *__restrict__ len, const uint8_t *__restrict__ flags, double budget,
                                                            BestKey *partial, unsigned *ticket, BestKey *key_out, long long k0)
{
   __shared__ BestKey sh[256];
   __shared__ bool last;
   const double L = P.hd[HD_LEN];
   BestKey best{0.0, 0.0, -1};
   for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x) {
      const double d = len[k];
      BestKey cur{0.0, d, k0 + k};      // global enumeration index
      bool ok;
      if (MODE == 0) {
         const vnsgpu_cand c = cands[k];
         cur.a = (d - L) / __ldg(I.reward + c.gain_cluster);                   // :1116-1117
         ok = d <= budget;                                                    // :1119 (NaN fails)
      } else {
         const uint8_t f = flags[k];
         ok = (f & VNSGPU_F_FEASIBLE) && !(f & VNSGPU_F_BAD);
         if (ok) {
            const vnsgpu_cand c = cands[k];
            double dR = 0.0;
            if (c.op == VNSGPU_OP_ADD) dR = __ldg(I.reward + c.gain_cluster);                                             // :581
            else if (c.op == VNSGPU_OP_REPLACE) dR = 0.0 + (__ldg(I.reward + c.gain_cluster) - __ldg(I.reward + P.tg[c.b]));   // :784
            cur.a = dR;
            // improving in the lexicographic objective (reward, then length): the reference's test (:619,:805-806) also
            // takes a shorter tour with LESS reward; a best-improvement descent must not
            ok = dR > 0 || (dR == 0 && (L - d) > VNS_MIN_CHANGE_EPS);
         }
      }
      if (ok && key_better<MODE>(cur, best)) best = cur;
   }
   sh[threadIdx.x] = best;
   __syncthreads();
   for (int s = 128; s > 0; s >>= 1) {
      if ((int)threadIdx.x < s && key_better<MODE>(sh[threadIdx.x + s], sh[threadIdx.x])) sh[threadIdx.x] = sh[threadIdx.x + s];
      __syncthreads();
   }
   if (threadIdx.x == 0) {
      partial[blockIdx.x] = sh[0];
      __threadfence();
      last = atomicAdd(ticket, 1u) == gridDim.x - 1;
   }
   __syncthreads();
   if (!last) return;
   __threadfence();
   best = BestKey{0.0, 0.0, -1};
   for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) {
      BestKey q;
      q.a = __ldcg(&partial[b].a); q.b = __ldcg(&partial[b].b); q.idx = __ldcg(&partial[b].idx);
      if (key_better<MODE>(q, best)) best = q;
   }
   sh[threadIdx.x] = best;
   __syncthreads();
   for (int s = 128; s > 0; s >>= 1) {
      if ((int)threadIdx.x < s && key_better<MODE>(sh[threadIdx.x + s], sh[threadIdx.x])) sh[threadIdx.x] = sh[threadIdx.x + s];
      __syncthreads();
   }
   if (threadIdx.x == 0) {
      *key_out = sh[0];
      *ticket = 0;
   }
}

// the best of nkeys keys (one per rank of a sharded neighbourhood, in rank order - the same array on every rank, so every
// rank picks the same move) as the record the apply kernels read
template <int MODE> static __global__ void k_best_pick(PathDev P, int kind, const BestKey *__restrict__ keys, int nkeys, BestRec *rec)
{
   const int T = P.hdr[HDR_T];
   BestKey best{0.0, 0.0, -1};
   for (int r = 0; r < nkeys; ++r)
      if (key_better<MODE>(keys[r], best)) best = keys[r];
   BestRec o;
   o.idx = best.idx; o.a = best.a; o.b = best.b; o.kind = kind; o.pad = 0; o.from = o.to = -1;
   if (o.idx >= 0) {
      const int ncol = kind == 2 ? T : T + 1;
      const int row = (int)(o.idx / ncol);
      o.to = (int)(o.idx - (long long)row * ncol);
      o.from = kind == 0 ? T + row : row;
   }
   *rec = o;
}

// ---- apply the chosen move (one block): the accept branches of insertRandom (:621-646) / exchangeRandom (:807-832), or
//      greedy_insertion's addPoint + VNS splice (:1133-1146) - all three are "VNS[from] goes to index `to`" -----------------
static __global__ void __launch_bounds__(VNS_CHAIN_THREADS) k_apply_move(InstDev I, PathDev P, const BestRec *rec, int from_arg, int to_arg, int kind_arg)
{
   int from = from_arg, to = to_arg, kind = kind_arg;
   if (rec) {
      if (rec->idx < 0) return;
      from = rec->from; to = rec->to; kind = rec->kind;
   }
   int T = P.hdr[HDR_T];
   const int nvns = P.hdr[HDR_NVNS];
   __syncthreads();
   if (kind == 2) cta_apply_exchange(I, P, T, from, to);
   else T = cta_apply_insert(I, P, T, nvns, from, to);
   __syncthreads();
   if (threadIdx.x == 0) { P.hdr[HDR_T] = T; P.hd[HD_REWARD] = path_reward(I, P.tg, T); }
}

// the same move as a splice of the VNS vector and the targets only (T, reward, the goal's zero row of B): the rows come
// from the segment products afterwards (path_rows_from_products) - all layers in parallel instead of the sequential re-scan
static __global__ void __launch_bounds__(VNS_CHAIN_THREADS) k_splice_move(InstDev I, PathDev P, const BestRec *rec, int from_arg, int to_arg, int kind_arg)
{
   __shared__ int s_c;
   int from = from_arg, to = to_arg, kind = kind_arg;
   if (rec) {
      if (rec->idx < 0) return;
      from = rec->from; to = rec->to; kind = rec->kind;
   }
   int T = P.hdr[HDR_T];
   if (threadIdx.x == 0) s_c = P.vns[from];
   __syncthreads();
   const int c = s_c;
   if (kind == 2) {
      if (threadIdx.x == 0) {
         const int ct = P.vns[to];
         P.vns[to] = c; P.vns[from] = ct;
         P.tg[to] = c;
         if (from < T) P.tg[from] = ct;
      }
   } else if (from < T) {
      if (to > from) {
         cta_shift(P.vns, from + 1, to - 1 - from, -1);
         cta_shift(P.tg, from + 1, to - 1 - from, -1);
         if (threadIdx.x == 0) { P.vns[to - 1] = c; P.tg[to - 1] = c; }
      } else {
         cta_shift(P.vns, to, from - to, +1);
         cta_shift(P.tg, to, from - to, +1);
         if (threadIdx.x == 0) { P.vns[to] = c; P.tg[to] = c; }
      }
   } else {
      cta_shift(P.vns, to, from - to, +1);
      cta_shift(P.tg, to, T - to, +1);
      if (threadIdx.x == 0) { P.vns[to] = c; P.tg[to] = c; }
      T += 1;
      const int mS = I.m[I.goal];
      for (int j = threadIdx.x; j < mS; j += blockDim.x) P.B[(size_t)(T + 1) * I.mp + j] = 0.0;   // the goal layer moved up by one
   }
   __syncthreads();
   if (threadIdx.x == 0) { P.hdr[HDR_T] = T; P.hd[HD_REWARD] = path_reward(I, P.tg, T); }
}

// ---- host side --------------------------------------------------------------------------------------------------------------
struct BestBufs {
   vnsgpu_cand *cands;
   double *len;
   uint8_t *flags;
   BestKey *partial;
   unsigned *ticket;
   BestRec *rec;
   BestKey *mykey, *keys;    // this rank's best, and every rank's (all-gather)
};

static vnsgpu_status best_bufs(vnsgpu_ctx *ctx, long long n, BestBufs *b)
{
   const size_t nb = 2048;
   const size_t o_len = al16(sizeof(vnsgpu_cand) * (size_t)n), o_fl = o_len + al16(sizeof(double) * (size_t)n),
                o_part = o_fl + al16((size_t)n), o_tick = o_part + al16(sizeof(BestKey) * nb), o_rec = o_tick + 16,
                o_my = o_rec + al16(sizeof(BestRec)), o_keys = o_my + al16(sizeof(BestKey)),
                total = o_keys + al16(sizeof(BestKey) * (size_t)std::max(1, ctx->comm_size));
   if (total > ctx->best_bytes) {
      if (ctx->d_best) { CK(cudaStreamSynchronize(ctx->stream)); CK(cudaFree(ctx->d_best)); }
      ctx->d_best = nullptr; ctx->best_bytes = 0;
      const size_t cap = total + total / 4;
      CK(cudaMalloc(&ctx->d_best, cap));
      ctx->best_bytes = cap;
   }
   // the offsets depend on n: the ticket must be zero whatever the previous call's layout was
   char *p = (char *)ctx->d_best;
   b->cands = (vnsgpu_cand *)p; b->len = (double *)(p + o_len); b->flags = (uint8_t *)(p + o_fl);
   b->partial = (BestKey *)(p + o_part); b->ticket = (unsigned *)(p + o_tick); b->rec = (BestRec *)(p + o_rec);
   b->mykey = (BestKey *)(p + o_my); b->keys = (BestKey *)(p + o_keys);
   CK(cudaMemsetAsync(b->ticket, 0, 16, ctx->stream));
   return VNSGPU_OK;
}

// evaluate neighbourhood `kind` (0 greedy grid, 1 insert, 2 exchange) of the path's current state and reduce it; the
// record stays on the device (b->rec) and is copied to *out
static vnsgpu_status best_step(vnsgpu_ctx *ctx, int path, int kind, double budget, BestBufs *b, BestRec *out, long long *n_out, bool fresh = false)
{
   PathHdr h;
   vnsgpu_status s = read_hdr(ctx, path, &h);
   if (s) return s;
   const int T = h.hdr[HDR_T], nvns = h.hdr[HDR_NVNS];
   long long n = 0;
   if (kind == 0) n = (long long)(nvns - T) * (T + 1);
   else if (kind == 1) n = nvns >= 3 ? (long long)nvns * (T + 1) : 0;                 // guard :498
   else n = (nvns >= 4 && T > 0) ? (long long)nvns * T : 0;                          // guard :711
   // with vnsgpu_set_shard_neighbourhoods and a communicator every rank evaluates one contiguous range of the enumeration
   // (table and path are replicated) and the ranks' best keys are all-gathered: one collective of 24 bytes per rank
   // not the greedy grid of a large instance: its rounds take the add-only endpoint-vector form (eval_mode_hinted), whose
   // per-state build is replicated and whose evaluation costs less than the collective (20 k nodes, 2 GPUs: 0.92 s sharded
   // against 0.66 s with every rank evaluating the whole grid, r7a)
   const bool cheap_round = kind == 0 && ctx->I.C >= 512 && ctx->use_vec && vnsgpu_int_path_active(ctx);
   const bool shard = ctx->shard_nbh && ctx->comm && ctx->comm_size > 1 && !cheap_round;
   const int R = shard ? ctx->comm_size : 1, r = shard ? ctx->comm_rank : 0;
   const long long k0 = n * r / R, nl = n * (r + 1) / R - k0;
   if (n_out) *n_out = nl;
   out->idx = -1; out->from = out->to = -1; out->a = out->b = 0; out->kind = kind;
   if (n <= 0) return VNSGPU_OK;
   s = best_bufs(ctx, std::max<long long>(nl, 1), b);
   if (s) return s;
   const PathDev P = ctx->paths[path].d;
   const int grid = (int)std::max<long long>(1, std::min<long long>((nl + 255) / 256, 2048));
   if (nl > 0) {
      k_nbh_cands<<<grid, 256, 0, ctx->stream>>>(P, kind, k0, nl, b->cands);
      LAUNCHED(ctx);
      ctx->seg_fresh_path = fresh ? path : -1;      // the products path_rows_from_products built for this very state
      ctx->batch_add_only = kind == 0;              // the greedy grid: tryToAdd only
      s = vnsgpu_eval_batch_dev(ctx, path, nl, b->cands, budget, 0, b->len, nullptr, kind == 0 ? nullptr : b->flags, nullptr, ctx->stream);
      ctx->seg_fresh_path = -1;
      ctx->batch_add_only = false;
      if (s) return s;
   }
   if (kind == 0)
      k_best_reduce<0><<<grid, 256, 0, ctx->stream>>>(ctx->I, P, kind, nl, b->cands, b->len, b->flags, budget, b->partial, b->ticket, b->mykey, k0);
   else
      k_best_reduce<1><<<grid, 256, 0, ctx->stream>>>(ctx->I, P, kind, nl, b->cands, b->len, b->flags, budget, b->partial, b->ticket, b->mykey, k0);
   LAUNCHED(ctx);
   if (shard) {
      s = comm_allgather_bytes(ctx, b->mykey, b->keys, sizeof(BestKey), ctx->stream);
      if (s) return s;
   }
   if (kind == 0) k_best_pick<0><<<1, 1, 0, ctx->stream>>>(P, kind, shard ? b->keys : b->mykey, R, b->rec);
   else k_best_pick<1><<<1, 1, 0, ctx->stream>>>(P, kind, shard ? b->keys : b->mykey, R, b->rec);
   LAUNCHED(ctx);
   CK(cudaMemcpyAsync(out, b->rec, sizeof(BestRec), cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   return VNSGPU_OK;
}

// true: the move was applied through the products, which are current for the next evaluation of this state
static bool apply_uses_products(const vnsgpu_ctx *ctx)
{
   // the sequential re-scan of one block is T dependent layer steps (~0.5 us each); the products cost a few launches:
   // worth it from a few hundred layers on
   if (const char *e = getenv("VNSGPU_APPLY_PRODUCTS")) return path_products_available(ctx) && atoi(e) != 0;   // tests: force either form
   return path_products_available(ctx) && ctx->I.C >= 256;
}

static vnsgpu_status apply_rec(vnsgpu_ctx *ctx, int path, const BestRec *d_rec, int from, int to, int kind, bool *fresh)
{
   *fresh = false;
   if (apply_uses_products(ctx)) {
      k_splice_move<<<1, VNS_CHAIN_THREADS, 0, ctx->stream>>>(ctx->I, ctx->paths[path].d, d_rec, from, to, kind);
      LAUNCHED(ctx);
      vnsgpu_status s = path_rows_from_products(ctx, path, ctx->stream);
      if (s) return s;
      *fresh = true;
      return VNSGPU_OK;
   }
   k_apply_move<<<1, VNS_CHAIN_THREADS, 0, ctx->stream>>>(ctx->I, ctx->paths[path].d, d_rec, from, to, kind);
   LAUNCHED(ctx);
   ctx->paths[path].hdr_ok = false;
   return VNSGPU_OK;
}

extern "C" {

vnsgpu_status vnsgpu_set_shard_neighbourhoods(vnsgpu_ctx *ctx, int enable)
{
   ctx->shard_nbh = enable != 0;
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_greedy_batched(vnsgpu_ctx *ctx, int path, double budget, int *rounds, int64_t *evals)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   int nr = 0;
   int64_t ne = 0;
   bool fresh = false;
   for (;;) {
      BestBufs b;
      BestRec r;
      long long n = 0;
      vnsgpu_status s = best_step(ctx, path, 0, budget, &b, &r, &n, fresh);
      if (s) return s;
      ne += n;
      if (r.idx < 0) break;                                              // :1129
      s = apply_rec(ctx, path, b.rec, 0, 0, 0, &fresh);
      if (s) return s;
      ++nr;
   }
   if (rounds) *rounds = nr;
   if (evals) *evals = ne;
   return check_path_err(ctx, path);
}

vnsgpu_status vnsgpu_best_move(vnsgpu_ctx *ctx, int path, int kind, double budget, vnsgpu_best_move_t *out)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   if (kind != 1 && kind != 2) return fail(ctx, VNSGPU_E_RANGE, "kind must be 1 (insert) or 2 (exchange)");
   BestBufs b;
   BestRec r;
   long long n = 0;
   vnsgpu_status s = best_step(ctx, path, kind, budget, &b, &r, &n);
   if (s) return s;
   if (out) {
      out->index = r.idx; out->from = r.from; out->to = r.to; out->kind = kind;
      out->added_reward = r.a; out->length = r.b; out->n_candidates = n;
   }
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_apply_move(vnsgpu_ctx *ctx, int path, int kind, int from, int to)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   if (kind != 1 && kind != 2) return fail(ctx, VNSGPU_E_RANGE, "kind must be 1 (insert) or 2 (exchange)");
   PathHdr h;
   vnsgpu_status s = read_hdr(ctx, path, &h);
   if (s) return s;
   const int T = h.hdr[HDR_T], nvns = h.hdr[HDR_NVNS], gap = kind == 1 ? 1 : 2, tohi = kind == 1 ? T : T - 1;
   if (from < 0 || from >= nvns || to < 0 || to > tohi || std::abs(from - to) <= gap) return fail(ctx, VNSGPU_E_RANGE, "move outside the neighbourhood");
   bool fresh = false;
   s = apply_rec(ctx, path, nullptr, from, to, kind, &fresh);
   if (s) return s;
   return check_path_err(ctx, path);
}

vnsgpu_status vnsgpu_best_improvement(vnsgpu_ctx *ctx, int path, double budget, int max_steps, int *steps, int64_t *evals)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   int ns = 0;
   int64_t ne = 0;
   int kind = 1, failed = 0;
   bool fresh = false;
   while (failed < 2 && (max_steps <= 0 || ns < max_steps)) {
      BestBufs b;
      BestRec r;
      long long n = 0;
      vnsgpu_status s = best_step(ctx, path, kind, budget, &b, &r, &n, fresh);
      if (s) return s;
      ne += n;
      if (r.idx < 0) { ++failed; kind = 3 - kind; continue; }      // the state did not change: the products stay current
      s = apply_rec(ctx, path, b.rec, 0, 0, 0, &fresh);
      if (s) return s;
      ++ns;
      failed = 0;
      kind = 1;                       // variable neighbourhood descent: back to the first neighbourhood after every move
   }
   if (steps) *steps = ns;
   if (evals) *evals = ne;
   return check_path_err(ctx, path);
}

} // extern "C"
