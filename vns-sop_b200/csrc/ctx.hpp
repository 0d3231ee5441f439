// Host-side context of libvnsgpu.so, shared by its translation units (vnsgpu.cu: context + distance table;
// api_path.cu: path handles + batched evaluation; api_search.cu: neighbourhood drivers; chain2.cu: the search chain).
#pragma once
#include <math_constants.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/vnsgpu.h"
#include "common.cuh"

static_assert(sizeof(vnsgpu_rng) == sizeof(RngState), "rng layout");
static_assert(sizeof(vnsgpu_cand) == 16, "cand layout");

#define VNS_PIPE_EVENTS 16

// host mirror of the path header (read back after every mutating call)
struct PathHdr {
   int hdr[HDR_INTS];
   double hd[HD_DBLS];
};

struct PathHost {
   PathHdr hdr_cache{};      // the header as last read back (every mutating call ends with a read); valid while hdr_ok
   bool hdr_ok = false;
   void *block = nullptr;
   size_t bytes = 0;         // size of the block (one cudaMalloc: header .. minB)
   PathDev d{};
};

// everything the stream operations of the host-buffer evaluation call bake in (see eval_batch_host)
struct HostGraphKey {
   InstDev I;
   PathDev P;
   const void *src, *len, *lb, *flags, *leni, *scratch, *D32, *D16, *vec, *veclb, *segmeta, *hist, *perm;
   cudaStream_t sk;
   int64_t n;
   double budget;
   size_t perm_cap;
   int skip, packed, mode, sort, use_int, G;
};

struct PhaseEvent {
   int phase;
   cudaEvent_t e0, e1;
};

struct vnsgpu_ctx {
   int device = 0;
   cudaStream_t own = nullptr, stream = nullptr;
   cudaStream_t s_in = nullptr, s_out = nullptr;                 // copy streams of the host-buffer pipeline
   cudaEvent_t ev_in[VNS_PIPE_EVENTS] = {}, ev_k[VNS_PIPE_EVENTS] = {}, ev_start = nullptr;
   InstDev I{};
   bool have_clusters = false, have_table = false;
   std::vector<int> m, off;
   std::vector<double> reward;
   int *d_m = nullptr, *d_off = nullptr, *d_clof = nullptr;
   double *d_reward = nullptr, *d_D = nullptr, *d_dmin = nullptr, *d_dmax = nullptr;
   int *d_D32 = nullptr;     // int32 copy of an all-integer table (DPX evaluation path), else null
   unsigned short *d_D16 = nullptr;   // uint16 copy when every entry is an integer below 65535: the endpoint-vector builds stream it
   bool use_u16 = true;      // vnsgpu_set_u16_table
   bool use_int = true;      // vnsgpu_set_int_path
   bool sort_batches = true; // evaluate batches in chain-length order (vnsgpu_set_batch_sort)
   bool use_seg = true;      // exact segment products on integer tables (vnsgpu_set_segment_cache)
   bool use_seg_tiles_always = false;   // enable = 3: the two-level build whatever the size (tests)
   bool use_seg_tiles = true; // two-level build of the segment products on long tours (enable = 2 of vnsgpu_set_segment_cache: off)
   bool seg_ok = false;      // the int32 buffer has room for them and offsets fit 32 bits
   int *d_segmeta = nullptr;
   bool use_vec_staged = false; // staged (bulk-copy) evaluation kernel on tables beyond the caches: opt-in, VNSGPU_VEC_STAGED=1 (slower
                                // than the prefetch form so far: 1.10 ms against 0.73 ms on the 20 k-node table, profiles/r3d_*)
   bool use_vec_pack = true; // several small clusters per warp in the endpoint-vector build (vnsgpu_set_vector_cache(ctx, 3): off)
   int use_vec = 1;          // 0 off, 1 when the batch is large enough, 2 always; endpoint vectors on top of the segment products (vnsgpu_set_vector_cache, vec.cuh)
   int *d_vec = nullptr, *d_veclb = nullptr;
   int group_lanes = 8;      // lanes per candidate of k_eval_groups, from the mean cluster size
   unsigned *d_hist = nullptr, *d_perm = nullptr;
   size_t perm_cap = 0;
   unsigned long long *d_info = nullptr;
   std::vector<PathHost> paths;
   // scratch
   void *d_scratch = nullptr;
   size_t scratch_bytes = 0;
   int *d_seg = nullptr;
   RngState *d_rng = nullptr;
   LsStats *d_stats = nullptr;
   long long *d_first = nullptr;
   VnsResultDev *d_res = nullptr;
   long long *d_prof = nullptr;   // phase cycles of the shared-memory chain
   int *d_stop = nullptr;         // stop flag of the multi-start race
   double target_reward = 0.0;    // vnsgpu_set_target_reward
   float table_kernel_ms = 0.f;   // device time of the last k_table_pairs launch
   std::string err;
   int64_t launches = 0;
   int num_sms = 148;
   int max_smem = 48 * 1024;      // opt-in shared memory per block
   bool use_smem_chain = true;    // vnsgpu_set_smem_chain
   // second search chain (chain2.cuh): exact-int tables, lower bound first, 1 .. num_sms cooperating thread blocks
   int chain_form = -1;           // vnsgpu_set_chain_form: -1 from the instance size, 0 second form, 1 first form
   int chain_blocks = 0;          // vnsgpu_set_chain_blocks: 0 = from the instance size
   int *d_dmin32 = nullptr;       // cluster-pair minima as int32
   void *d_cinfo = nullptr;       // int2 {first sample, size} per cluster
   void *d_c2grid = nullptr;      // grid barrier + control words, slots, RNG jump matrices
   int c2_ncta = 0;
   // multi-GPU exchange steps (comm.cu): NCCL communicator, created from a unique id or adopted from the caller
   void *comm = nullptr;
   bool comm_owned = false;
   int comm_rank = 0, comm_size = 1;
   // device-side candidate generation (gen.cu)
   unsigned *d_genjump = nullptr;
   size_t genjump_lo_words = 0, genjump_mid_words = 0;
   void *h_pick = nullptr;        // pinned record vnsgpu_propose_eval reads back
   void *d_gen = nullptr, *d_gencands = nullptr;
   size_t gen_bytes = 0, gencands_bytes = 0;
   bool batch_add_only = false;   // set by best.cu right before an evaluation call: the batch holds tryToAdd candidates only
   bool vec_skip_pairs = false;   // while such a batch is evaluated in endpoint-vector form: no pair vectors, no sort
   int seg_fresh_path = -1;       // set by best.cu right before an evaluation call: the products of this path's state are current
   bool shard_nbh = false;        // vnsgpu_set_shard_neighbourhoods: full-neighbourhood modes split every neighbourhood over the communicator's ranks
   void *d_best = nullptr;        // full-neighbourhood modes (best.cu): descriptors, lengths, flags, partial keys
   size_t best_bytes = 0;
   bool phase_timing = false;     // vnsgpu_set_phase_timing: CUDA events around the phases of the batched evaluation call
   std::vector<PhaseEvent> phase_events;
   int chain_tree = -1;           // vnsgpu_set_chain_tree: -1 from the instance size, 0 off, 1 on
   int *d_c2tree = nullptr;       // product-tree replicas
   size_t c2_tree_bytes = 0;
   int *d_c2rows = nullptr;       // F / Brev replicas of instances beyond shared memory
   size_t c2_rows_bytes = 0;
   size_t c2_slots_off = 0, c2_jump_off = 0;
   // the host-buffer evaluation call as a CUDA graph (eval_batch_host)
   bool use_graph = true, hg_seen = false;
   HostGraphKey hg_key, hg_last;
   cudaGraphExec_t hg_exec = nullptr;
   int64_t hg_launches = 0;
   long long *h_first = nullptr;  // pinned
   cudaEvent_t ev_out = nullptr;
};

#define CK(call)                                                                                   \
   do {                                                                                            \
      cudaError_t e__ = (call);                                                                    \
      if (e__ != cudaSuccess) {                                                                    \
         ctx->err = std::string(#call) + ": " + cudaGetErrorString(e__);                           \
         return VNSGPU_E_CUDA;                                                                     \
      }                                                                                            \
   } while (0)

#define LAUNCHED(ctx)                                                                              \
   do {                                                                                            \
      (ctx)->launches++;                                                                           \
      cudaError_t e__ = cudaGetLastError();                                                        \
      if (e__ != cudaSuccess) {                                                                    \
         (ctx)->err = std::string("kernel launch: ") + cudaGetErrorString(e__);                   \
         return VNSGPU_E_CUDA;                                                                     \
      }                                                                                            \
   } while (0)

static inline vnsgpu_status fail(vnsgpu_ctx *ctx, vnsgpu_status s, const char *msg)
{
   ctx->err = msg;
   return s;
}

static inline vnsgpu_status need_scratch(vnsgpu_ctx *ctx, size_t bytes)
{
   if (bytes <= ctx->scratch_bytes) return VNSGPU_OK;
   if (ctx->d_scratch) CK(cudaFree(ctx->d_scratch));
   ctx->d_scratch = nullptr;
   ctx->scratch_bytes = 0;
   CK(cudaMalloc(&ctx->d_scratch, bytes));
   ctx->scratch_bytes = bytes;
   return VNSGPU_OK;
}


static inline bool group_kernel_ok(const vnsgpu_ctx *ctx)
{
   const int G = ctx->group_lanes, nr = (ctx->I.mp + G - 1) / G;
   return nr <= (G == 8 ? 4 : (G == 16 ? 2 : 1)) && (long long)ctx->I.N * ctx->I.N < (1ll << 32);
}

// ---- helpers shared by the API units ---------------------------------------------------------------
static inline bool path_ok(vnsgpu_ctx *ctx, int path)
{
   if (!ctx->have_table) { ctx->err = "no distance table"; return false; }
   if (path < 0 || path >= (int)ctx->paths.size() || !ctx->paths[path].block) { ctx->err = "bad path handle"; return false; }
   return true;
}

static inline size_t al16(size_t x) { return (x + 15) & ~(size_t)15; }

static inline vnsgpu_status read_hdr(vnsgpu_ctx *ctx, int path, PathHdr *h)
{
   const PathDev &d = ctx->paths[path].d;
   // hdr and hd are adjacent at the start of the block
   CK(cudaMemcpyAsync(h, d.hdr, sizeof(PathHdr), cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   ctx->paths[path].hdr_cache = *h;
   ctx->paths[path].hdr_ok = true;
   return VNSGPU_OK;
}

static inline vnsgpu_status check_path_err(vnsgpu_ctx *ctx, int path)
{
   PathHdr h;
   vnsgpu_status s = read_hdr(ctx, path, &h);
   if (s) return s;
   if (h.hdr[HDR_ERR] == 4) return fail(ctx, VNSGPU_E_INCONSISTENT, "forward/backward path length mismatch");
   if (h.hdr[HDR_ERR] == 3) return fail(ctx, VNSGPU_E_STATE, "proposal rejection loop did not terminate");
   if (h.hdr[HDR_ERR]) return fail(ctx, VNSGPU_E_STATE, "device-side error flag");
   return VNSGPU_OK;
}

static inline size_t chain_smem(const vnsgpu_ctx *ctx) { return sizeof(double) * 2 * ctx->I.mp * VNS_CHAIN_WARPS; }


// phase timing (measure.cu): PhaseTimer t(ctx, phase, stream) records an event pair around the launches in its scope
struct PhaseTimer {
   vnsgpu_ctx *ctx;
   cudaStream_t st;
   PhaseEvent pe;
   bool on;
   PhaseTimer(vnsgpu_ctx *c, int phase, cudaStream_t s) : ctx(c), st(s), on(c->phase_timing)
   {
      if (!on) return;
      pe.phase = phase;
      if (cudaEventCreate(&pe.e0) != cudaSuccess || cudaEventCreate(&pe.e1) != cudaSuccess) { on = false; cudaGetLastError(); return; }
      cudaEventRecord(pe.e0, st);
   }
   ~PhaseTimer()
   {
      if (!on) return;
      cudaEventRecord(pe.e1, st);
      ctx->phase_events.push_back(pe);
   }
};

// comm.cu: all-gather of a few bytes per rank (best.cu)
vnsgpu_status comm_allgather_bytes(vnsgpu_ctx *ctx, const void *d_send, void *d_recv, size_t bytes, cudaStream_t st);
// internal entry points of api_path.cu used by best.cu: every F / B row of a path from the segment products of its current
// targets (all layers in parallel; integer tables)
bool path_products_available(const vnsgpu_ctx *ctx);
vnsgpu_status path_rows_from_products(vnsgpu_ctx *ctx, int path, cudaStream_t st, bool init_ends = false);
// internal entry points of api_path.cu used by gen.cu: evaluation of device candidates for the first accepted index only;
// format 0 = vnsgpu_cand, 2 = vnsgpu_pack_cand32 words (only when eval_dev_accepts_packed32 says so)
bool eval_dev_accepts_packed32(const vnsgpu_ctx *ctx, int64_t n);
vnsgpu_status eval_dev_first(vnsgpu_ctx *ctx, int path, int64_t n, const void *d_cands, int format, double budget, int skip, long long *d_first,
                             cudaStream_t st);
