/* vnsgpu.h - C ABI of the B200 (sm_100a) neighbourhood-evaluation engine for VNS-SOP.
 *
 * Drop-in seam: the reference (ctu-mrs/vns-sop) has no plugin/FFI layer; the boundary is the public
 * interface of op::VnsSopPath (sources/vns_sop_path.h:22-194) + its process-global tables
 * (setSamplesDistances, vns_sop_path.h:30-31,157-160) and the neighbourhood drivers of
 * op::VnsSopSolver that call it (sources/vns_sop_solver.cpp:364-385,469-864,931-1155).
 * Each entry point below names the reference interface it replaces.  INTEGRATION.md shows the
 * binding a maintainer adds to sources/vns_sop_solver.cpp.
 *
 * Conventions: plain pointers and sizes only; every function returns a vnsgpu_status (0 = ok) and
 * never exits the process (the reference exit(1)s on range errors, e.g. vns_sop_path.cpp:69-72);
 * host pointers are copied during the call; the context owns all device memory; one host thread
 * per context.  "_dev" variants take DEVICE pointers and a CUDA stream handle and do not
 * synchronise.  Clusters are numbered as after the reference loader's renumbering
 * (sop_loader.cpp:118-156); samples are stored cluster-contiguously.
 */
#ifndef VNSGPU_H
#define VNSGPU_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct vnsgpu_ctx vnsgpu_ctx;

typedef enum {
   VNSGPU_OK = 0,
   VNSGPU_E_RANGE = 1,        /* index outside the range the reference exit(1)s on */
   VNSGPU_E_CUDA = 2,         /* CUDA runtime error, see vnsgpu_last_error */
   VNSGPU_E_STATE = 3,        /* call order (no table, bad handle, ...) */
   VNSGPU_E_INCONSISTENT = 4, /* forward/backward length mismatch (vns_sop_path.cpp:80-88) or
                                 over-budget improved tour (vns_sop_solver.cpp:229-237) */
   VNSGPU_E_NOMEM = 5,
   VNSGPU_E_TOO_LARGE = 6,    /* the instance does not fit the requested form (e.g. vnsgpu_vns_run_batch: search state
                                 beyond one SM's shared memory); the caller may fall back to another entry point */
   VNSGPU_E_NCCL = 7          /* NCCL error in a vnsgpu_comm_* call, see vnsgpu_last_error */
} vnsgpu_status;

/* edge-weight rules of SOPLoader (sop_loader.cpp:278,360,474-486) */
enum { VNSGPU_RULE_RAW = 0, VNSGPU_RULE_ROUND = 1, VNSGPU_RULE_CEIL = 2 };

/* candidate operations = the const queries of VnsSopPath */
enum {
   VNSGPU_OP_ADD = 0,      /* tryToAdd(a = cluster, b = idx)           vns_sop_path.cpp:245-304  */
   VNSGPU_OP_REPLACE = 1,  /* tryToReplace(a = cluster, b = idx, 1)    vns_sop_path.cpp:456-513  */
   VNSGPU_OP_REMOVE = 2,   /* tryToRemove(a = idx)                     vns_sop_path.cpp:515-535  */
   VNSGPU_OP_MOVE = 3,     /* tryToMove(a = what, b = where)           vns_sop_path.cpp:932-1154 */
   VNSGPU_OP_EXCHANGE = 4  /* tryToExchange(a = idx1, b = idx2)        vns_sop_path.cpp:635-797  */
};

typedef struct {
   int32_t op, a, b;
   int32_t gain_cluster; /* cluster whose reward is gained (ADD/REPLACE) or -1; used for flags */
} vnsgpu_cand;

/* per-candidate result flags */
enum {
   VNSGPU_F_FEASIBLE = 1, /* len <= budget */
   VNSGPU_F_ACCEPT = 2,   /* len <= budget && (dReward > 0 || L - len > 1e-5)
                             (vns_sop_solver.cpp:619,805-806) */
   VNSGPU_F_LB_SKIP = 4,  /* lower bound > budget: the reference skips the evaluation
                             (vns_sop_solver.cpp:539,574,749,779) */
   VNSGPU_F_BAD = 8       /* indices out of range: len = NaN */
};

/* glibc rand() state (random_r TYPE_3), replayed bit-exactly: math_common.cpp:21-28 */
typedef struct {
   int32_t r[31];
   int32_t f, b;
   int32_t pad;
} vnsgpu_rng;

typedef struct {
   int64_t proposals, lb_skips, evals, accepts, rounds;
} vnsgpu_ls_stats;

typedef struct {
   int32_t num_iterations;            /* sop.cfg num-iterations            (vns_sop_solver.cpp:64) */
   int32_t num_iterations_unimproved; /* sop.cfg num-iterations-unimproved (:65) */
   int32_t lower_bound_skip;          /* lower-bound-above-budget-then-skip (:68) */
   int32_t greedy_initial;            /* initial-greedy-solution (:70) */
   double budget;                     /* TMAX */
   double max_time_ms;                /* maximal-calculation-time-sec * 1000 (:67); <=0: none */
   uint32_t seed;                     /* srand(seed) hook (math_common.cpp:9-14) */
   int32_t max_improvements;          /* capacity of the improvement log arrays */
} vnsgpu_vns_params;

typedef struct {
   int32_t num_iters, num_iters_last_impr, n_improvements, status;
   double reward, length;
   double time_ms, time_last_impr_ms;
   int64_t proposals, evals, accepts;
} vnsgpu_vns_result;

/* ---- context ------------------------------------------------------------------------------ */
vnsgpu_status vnsgpu_create(vnsgpu_ctx **out, int device);
void vnsgpu_destroy(vnsgpu_ctx *ctx);
const char *vnsgpu_last_error(const vnsgpu_ctx *ctx);
const char *vnsgpu_version(void);
/* number of kernels this context has launched so far (bench.py's gpu_launches) */
int64_t vnsgpu_launch_count(const vnsgpu_ctx *ctx);
/* run everything on this CUDA stream (cudaStream_t as void*); NULL = the context's own stream */
vnsgpu_status vnsgpu_set_stream(vnsgpu_ctx *ctx, void *stream);
vnsgpu_status vnsgpu_sync(vnsgpu_ctx *ctx);
/* Tables whose entries are all integers (every shipped dataset: CEIL_2D / EUC_2D / integer EXPLICIT)
 * are evaluated in exact int32 arithmetic on a 4-byte copy of the table (bit-identical results);
 * enable = 0 forces the IEEE double path.  Default: enabled. */
vnsgpu_status vnsgpu_set_int_path(vnsgpu_ctx *ctx, int enable);
int vnsgpu_int_path_active(const vnsgpu_ctx *ctx);
/* Batches of >= 8192 candidates are evaluated in chain-length order (a device-side counting sort of the
 * descriptors; results still land at the caller's indices).  The endpoint-vector form keeps the caller's order on
 * cache-resident tables with small clusters (nothing to gain there) and sorts by product row on tables beyond the
 * caches (L2 locality).  enable = 0 keeps the caller's order everywhere. */
vnsgpu_status vnsgpu_set_batch_sort(vnsgpu_ctx *ctx, int enable);
/* Integer tables, batches of >= 16384: the min-plus product of every run of consecutive path layers is
 * built once per call and shared by all candidates (exactly associative on integers: bit-identical
 * lengths); a MOVE / EXCHANGE chain of `span` blocks then costs at most 5.  enable = 0: block by block. */
vnsgpu_status vnsgpu_set_segment_cache(vnsgpu_ctx *ctx, int enable);
/* Integer tables, batches of >= C^2 / 4 candidates (on top of the segment products): the two ends of every
 * candidate's chain depend only on (cluster, layer) - they are built once per call as vectors, and a candidate
 * costs at most ONE block step (exact: integer min-plus is associative).  enable = 0: segment products only;
 * 2: also for smaller batches. */
vnsgpu_status vnsgpu_set_vector_cache(vnsgpu_ctx *ctx, int enable);
/* All-integer tables whose entries are below 65535 also get a uint16 copy (2 N^2 bytes); the endpoint-vector builds,
 * which stream whole row bands of the table, read that copy (half the HBM bytes; the values are the same integers).
 * enable = 0 makes them read the int32 copy (tests run both).  Default: enabled. */
vnsgpu_status vnsgpu_set_u16_table(vnsgpu_ctx *ctx, int enable);
/* vnsgpu_local_search keeps the whole search state in one SM's shared memory when it fits (clusters <=
 * 510, samples per cluster <= 32); enable = 0 forces the global-memory chain (tests run both). */
vnsgpu_status vnsgpu_set_smem_chain(vnsgpu_ctx *ctx, int enable);
/* Form of the shared-memory chain on all-integer tables.  -1 (default): the third form below 96 clusters, or when the whole
 * distance table fits one SM's shared memory (N^2 int32 or uint16 entries + the search state below ~220 KB: every shipped
 * OPN / DOP instance up to 12 headings and the ~200-node GTSP instances); the second form from 96 clusters on; the first
 * form for what is left (cluster sizes beyond 32, real-valued tables).  0 (or 2): the second form - the lower bound of every
 * proposal first (one thread each, O(1) from prefix sums along the path), only the survivors are evaluated, up to 256
 * draw pairs per thread block and round, shake + fitDOPtoBudget in shared memory as well, product tree, cooperating
 * thread blocks, instances beyond shared memory.  1: the first form (one sub-warp group per proposal, 64 proposals per
 * round).  3: the third form - table, cluster-pair minima and state all in shared memory, every warp derives the
 * round's window itself (no compaction barrier), bookkeeping replicated in registers, generator running ahead during
 * the re-scan; below 96 clusters a table that does not fit stays in global memory (read-only path) and the rest is the same;
 * otherwise falls back to the default choice.  Same trajectory in every form (tests
 * run all of them). */
vnsgpu_status vnsgpu_set_chain_form(vnsgpu_ctx *ctx, int form);
/* Thread blocks that cooperate on ONE trajectory-exact chain (second form; cooperative launch).  Every block holds a
 * replica of the search state and evaluates its own slice of 256 draw pairs of the round's window (it reaches its
 * slice of the glibc rand() stream with one 31 x 31 jump matrix); one grid barrier per round; the first slice with an
 * accepted proposal wins, exactly the proposal the reference's sequential loop would have applied.  0 (default): from the
 * instance size (1 block up to ~128 clusters, then |clusters|^2 / 2048, at most one per SM). */
vnsgpu_status vnsgpu_set_chain_blocks(vnsgpu_ctx *ctx, int blocks);
/* Product tree of the second chain form: an aligned segment tree of exact min-plus products over the path's edges.  A run
 * of consecutive layers inside a tryToMove / tryToExchange chain costs at most 2 log2(T) steps instead of `span`, and
 * after an accepted move every layer's shortest / shortest_back row is recomputed independently through the tree
 * (updateAfterInsert / updateAfterRemove, vns_sop_path.cpp:1406-1540, are T dependent steps).  Exact on integer tables
 * (associativity): same trajectory.  mode -1 (default): on from 96 clusters; 0 off; 1 on (tests run both). */
vnsgpu_status vnsgpu_set_chain_tree(vnsgpu_ctx *ctx, int mode);

/* ---- instance + distance table: replaces VnsSopPath::setSamplesDistances (vns_sop_path.cpp:38-48)
 *      fed by SOPLoader::getSOPFromNodes / getSOPFromWeightMatrix / getOP / getOPN ---------------- */
vnsgpu_status vnsgpu_set_clusters(vnsgpu_ctx *ctx, int C, const int *m, const double *reward, int start,
                                  int goal);
/* a1/a3: D = rule(sqrt(dx*dx+dy*dy)); xy = N x 2 doubles in sample order (sop_loader.cpp:457-491) */
vnsgpu_status vnsgpu_build_table_euclid(vnsgpu_ctx *ctx, const double *xy, int rule);
/* a4 (new; reference ships the result as EXPLICIT matrices): every cluster has S heading samples
 * theta_k = 2 pi k / S at its centre; D = ceil(Dubins length) or the raw length if !do_ceil */
vnsgpu_status vnsgpu_build_table_dubins(vnsgpu_ctx *ctx, const double *centers, int S, double rho, int do_ceil);
/* a2: gather from an explicit N x N row-major matrix in sample order (sop_loader.cpp:413-427) */
vnsgpu_status vnsgpu_set_table_dense(vnsgpu_ctx *ctx, const double *W);
vnsgpu_status vnsgpu_get_table_dense(vnsgpu_ctx *ctx, double *W);
/* a5: calcClusterDistancesBounds (vns_sop_solver.cpp:262-297); C x C each; either may be NULL */
vnsgpu_status vnsgpu_get_bounds(vnsgpu_ctx *ctx, double *dmin, double *dmax);
/* device time (ms, CUDA events) of the last table build kernel: one fused pass writing the table as double (8 N^2 B) and,
 * for integer-valued rules, as int32 (4 N^2 B), with the a5 bounds and the table classification in its epilogue */
double vnsgpu_table_kernel_ms(const vnsgpu_ctx *ctx);
/* bytes of the device-resident table (8 N^2) */
int64_t vnsgpu_table_bytes(const vnsgpu_ctx *ctx);

/* ---- paths: value-type VnsSopPath objects become handles (copy = device memcpy) -------------- */
vnsgpu_status vnsgpu_path_create(vnsgpu_ctx *ctx, int *path);
/* ~VnsSopPath: frees the handle's device memory; the handle value may be handed out again */
vnsgpu_status vnsgpu_path_destroy(vnsgpu_ctx *ctx, int path);
vnsgpu_status vnsgpu_path_copy(vnsgpu_ctx *ctx, int dst, int src);
/* targets := vns[0..T) and full update() (vns_sop_path.cpp:1321-1404); vns = the solver's VNS vector */
vnsgpu_status vnsgpu_path_set(vnsgpu_ctx *ctx, int path, const int *vns, int nvns, int T);
/* addPoint(c, idx) / removePoint(idx) + updateAfterInsert/Remove (vns_sop_path.cpp:67-103,154-180,
 * 1406-1540); these touch targets only, like the reference (the VNS vector is the caller's) */
vnsgpu_status vnsgpu_path_add_point(vnsgpu_ctx *ctx, int path, int cluster, int idx);
vnsgpu_status vnsgpu_path_remove_point(vnsgpu_ctx *ctx, int path, int idx);
/* getNumTargets/getAllTargets/getPathLength/getPathLengthRear/getReward (+ the VNS vector);
 * any output pointer may be NULL */
vnsgpu_status vnsgpu_path_get(vnsgpu_ctx *ctx, int path, int *T, int *targets, int *nvns, int *vns,
                              double *length, double *length_rear, double *reward);
/* getPath (vns_sop_path.cpp:1542-1599) + getPathLengthCalculate (:1700-1741): T+2 entries */
vnsgpu_status vnsgpu_path_get_solution(vnsgpu_ctx *ctx, int path, int *cluster, int *node, double *length_calc);
/* shortest / shortest_back as (T+2) x stride doubles (tests; stride >= max cluster size) */
vnsgpu_status vnsgpu_path_get_dp(vnsgpu_ctx *ctx, int path, double *F, double *B, int stride);

/* ---- batched candidate evaluation (a8-a13): all tryTo* + lower bounds + accept predicate ------ */
/* host buffers; len/lb/flags may be NULL; first_accept (may be NULL) receives the smallest index
 * whose ACCEPT flag is set and (when skip != 0) whose LB_SKIP flag is not, or -1 */
vnsgpu_status vnsgpu_eval_batch(vnsgpu_ctx *ctx, int path, int64_t n, const vnsgpu_cand *cands, double budget,
                                int skip, double *len, double *lb, uint8_t *flags, int64_t *first_accept);
/* The same call with 8-byte descriptors (halves the host-to-device bytes; the library expands them on the
 * device).  Indices must lie in [0, 2^20); gain_cluster in [-1, 2^21 - 2]. */
static inline uint64_t vnsgpu_pack_cand(int op, int a, int b, int gain_cluster)
{
   return (uint64_t)(op & 7) | ((uint64_t)(a & 0xfffff) << 3) | ((uint64_t)(b & 0xfffff) << 23) | ((uint64_t)(gain_cluster + 1) << 43);
}
vnsgpu_status vnsgpu_eval_batch_packed(vnsgpu_ctx *ctx, int path, int64_t n, const uint64_t *packed, double budget,
                                       int skip, double *len, double *lb, uint8_t *flags, int64_t *first_accept);
/* Compact wire format for link-bound batches on ALL-INTEGER tables (vnsgpu_int_path_active): 4-byte descriptors in,
 * int32 lengths out - 4 + 5 bytes per candidate over PCIe instead of 8 + 9.  Every length of an integer table is an
 * integer below 2^31, so (double)len[i] is the reference's tryTo* value exactly; bad indices give INT32_MIN (the
 * double calls give NaN).  Indices must lie in [0, 16384); gain_is_a != 0: the candidate's gain cluster is a (what
 * insertRandom / exchangeRandom pass for ADD / REPLACE), else none.  VNSGPU_E_STATE on a real-valued table. */
static inline uint32_t vnsgpu_pack_cand32(int op, int a, int b, int gain_is_a)
{
   return (uint32_t)(op & 7) | ((uint32_t)(a & 0x3fff) << 3) | ((uint32_t)(b & 0x3fff) << 17) | ((uint32_t)(gain_is_a != 0) << 31);
}
vnsgpu_status vnsgpu_eval_batch_packed32(vnsgpu_ctx *ctx, int path, int64_t n, const uint32_t *packed, double budget,
                                         int skip, int32_t *len, uint8_t *flags, int64_t *first_accept);
/* The same call with ONE 4-byte result word per candidate: length | flags << 29 (FEASIBLE / ACCEPT / LB_SKIP in bits
 * 29-31; every length of the exact int32 path is below 2^29), 0xffffffff for bad indices - 4 + 4 bytes per candidate
 * over the link.  vnsgpu_result_len / vnsgpu_result_flags take a word apart. */
static inline int32_t vnsgpu_result_len(uint32_t w)
{
   return (int32_t)(w & 0x1fffffffu);
}
static inline int vnsgpu_result_flags(uint32_t w)
{
   return w == 0xffffffffu ? VNSGPU_F_BAD : (int)(w >> 29);
}
vnsgpu_status vnsgpu_eval_batch_packed32r(vnsgpu_ctx *ctx, int path, int64_t n, const uint32_t *packed, double budget,
                                          int skip, uint32_t *result, int64_t *first_accept);
/* Allocation-free variant of vnsgpu_eval_batch.  It has no counterpart in the reference and was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, results and flags as vnsgpu_eval_batch; the staged descriptors and results
 * live in a file-scope static __device__ array of the library (about 0.5 MB that every GPU which loads the library holds
 * until exit) instead of context scratch, the candidates are evaluated block by block (no segment products, endpoint
 * vectors or sort: those caches are allocated on demand by the other entry points), and the call makes no cudaMalloc,
 * cudaFree or other allocating call.  n > VNSGPU_NOALLOC_MAX_CANDS returns VNSGPU_E_RANGE with a message in
 * vnsgpu_last_error.  All calls on one device share the array, which keeps the last call's contents: the call clears
 * what has to start clean, and calls that overlap on the GPU (several contexts or streams on one device) race - keep
 * them serial. */
#define VNSGPU_NOALLOC_MAX_CANDS 16384
vnsgpu_status vnsgpu_eval_batch_noalloc(vnsgpu_ctx *ctx, int path, int64_t n, const vnsgpu_cand *cands, double budget,
                                        int skip, double *len, double *lb, uint8_t *flags, int64_t *first_accept);
int64_t vnsgpu_noalloc_max_cands(void);
/* device buffers, asynchronous on `stream` (NULL = context stream); d_first_accept: one int64 that
 * the call resets and fills like first_accept above (-1 = none), or NULL.  The segment products, endpoint vectors and
 * sort scratch a call builds belong to the CONTEXT: at most one evaluation call of a context may be in flight at a time
 * (a second one on another stream, or one racing vnsgpu_path_copy of the same path, overwrites them) - order such calls
 * with an event, or use one context per concurrent evaluation. */
vnsgpu_status vnsgpu_eval_batch_dev(vnsgpu_ctx *ctx, int path, int64_t n, const vnsgpu_cand *d_cands,
                                    double budget, int skip, double *d_len, double *d_lb, uint8_t *d_flags,
                                    long long *d_first_accept, void *stream);
/* Multi-GPU form (candidates sharded over ranks, table and path replicated on each): d_first_key receives index_base +
 * the local index of the shard's first accepted candidate, or VNSGPU_NO_KEY; an all-reduce(MIN) of that one int64 over
 * the ranks is the index the sequential loop would have applied, identical for any number of GPUs. */
#define VNSGPU_NO_KEY 0x7f7f7f7f7f7f7f7fLL
vnsgpu_status vnsgpu_eval_batch_dev_shard(vnsgpu_ctx *ctx, int path, int64_t n, const vnsgpu_cand *d_cands,
                                          double budget, int skip, double *d_len, double *d_lb, uint8_t *d_flags,
                                          long long *d_first_key, int64_t index_base, void *stream);
/* draws n proposals with the reference's proposal distribution for neighbourhood `kind`
 * (1 = insertRandom :507-519 -> MOVE / ADD, 2 = exchangeRandom :723-730 -> EXCHANGE / REPLACE) against
 * the path's current state, advancing *rng exactly as the reference's draws would; HOST buffer */
vnsgpu_status vnsgpu_gen_proposals(vnsgpu_ctx *ctx, int path, int kind, int64_t n, vnsgpu_rng *rng,
                                   vnsgpu_cand *cands);
/* The same proposals generated ON THE DEVICE, in draw order, into a device buffer: format 0 = vnsgpu_cand[n], 2 =
 * vnsgpu_pack_cand32 words.  glibc's stream is linear, so every thread block jumps to its own slice of draws (31 x 31
 * jump matrices) and a scan of the valid-pair counts gives each proposal its index.  d_draws (may be NULL): per proposal
 * the number of rand() calls consumed up to and including it - vnsgpu_rng_advance(&rng_at_entry, d_draws[k]) is the
 * generator just after proposal k, what the reference's loop leaves behind when it accepts k.  *rng (host) advances past
 * proposal n - 1; the call synchronises `stream` (NULL = context stream).  Replaces the draws of insertRandom
 * (vns_sop_solver.cpp:507-519) and exchangeRandom (:723-730) for a whole window at once. */
vnsgpu_status vnsgpu_gen_proposals_dev(vnsgpu_ctx *ctx, int path, int kind, int64_t n, vnsgpu_rng *rng, void *d_out, int format,
                                       uint32_t *d_draws, void *stream);
/* rand() called ndraws times (jump-ahead: O(log ndraws) 31 x 31 matrix-vector products on the host) */
void vnsgpu_rng_advance(vnsgpu_rng *rng, uint64_t ndraws);
/* One round of a host-driven insertRandom / exchangeRandom: n proposals drawn on the device, evaluated, the first accepted
 * one in draw order returned (*first_accept = its index or -1, *accepted = its descriptor), *rng rewound to just after
 * its draws (past all n when none is accepted).  132 bytes in, ~30 bytes out, whatever n. */
vnsgpu_status vnsgpu_propose_eval(vnsgpu_ctx *ctx, int path, int kind, int64_t n, vnsgpu_rng *rng, double budget, int skip,
                                  int64_t *first_accept, vnsgpu_cand *accepted);
/* How vnsgpu_eval_batch_dev would evaluate a batch of n candidates right now: bits 0-1 = form (0 block by block, 1 exact
 * segment products, 2 endpoint vectors + segment products), bit 4 = the batch is sorted on the device first, bit 5 = the
 * table is beyond the caches (HBM form of the endpoint-vector kernel), bits 8-15 = lanes per candidate group */
int vnsgpu_eval_mode(const vnsgpu_ctx *ctx, int64_t n);
/* algorithmic bytes of a candidate list: w*sum m_src*m_dst + 8*(m_first + m_last) + 8,
 * w = bytes per table entry the active evaluation path reads (8 double, 4 exact-int32) */
vnsgpu_status vnsgpu_cand_bytes(vnsgpu_ctx *ctx, int path, int64_t n, const vnsgpu_cand *cands, int64_t *bytes);

/* ---- full-neighbourhood modes on the batched evaluation --------------------------------------------
 * greedy_insertion (vns_sop_solver.cpp:1091-1155) with every round's (|VNS| - T) x (T + 1) tryToAdd grid evaluated by ONE
 * batched call and reduced on the device: strict minimum of (d - L) / reward among d <= budget, first grid index on ties
 * (:1119).  Same tour as vnsgpu_greedy and the reference.  rounds / evals (may be NULL): insertions made / candidates
 * evaluated. */
vnsgpu_status vnsgpu_greedy_batched(vnsgpu_ctx *ctx, int path, double budget, int *rounds, int64_t *evals);
/* Opt-in local search that is NOT trajectory-exact: the WHOLE insert (kind 1: tryToMove / tryToAdd of every (from, to)
 * pair insertRandom can draw, :507-519) or exchange (kind 2: tryToExchange / tryToReplace, :723-730) neighbourhood is
 * evaluated at once and the BEST improving candidate is returned instead of the first drawn one: most added reward,
 * then shortest tour, then lowest index (from-major).  Improving = feasible and (added reward > 0, or the same reward and
 * at least 1e-5 shorter); unlike the reference's acceptance test (:619,:805-806) a shorter tour with LESS reward does not
 * qualify, so a descent terminates.  index = -1: no improving candidate. */
typedef struct {
   int64_t index;          /* from * (tohi + 1) + to, tohi = T (insert) or T - 1 (exchange); -1: none */
   int32_t from, to, kind, pad;
   double added_reward, length;
   int64_t n_candidates;
} vnsgpu_best_move_t;
vnsgpu_status vnsgpu_best_move(vnsgpu_ctx *ctx, int path, int kind, double budget, vnsgpu_best_move_t *out);
/* the accept branch of insertRandom (:621-646, kind 1) / exchangeRandom (:807-832, kind 2) for the pair (from, to):
 * VNS[from] moves in front of index `to` / VNS[from] and VNS[to] swap; F / B re-scanned */
vnsgpu_status vnsgpu_apply_move(vnsgpu_ctx *ctx, int path, int kind, int from, int to);
/* Multi-GPU form of the three calls above and of vnsgpu_greedy_batched: with enable != 0 and a communicator (vnsgpu_comm_*)
 * every rank evaluates one contiguous range of each neighbourhood's enumeration against its replica of the table and the
 * path, the ranks' best keys (24 bytes each) are all-gathered and every rank picks - and applies - the same move: the
 * result is the single-GPU one for any number of ranks.  The calls become collective: every rank makes them in the same
 * order on paths in the same state.  evals / n_candidates then count this rank's share. */
vnsgpu_status vnsgpu_set_shard_neighbourhoods(vnsgpu_ctx *ctx, int enable);
/* variable-neighbourhood descent with the two calls above: best insert move until there is none, then best exchange move,
 * back to insert after every applied move; ends when neither neighbourhood improves (or after max_steps > 0 moves) */
vnsgpu_status vnsgpu_best_improvement(vnsgpu_ctx *ctx, int path, double budget, int max_steps, int *steps, int64_t *evals);

/* ---- multi-GPU exchange steps: candidates shard over ranks, table and incumbent are replicated; the only
 *      collectives are the all-reduce(MIN) of the first-accepted key and the broadcast of the incumbent.  One context per
 *      GPU and rank; the communicator is NCCL's (libnccl.so.2 is bound with dlopen at the first call - VNSGPU_E_NCCL when it
 *      cannot be loaded or a collective fails).  No reference counterpart: the reference is a single-threaded CPU program. */
#define VNSGPU_COMM_ID_BYTES 128
/* rank 0: ncclGetUniqueId; hand the 128 bytes to the other ranks out of band (file, socket, MPI, torch.distributed) */
vnsgpu_status vnsgpu_comm_unique_id(vnsgpu_ctx *ctx, void *id128);
/* every rank, collectively: ncclCommInitRank on the context's device */
vnsgpu_status vnsgpu_comm_init(vnsgpu_ctx *ctx, const void *id128, int rank, int nranks);
/* use a communicator the caller owns (ncclComm_t as void*); the context never destroys it */
vnsgpu_status vnsgpu_comm_adopt(vnsgpu_ctx *ctx, void *nccl_comm, int rank, int nranks);
vnsgpu_status vnsgpu_comm_destroy(vnsgpu_ctx *ctx);
vnsgpu_status vnsgpu_comm_info(const vnsgpu_ctx *ctx, int *rank, int *nranks);
/* in place all-reduce(MIN) of the key vnsgpu_eval_batch_dev_shard wrote: the chosen move, identical on every rank and for
 * any number of ranks; asynchronous on `stream` (NULL = context stream).  Without a communicator (1 GPU): no-op */
vnsgpu_status vnsgpu_reduce_first_key(vnsgpu_ctx *ctx, long long *d_key, void *stream);
/* the incumbent: every rank's `path` becomes a copy of root's (targets, VNS vector, shortest / shortest_back rows, minima:
 * one contiguous device block, one ncclBroadcast) */
vnsgpu_status vnsgpu_path_broadcast(vnsgpu_ctx *ctx, int path, int root, void *stream);
vnsgpu_status vnsgpu_bcast_bytes(vnsgpu_ctx *ctx, void *d_buf, int64_t nbytes, int root, void *stream);

/* ---- measurement support (bench.py; SURVEY.md 8d: "report achieved / peak for whichever level holds the working set") ---
 * Streaming-read bandwidth of the whole device with the data resident in L1 (each thread block re-reads its own 64 KB),
 * in L2 (all blocks re-read one 32 MB buffer) and in HBM (4 GiB read once): gbs[0..2] in GB/s, best of 5. */
vnsgpu_status vnsgpu_measure_read_peaks(vnsgpu_ctx *ctx, double *gbs);
/* Device time of the phases of vnsgpu_eval_batch_dev / _shard (CUDA events on the launching stream): segment products
 * (k_seg_offsets + k_seg_build), endpoint vectors (k_vec_uw + k_vec_pairs), counting sort, evaluation kernel. */
enum { VNSGPU_PHASE_SEG = 0, VNSGPU_PHASE_VEC = 1, VNSGPU_PHASE_SORT = 2, VNSGPU_PHASE_EVAL = 3, VNSGPU_PHASES = 4 };
vnsgpu_status vnsgpu_set_phase_timing(vnsgpu_ctx *ctx, int enable);
/* ms[VNSGPU_PHASES], count[VNSGPU_PHASES]: totals since the last call; synchronise the stream first */
vnsgpu_status vnsgpu_phase_times(vnsgpu_ctx *ctx, double *ms, int64_t *count);

/* ---- neighbourhood drivers (a15, a16) ---------------------------------------------------------- */
void vnsgpu_rng_seed(vnsgpu_rng *rng, uint32_t seed); /* srand(seed) */
int32_t vnsgpu_rng_next(vnsgpu_rng *rng);             /* rand() */
/* insertRandom (kind 1, vns_sop_solver.cpp:469-684) / exchangeRandom (kind 2, :686-864): the whole
 * call runs on the device, trajectory-exact; path's targets/VNS vector and *rng are updated */
vnsgpu_status vnsgpu_local_search(vnsgpu_ctx *ctx, int path, int kind, int num_changes, double budget, int skip,
                                  vnsgpu_rng *rng, vnsgpu_ls_stats *stats);
/* SM cycles the shared-memory chain spent per phase since the last reset: drawing proposals, evaluating
 * them, selecting the first accepted one, applying it (move + DP re-scan).  Profiling aid. */
vnsgpu_status vnsgpu_chain_cycles(vnsgpu_ctx *ctx, int64_t *cycles4, int reset);
/* pathInsert (kind 1, :931-999) / pathExchange (kind 2, :1001-1089) incl. fitDOPtoBudget */
vnsgpu_status vnsgpu_shake(vnsgpu_ctx *ctx, int path, int kind, double budget, vnsgpu_rng *rng);
/* fitDOPtoBudget (:364-385) */
vnsgpu_status vnsgpu_fit_to_budget(vnsgpu_ctx *ctx, int path, double budget);
/* greedy_insertion (:1091-1155) */
vnsgpu_status vnsgpu_greedy(vnsgpu_ctx *ctx, int path, double budget);
/* VnsSopSolver::iterate (:82-260) as ONE persistent device chain; `path` enters as any state with the
 * VNS vector set (normally T = 0) and leaves as the best tour.  impr_* receive the improvement log
 * (reward, length, ms) up to params->max_improvements entries; may be NULL */
vnsgpu_status vnsgpu_vns_run(vnsgpu_ctx *ctx, int path, const vnsgpu_vns_params *params,
                             vnsgpu_vns_result *result, double *impr_reward, double *impr_length,
                             double *impr_time_ms);

/* Multi-start = the reference's --batch=N (N independent solves with different seeds, crl
 * algorithm.cc:120-161) as ONE launch: one thread block per run, each a trajectory-exact replica of a
 * seeded reference run.  Every run starts from `path`'s state; results[r] belongs to seeds[r]; the best
 * tour (max reward, then min length) is copied into `path` and its index returned in *best_run.
 * tour_paths (may be NULL): nruns handles that receive every run's final tour; the caller destroys them. */
vnsgpu_status vnsgpu_vns_run_batch(vnsgpu_ctx *ctx, int path, int nruns, const uint32_t *seeds, const vnsgpu_vns_params *params,
                                   vnsgpu_vns_result *results, int *best_run, int *tour_paths);
/* Race mode of vnsgpu_vns_run_batch ("wall time to a given reward"): with target > 0 the first run whose tour reaches
 * reward >= target raises a device flag and every run returns at its next shake, so the launch ends as soon as ANY seed
 * got there (results[r].time_ms of that run = time to the target).  Each run is still a trajectory-exact prefix of the
 * seeded reference run.  target <= 0 (default): off, only the reference's stopping rules (vns_sop_solver.cpp:178-193). */
vnsgpu_status vnsgpu_set_target_reward(vnsgpu_ctx *ctx, double target);

#ifdef __cplusplus
}
#endif
#endif /* VNSGPU_H */
