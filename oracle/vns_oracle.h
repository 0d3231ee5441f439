/* TEST INFRASTRUCTURE ONLY - never linked, imported or executed by the product path.
 *
 * Plain-C restatement of the VNS-SOP hot path of ctu-mrs/vns-sop on the flat data model the
 * CUDA engine uses (cluster-contiguous samples, block-major distance table, padded DP rows).
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference/sources).  Parity is PINNED: tests/test_oracle_vs_ref.py checks every function
 * against the reference's own code compiled into oracle/_ref/liboracle_ref.so, and
 * tests/test_oracle_golden.py against vectors that library produced (tests/golden/), incl. the
 * 98 784 Dubins matrix entries shipped in datasets/dop_sop_dataset.
 */
#ifndef VNS_ORACLE_H
#define VNS_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VO_M 1.7976931348623157e308 /* numeric_limits<double>::max(), vns_sop_path.cpp:10 */

enum { VO_RULE_RAW = 0, VO_RULE_ROUND = 1, VO_RULE_CEIL = 2 };
enum { VO_ADD = 0, VO_REPLACE = 1, VO_REMOVE = 2, VO_MOVE = 3, VO_EXCHANGE = 4 };

typedef struct {
   int C;          /* clusters after the loader's renumbering (start = 0 ... goal = C-1 typ.) */
   int N;          /* sum of m[c] */
   int mp;         /* max m[c]: stride of DP rows */
   int start, goal;
   int *m, *off;   /* off[c] = first sample of cluster c; off[C] = N */
   double *reward; /* per cluster (allSamples[c][0].reward) */
   double *D;      /* block-major: block(c1,c2) at off[c1]*N + m[c1]*off[c2], row-major m1 x m2 */
   double *dmin, *dmax; /* C x C (vns_sop_solver.cpp:262-297) */
} vo_inst;

typedef struct {
   const vo_inst *I;
   int T;        /* targets in path */
   int *targets; /* capacity C (VnsSopPath::targets) */
   int nvns;     /* length of the solver's VNS vector; at call boundaries targets == vns[0..T) */
   int *vns;     /* capacity C (actualVNS / vnsVector of vns_sop_solver.cpp) */
   double *F, *B; /* (T+2) rows used, capacity (C+2) rows, stride mp */
   int *pF, *pB;  /* back pointers (idClusterNode) */
} vo_path;

typedef struct { /* glibc random() TYPE_3 state, stdlib/random_r.c */
   int32_t r[34];
   int f, b;
} vo_rng;

typedef struct {
   long proposals, lb_skips, evals, accepts; /* per neighbourhood call */
} vo_ls_stats;

/* instance + distance table (a1-a5) */
vo_inst *vo_inst_new(int C, const int *m, const double *reward, int start, int goal);
void vo_inst_free(vo_inst *I);
static inline long vo_block(const vo_inst *I, int c1, int c2)
{
   return (long)I->off[c1] * I->N + (long)I->m[c1] * I->off[c2];
}
void vo_table_euclid(vo_inst *I, const double *xy, int rule);
void vo_table_dense(vo_inst *I, const double *W);
void vo_table_dubins(vo_inst *I, const double *centers, int S, double rho, int do_ceil);
extern double vo_dubins_eps; /* CSC words with p^2 >= -eps are kept (p^2 clamped to 0); default 1e-9 */
double vo_dubins_length(double x0, double y0, double th0, double x1, double y1, double th1, double rho);
void vo_bounds(vo_inst *I);
void vo_get_dense(const vo_inst *I, double *W);

/* path evaluator (a6-a14) */
vo_path *vo_path_new(const vo_inst *I);
void vo_path_free(vo_path *p);
void vo_path_copy(vo_path *dst, const vo_path *src);
void vo_path_set(vo_path *p, const int *vns, int nvns, int T);
void vo_update(vo_path *p);
void vo_add_point(vo_path *p, int c, int idx);
void vo_add_points(vo_path *p, const int *cs, int n, int idx);
void vo_remove_point(vo_path *p, int idx);
void vo_remove_range(vo_path *p, int a, int b);
double vo_try_add(const vo_path *p, int c, int idx);
double vo_try_add_lb(const vo_path *p, int c, int idx);
double vo_try_replace(const vo_path *p, int c, int idx);
double vo_try_replace_lb(const vo_path *p, int c, int idx);
double vo_try_remove(const vo_path *p, int idx);
double vo_try_move(const vo_path *p, int what, int where);
double vo_try_move_lb(const vo_path *p, int what, int where);
double vo_try_exchange(const vo_path *p, int i1, int i2);
double vo_try_exchange_lb(const vo_path *p, int i1, int i2);
double vo_length(const vo_path *p);
double vo_length_rear(const vo_path *p);
double vo_length_calc(const vo_path *p);
double vo_reward(const vo_path *p);
int vo_get_path(const vo_path *p, int *cluster, int *node);
void vo_eval_batch(const vo_path *p, int n, const int *cands, double *len, double *lb);

/* RNG (a16) */
void vo_srand(vo_rng *g, unsigned seed);
int vo_rand(vo_rng *g);
int vo_rand_int(vo_rng *g, int lo, int hi);

/* neighbourhood drivers (a15); vns is p->vns */
void vo_insert_random(vo_path *p, int num_changes, double budget, int skip, vo_rng *g, vo_ls_stats *st);
void vo_exchange_random(vo_path *p, int num_changes, double budget, int skip, vo_rng *g, vo_ls_stats *st);
void vo_fit(vo_path *p, double budget);
void vo_path_insert(vo_path *p, double budget, vo_rng *g);
void vo_path_exchange(vo_path *p, double budget, vo_rng *g);
void vo_greedy(vo_path *p, double budget);

typedef struct {
   int num_iterations, num_iterations_unimproved, skip, greedy;
   double budget;
   unsigned seed;
} vo_params;
typedef struct {
   int num_iters, num_iters_last_impr, n_improvements;
   double reward, length;
   double impr_reward[4096], impr_length[4096];
} vo_result;
/* VnsSopSolver::iterate (vns_sop_solver.cpp:82-260); best path left in p */
void vo_solve(vo_path *p, const vo_params *prm, vo_rng *g, vo_result *res);

#ifdef __cplusplus
}
#endif
#endif
