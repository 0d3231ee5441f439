/* TEST INFRASTRUCTURE ONLY - see vns_oracle.h.  Plain C99, single thread, libm only. */
#include "vns_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define TWO_PI 6.283185307179586476925286766559
/* a CSC word whose straight segment has length 0 is feasible; rounding in sin/cos can push p^2 a few
 * ulps below 0 there (any libm), so p^2 in [-eps, 0) is clamped to 0 instead of dropping the word */
#define VNS_DUBINS_EPS vo_dubins_eps
double vo_dubins_eps = 1e-9; /* tests set -1e-9 to get the "word dropped" alternative of knife-edge entries */

/* ------------------------------------------------------------------ instance + tables ---- */

vo_inst *vo_inst_new(int C, const int *m, const double *reward, int start, int goal)
{
   vo_inst *I = (vo_inst *)calloc(1, sizeof(vo_inst));
   I->C = C;
   I->start = start;
   I->goal = goal;
   I->m = (int *)malloc(sizeof(int) * C);
   I->off = (int *)malloc(sizeof(int) * (C + 1));
   I->reward = (double *)malloc(sizeof(double) * C);
   I->off[0] = 0;
   for (int c = 0; c < C; ++c) {
      I->m[c] = m[c];
      I->reward[c] = reward[c];
      I->off[c + 1] = I->off[c] + m[c];
      if (m[c] > I->mp) I->mp = m[c];
   }
   I->N = I->off[C];
   I->D = (double *)malloc(sizeof(double) * (size_t)I->N * I->N);
   I->dmin = (double *)malloc(sizeof(double) * (size_t)C * C);
   I->dmax = (double *)malloc(sizeof(double) * (size_t)C * C);
   return I;
}

void vo_inst_free(vo_inst *I)
{
   if (!I) return;
   free(I->m); free(I->off); free(I->reward); free(I->D); free(I->dmin); free(I->dmax);
   free(I);
}

/* SOPLoader::getSOPFromNodes sop_loader.cpp:457-491 (+ GraphNode::distanceTo heuristic_types.h:67-75)
 * and getOP/getOPN sop_loader.cpp:278,360 (raw rule).  xy = N x 2, cluster-contiguous samples. */
void vo_table_euclid(vo_inst *I, const double *xy, int rule)
{
   for (int c1 = 0; c1 < I->C; ++c1)
      for (int c2 = 0; c2 < I->C; ++c2) {
         double *blk = I->D + vo_block(I, c1, c2);
         for (int i = 0; i < I->m[c1]; ++i)
            for (int j = 0; j < I->m[c2]; ++j) {
               const double *a = xy + 2 * (I->off[c1] + i), *b = xy + 2 * (I->off[c2] + j);
               volatile double dx = b[0] - a[0], dy = b[1] - a[1];
               volatile double sx = dx * dx, sy = dy * dy; /* no FMA contraction */
               double d = sqrt(sx + sy);
               if (rule == VO_RULE_ROUND) d = round(d);
               else if (rule == VO_RULE_CEIL) d = ceil(d);
               blk[i * I->m[c2] + j] = d;
            }
      }
}

/* SOPLoader::getSOPFromWeightMatrix sop_loader.cpp:413-427; W = N x N row-major in sample order */
void vo_table_dense(vo_inst *I, const double *W)
{
   for (int c1 = 0; c1 < I->C; ++c1)
      for (int c2 = 0; c2 < I->C; ++c2) {
         double *blk = I->D + vo_block(I, c1, c2);
         for (int i = 0; i < I->m[c1]; ++i)
            for (int j = 0; j < I->m[c2]; ++j)
               blk[i * I->m[c2] + j] = W[(size_t)(I->off[c1] + i) * I->N + I->off[c2] + j];
      }
}

void vo_get_dense(const vo_inst *I, double *W)
{
   for (int c1 = 0; c1 < I->C; ++c1)
      for (int c2 = 0; c2 < I->C; ++c2) {
         const double *blk = I->D + vo_block(I, c1, c2);
         for (int i = 0; i < I->m[c1]; ++i)
            for (int j = 0; j < I->m[c2]; ++j)
               W[(size_t)(I->off[c1] + i) * I->N + I->off[c2] + j] = blk[i * I->m[c2] + j];
      }
}

static double mod2pi(double t) { return t - TWO_PI * floor(t / TWO_PI); }

/* Shortest Dubins path length.  No Dubins code exists in the reference tree;
 * this restates the standard normalised six-word formulation whose ceil() reproduces the shipped
 * matrices datasets/dop_sop_dataset/ *.sop. */
double vo_dubins_length(double x0, double y0, double th0, double x1, double y1, double th1, double rho)
{
   double dx = x1 - x0, dy = y1 - y0;
   double d = sqrt(dx * dx + dy * dy) / rho;
   double phi = d > 0 ? atan2(dy, dx) : 0.0;
   double al = mod2pi(th0 - phi), be = mod2pi(th1 - phi);
   double sa = sin(al), sb = sin(be), ca = cos(al), cb = cos(be), cab = cos(al - be);
   double best = INFINITY, psq, p, tau, u, t;
   /* LSL */
   psq = 2 + d * d - 2 * cab + 2 * d * (sa - sb);
   if (psq >= -VNS_DUBINS_EPS) {
      psq = psq > 0 ? psq : 0;
      tau = atan2(cb - ca, d + sa - sb);
      t = mod2pi(tau - al) + sqrt(psq) + mod2pi(be - tau);
      if (t < best) best = t;
   }
   /* RSR */
   psq = 2 + d * d - 2 * cab + 2 * d * (sb - sa);
   if (psq >= -VNS_DUBINS_EPS) {
      psq = psq > 0 ? psq : 0;
      tau = atan2(ca - cb, d - sa + sb);
      t = mod2pi(al - tau) + sqrt(psq) + mod2pi(tau - be);
      if (t < best) best = t;
   }
   /* LSR */
   psq = -2 + d * d + 2 * cab + 2 * d * (sa + sb);
   if (psq >= -VNS_DUBINS_EPS) {
      psq = psq > 0 ? psq : 0;
      p = sqrt(psq);
      tau = atan2(-ca - cb, d + sa + sb) - atan2(-2.0, p);
      t = mod2pi(tau - al) + p + mod2pi(tau - mod2pi(be));
      if (t < best) best = t;
   }
   /* RSL */
   psq = -2 + d * d + 2 * cab - 2 * d * (sa + sb);
   if (psq >= -VNS_DUBINS_EPS) {
      psq = psq > 0 ? psq : 0;
      p = sqrt(psq);
      tau = atan2(ca + cb, d - sa - sb) - atan2(2.0, p);
      t = mod2pi(al - tau) + p + mod2pi(be - tau);
      if (t < best) best = t;
   }
   /* RLR */
   u = (6 - d * d + 2 * cab + 2 * d * (sa - sb)) / 8;
   if (fabs(u) <= 1) {
      p = mod2pi(TWO_PI - acos(u));
      t = mod2pi(al - atan2(ca - cb, d - sa + sb) + p / 2);
      t = t + p + mod2pi(al - be - t + p);
      if (t < best) best = t;
   }
   /* LRL */
   u = (6 - d * d + 2 * cab + 2 * d * (sb - sa)) / 8;
   if (fabs(u) <= 1) {
      p = mod2pi(TWO_PI - acos(u));
      t = mod2pi(-al - atan2(ca - cb, d + sa - sb) + p / 2);
      t = t + p + mod2pi(mod2pi(be) - al - t + mod2pi(p));
      if (t < best) best = t;
   }
   return best * rho;
}

/* DOP table: every cluster carries S heading samples theta_k = 2 pi k / S at its centre;
 * entry = ceil(L).  centers = C x 2. */
void vo_table_dubins(vo_inst *I, const double *centers, int S, double rho, int do_ceil)
{
   for (int c1 = 0; c1 < I->C; ++c1)
      for (int c2 = 0; c2 < I->C; ++c2) {
         double *blk = I->D + vo_block(I, c1, c2);
         for (int i = 0; i < I->m[c1]; ++i)
            for (int j = 0; j < I->m[c2]; ++j) {
               double L = vo_dubins_length(centers[2 * c1], centers[2 * c1 + 1], TWO_PI * i / S,
                                           centers[2 * c2], centers[2 * c2 + 1], TWO_PI * j / S, rho);
               blk[i * I->m[c2] + j] = do_ceil ? ceil(L) : L;
            }
      }
}

/* VnsSopSolver::calcClusterDistancesBounds vns_sop_solver.cpp:262-297 */
void vo_bounds(vo_inst *I)
{
   for (int c1 = 0; c1 < I->C; ++c1)
      for (int c2 = 0; c2 < I->C; ++c2) {
         const double *blk = I->D + vo_block(I, c1, c2);
         double lo = VO_M, hi = -VO_M;
         for (int k = 0; k < I->m[c1] * I->m[c2]; ++k) {
            if (blk[k] < lo) lo = blk[k];
            if (blk[k] > hi) hi = blk[k];
         }
         I->dmin[(size_t)c1 * I->C + c2] = lo;
         I->dmax[(size_t)c1 * I->C + c2] = hi;
      }
}

/* ------------------------------------------------------------------ path evaluator ------- */
/* Layers l = 0..S, S = T+1: layer 0 = start, layer l = targets[l-1], layer S = goal
 * (VnsSopPath::update vns_sop_path.cpp:1327-1336). */

/* row capacity: the shakes add a range before removing one, so T can transiently reach 2(C-2) */
#define VO_ROWS(I) ((size_t)2 * (I)->C + 2)

static inline int lc(const vo_path *p, int l)
{
   return l == 0 ? p->I->start : (l == p->T + 1 ? p->I->goal : p->targets[l - 1]);
}
static inline const double *blk(const vo_path *p, int ca, int cb) { return p->I->D + vo_block(p->I, ca, cb); }

vo_path *vo_path_new(const vo_inst *I)
{
   vo_path *p = (vo_path *)calloc(1, sizeof(vo_path));
   size_t rows = VO_ROWS(I);
   p->I = I;
   p->targets = (int *)calloc(rows, sizeof(int));
   p->vns = (int *)calloc(rows, sizeof(int));
   p->F = (double *)calloc(rows * I->mp, sizeof(double));
   p->B = (double *)calloc(rows * I->mp, sizeof(double));
   p->pF = (int *)calloc(rows * I->mp, sizeof(int));
   p->pB = (int *)calloc(rows * I->mp, sizeof(int));
   vo_update(p);
   return p;
}

void vo_path_free(vo_path *p)
{
   if (!p) return;
   free(p->targets); free(p->vns); free(p->F); free(p->B); free(p->pF); free(p->pB);
   free(p);
}

void vo_path_copy(vo_path *d, const vo_path *s)
{
   const vo_inst *I = s->I;
   size_t rows = VO_ROWS(I);
   d->I = I; d->T = s->T; d->nvns = s->nvns;
   memcpy(d->targets, s->targets, sizeof(int) * rows);
   memcpy(d->vns, s->vns, sizeof(int) * rows);
   memcpy(d->F, s->F, sizeof(double) * rows * I->mp);
   memcpy(d->B, s->B, sizeof(double) * rows * I->mp);
   memcpy(d->pF, s->pF, sizeof(int) * rows * I->mp);
   memcpy(d->pB, s->pB, sizeof(int) * rows * I->mp);
}

void vo_path_set(vo_path *p, const int *vns, int nvns, int T)
{
   p->nvns = nvns; p->T = T;
   memcpy(p->vns, vns, sizeof(int) * nvns);
   memcpy(p->targets, vns, sizeof(int) * T);
   vo_update(p);
}

/* one forward layer: F[l] = F[l-1] (x) D(c_{l-1}, c_l); strict '<' with the predecessor in the
 * outer loop => smallest predecessor index wins ties (vns_sop_path.cpp:1353-1373,1426-1448) */
static void fwd_layer(vo_path *p, int l)
{
   const vo_inst *I = p->I;
   int ca = lc(p, l - 1), cb = lc(p, l), ma = I->m[ca], mb = I->m[cb];
   const double *d = blk(p, ca, cb);
   double *Fp = p->F + (size_t)(l - 1) * I->mp, *Fl = p->F + (size_t)l * I->mp;
   int *pl = p->pF + (size_t)l * I->mp;
   for (int j = 0; j < mb; ++j) { Fl[j] = VO_M; pl[j] = -1; }
   for (int i = 0; i < ma; ++i)
      for (int j = 0; j < mb; ++j) {
         double n = Fp[i] + d[i * mb + j];
         if (n < Fl[j]) { Fl[j] = n; pl[j] = i; }
      }
}

/* one backward layer: B[l][i] = min_j B[l+1][j] + D(c_l, c_{l+1})[i][j], smallest successor
 * index wins ties (vns_sop_path.cpp:1380-1399) */
static void bwd_layer(vo_path *p, int l)
{
   const vo_inst *I = p->I;
   int ca = lc(p, l), cb = lc(p, l + 1), ma = I->m[ca], mb = I->m[cb];
   const double *d = blk(p, ca, cb);
   double *Bn = p->B + (size_t)(l + 1) * I->mp, *Bl = p->B + (size_t)l * I->mp;
   int *pl = p->pB + (size_t)l * I->mp;
   for (int i = 0; i < ma; ++i) {
      Bl[i] = VO_M; pl[i] = -1;
      for (int j = 0; j < mb; ++j) {
         double n = Bn[j] + d[i * mb + j];
         if (n < Bl[i]) { Bl[i] = n; pl[i] = j; }
      }
   }
}

/* VnsSopPath::update vns_sop_path.cpp:1321-1404 */
void vo_update(vo_path *p)
{
   const vo_inst *I = p->I;
   int S = p->T + 1;
   for (int j = 0; j < I->m[lc(p, 0)]; ++j) { p->F[j] = 0; p->pF[j] = -1; }
   for (int l = 1; l <= S; ++l) fwd_layer(p, l);
   for (int j = 0; j < I->m[lc(p, S)]; ++j) { p->B[(size_t)S * I->mp + j] = 0; p->pB[(size_t)S * I->mp + j] = -1; }
   for (int l = S - 1; l >= 0; --l) bwd_layer(p, l);
}

static void shift_rows(vo_path *p, int from_row, int old_rows, int delta)
{ /* B rows [from_row, old_rows) move by delta rows (F rows there are recomputed anyway) */
   size_t mp = p->I->mp;
   memmove(p->B + (size_t)(from_row + delta) * mp, p->B + (size_t)from_row * mp, sizeof(double) * mp * (old_rows - from_row));
   memmove(p->pB + (size_t)(from_row + delta) * mp, p->pB + (size_t)from_row * mp, sizeof(int) * mp * (old_rows - from_row));
}

/* addPoint(vector,idx) + updateAfterInsert vns_sop_path.cpp:105-136,1406-1476 */
void vo_add_points(vo_path *p, const int *cs, int n, int idx)
{
   int oldS = p->T + 1;
   memmove(p->targets + idx + n, p->targets + idx, sizeof(int) * (p->T - idx));
   memcpy(p->targets + idx, cs, sizeof(int) * n);
   p->T += n;
   int S = p->T + 1;
   shift_rows(p, idx + 1, oldS + 1, n);
   for (int l = idx + 1; l <= S; ++l) fwd_layer(p, l);          /* :1426 */
   for (int l = idx + n; l >= 0; --l) bwd_layer(p, l);          /* :1453 (idxEnd+1 = idx+n) */
}
void vo_add_point(vo_path *p, int c, int idx) { vo_add_points(p, &c, 1, idx); } /* :67-103 */

/* removePoint(a,b) + updateAfterRemove vns_sop_path.cpp:182-205,1478-1540 */
void vo_remove_range(vo_path *p, int a, int b)
{
   int n = b - a + 1, oldS = p->T + 1;
   memmove(p->targets + a, p->targets + b + 1, sizeof(int) * (p->T - b - 1));
   p->T -= n;
   int S = p->T + 1;
   shift_rows(p, b + 2, oldS + 1, -n);
   for (int l = a + 1; l <= S; ++l) fwd_layer(p, l);            /* :1488 */
   for (int l = a; l >= 0; --l) bwd_layer(p, l);                /* :1515 */
}
void vo_remove_point(vo_path *p, int idx) { vo_remove_range(p, idx, idx); } /* :154-180 */

static double row_min(const double *r, int m)
{
   double v = VO_M;
   for (int i = 0; i < m; ++i) if (r[i] < v) v = r[i];
   return v;
}
static inline const double *Frow(const vo_path *p, int l) { return p->F + (size_t)l * p->I->mp; }
static inline const double *Brow(const vo_path *p, int l) { return p->B + (size_t)l * p->I->mp; }
static inline double minF(const vo_path *p, int l) { return row_min(Frow(p, l), p->I->m[lc(p, l)]); }
static inline double minB(const vo_path *p, int l) { return row_min(Brow(p, l), p->I->m[lc(p, l)]); }
static inline double dmn(const vo_path *p, int ca, int cb) { return p->I->dmin[(size_t)ca * p->I->C + cb]; }

double vo_length(const vo_path *p) { return minF(p, p->T + 1); }       /* :1660-1678 */
double vo_length_rear(const vo_path *p) { return minB(p, 0); }         /* :1680-1698 */

/* out = v (x) D(ca,cb):  out[j] = min_i (v[i] + d[i][j])  (every "pathPartShortest" step) */
static void minplus(const vo_path *p, const double *v, int ca, int cb, double *out)
{
   int ma = p->I->m[ca], mb = p->I->m[cb];
   const double *d = blk(p, ca, cb);
   for (int j = 0; j < mb; ++j) out[j] = VO_M;
   for (int i = 0; i < ma; ++i)
      for (int j = 0; j < mb; ++j) {
         double n = v[i] + d[i * mb + j];
         if (n < out[j]) out[j] = n;
      }
}
/* closing block with association (d + B) + v: tryToAdd :295, tryToReplace :504, tryToRemove :526 */
static double close_dBv(const vo_path *p, const double *v, int ca, int cb, const double *Bt)
{
   int ma = p->I->m[ca], mb = p->I->m[cb];
   const double *d = blk(p, ca, cb);
   double best = VO_M;
   for (int i = 0; i < ma; ++i)
      for (int j = 0; j < mb; ++j) {
         double n = d[i * mb + j] + Bt[j] + v[i];
         if (n < best) best = n;
      }
   return best;
}
/* closing block with association (v + d) + B: tryToMove :1066,:1145, tryToExchange :789 */
static double close_vdB(const vo_path *p, const double *v, int ca, int cb, const double *Bt)
{
   int ma = p->I->m[ca], mb = p->I->m[cb];
   const double *d = blk(p, ca, cb);
   double best = VO_M;
   for (int i = 0; i < ma; ++i)
      for (int j = 0; j < mb; ++j) {
         double n = v[i] + d[i * mb + j] + Bt[j];
         if (n < best) best = n;
      }
   return best;
}

/* tryToAdd(int,int) vns_sop_path.cpp:245-304 */
double vo_try_add(const vo_path *p, int c, int idx)
{
   double act[p->I->mp];
   minplus(p, Frow(p, idx), lc(p, idx), c, act);
   return close_dBv(p, act, c, lc(p, idx + 1), Brow(p, idx + 1));
}
/* tryToAddLowerBound :207-243 */
double vo_try_add_lb(const vo_path *p, int c, int idx)
{
   double add = dmn(p, lc(p, idx), c) + dmn(p, c, lc(p, idx + 1));
   return minF(p, idx) + add + minB(p, idx + 1);
}
/* tryToReplace(int,int,1) :456-513 */
double vo_try_replace(const vo_path *p, int c, int idx)
{
   double nw[p->I->mp];
   minplus(p, Frow(p, idx), lc(p, idx), c, nw);
   return close_dBv(p, nw, c, lc(p, idx + 2), Brow(p, idx + 2));
}
/* tryToReplaceLowerBound :423-454 */
double vo_try_replace_lb(const vo_path *p, int c, int idx)
{
   double between = dmn(p, lc(p, idx), c) + dmn(p, c, lc(p, idx + 2));
   return minF(p, idx) + between + minB(p, idx + 2);
}
/* tryToRemove(int) :515-535 */
double vo_try_remove(const vo_path *p, int idx)
{
   return close_dBv(p, Frow(p, idx), lc(p, idx), lc(p, idx + 2), Brow(p, idx + 2));
}

/* tryToMove :932-1154 */
double vo_try_move(const vo_path *p, int what, int where)
{
   double va[p->I->mp], vb[p->I->mp], *v = va, *w = vb, *t;
   int lo = what < where ? what : where, hi = what < where ? where : what;
   int i = lo + 1, j = hi + 1;
   if (where > what) { /* up: i-1 -> i+1 -> ... -> j-1 -> i -> j   (:988-1071) */
      minplus(p, Frow(p, i - 1), lc(p, i - 1), lc(p, i + 1), v);
      for (int l = i + 1; l < j - 1; ++l) { minplus(p, v, lc(p, l), lc(p, l + 1), w); t = v; v = w; w = t; }
      minplus(p, v, lc(p, j - 1), lc(p, i), w);
      return close_vdB(p, w, lc(p, i), lc(p, j), Brow(p, j));
   }
   /* down: i-1 -> j -> i -> ... -> j-1 -> j+1   (:1074-1151) */
   minplus(p, Frow(p, i - 1), lc(p, i - 1), lc(p, j), v);
   minplus(p, v, lc(p, j), lc(p, i), w); t = v; v = w; w = t;
   for (int l = i; l < j - 1; ++l) { minplus(p, v, lc(p, l), lc(p, l + 1), w); t = v; v = w; w = t; }
   return close_vdB(p, v, lc(p, j - 1), lc(p, j + 1), Brow(p, j + 1));
}

/* tryToMoveLowerBound :799-930 (association of dist_between exactly as written) */
double vo_try_move_lb(const vo_path *p, int what, int where)
{
   double between = 0;
   if (where > what) {
      int i = what + 1, j = where + 1;
      between += dmn(p, lc(p, i - 1), lc(p, i + 1));
      for (int l = i + 2; l <= j - 1; ++l) between += dmn(p, lc(p, l - 1), lc(p, l));
      between += dmn(p, lc(p, j - 1), lc(p, i)) + dmn(p, lc(p, i), lc(p, j));
      return minF(p, i - 1) + between + minB(p, j);
   }
   int j = what + 1, i = where + 1;
   between += dmn(p, lc(p, i - 1), lc(p, j)) + dmn(p, lc(p, j), lc(p, i));
   for (int l = i + 1; l <= j - 1; ++l) between += dmn(p, lc(p, l - 1), lc(p, l));
   between += dmn(p, lc(p, j - 1), lc(p, j + 1));
   return minF(p, i - 1) + between + minB(p, j + 1);
}

/* tryToExchange :635-797: i -> j -> i+2 -> ... -> j-1 -> i+1 -> j+1 */
double vo_try_exchange(const vo_path *p, int i1, int i2)
{
   double va[p->I->mp], vb[p->I->mp], *v = va, *w = vb, *t;
   int i = i1 < i2 ? i1 : i2, j = (i1 < i2 ? i2 : i1) + 1;
   minplus(p, Frow(p, i), lc(p, i), lc(p, j), v);
   minplus(p, v, lc(p, j), lc(p, i + 2), w); t = v; v = w; w = t;
   for (int l = i + 2; l <= j - 2; ++l) { minplus(p, v, lc(p, l), lc(p, l + 1), w); t = v; v = w; w = t; }
   minplus(p, v, lc(p, j - 1), lc(p, i + 1), w);
   return close_vdB(p, w, lc(p, i + 1), lc(p, j + 1), Brow(p, j + 1));
}

/* tryToExchangeLowerBound :573-633 */
double vo_try_exchange_lb(const vo_path *p, int i1, int i2)
{
   int i = i1 < i2 ? i1 : i2, j = (i1 < i2 ? i2 : i1) + 1;
   double between = 0;
   between += dmn(p, lc(p, i), lc(p, j));
   between += dmn(p, lc(p, j), lc(p, i + 2));
   for (int l = i + 3; l <= j - 1; ++l) between += dmn(p, lc(p, l - 1), lc(p, l));
   between += dmn(p, lc(p, j - 1), lc(p, i + 1));
   between += dmn(p, lc(p, i + 1), lc(p, j + 1));
   return minF(p, i) + between + minB(p, j + 1);
}

/* getReward :2016-2031 */
double vo_reward(const vo_path *p)
{
   const vo_inst *I = p->I;
   double r = 0;
   for (int t = 0; t < p->T; ++t) r += I->reward[p->targets[t]];
   if (I->reward[I->start] > 0) {
      r += I->reward[I->start];
      if (I->start != I->goal) r += I->reward[I->goal];
   }
   return r;
}

/* getPath :1542-1599: arg-min at the goal layer (first index on ties), follow F back pointers */
int vo_get_path(const vo_path *p, int *cluster, int *node)
{
   int S = p->T + 1, best = 0;
   double mn = VO_M;
   const double *Fs = Frow(p, S);
   for (int i = 0; i < p->I->m[lc(p, S)]; ++i) if (Fs[i] < mn) { mn = Fs[i]; best = i; }
   for (int l = S; l >= 0; --l) {
      cluster[l] = lc(p, l);
      node[l] = best;
      if (l > 0) best = p->pF[(size_t)l * p->I->mp + best];
   }
   return S + 1;
}

/* getPathLengthCalculate :1700-1741: re-sum raw distances along the back pointers, goal -> start */
double vo_length_calc(const vo_path *p)
{
   int n = p->T + 2, cl[n], nd[n];
   vo_get_path(p, cl, nd);
   double sum = 0;
   for (int l = n - 1; l > 0; --l)
      sum += blk(p, cl[l - 1], cl[l])[nd[l - 1] * p->I->m[cl[l]] + nd[l]];
   return sum;
}

void vo_eval_batch(const vo_path *p, int n, const int *cands, double *len, double *lb)
{
   for (int k = 0; k < n; ++k) {
      int o = cands[3 * k], a = cands[3 * k + 1], b = cands[3 * k + 2];
      double l = 0, bound = 0;
      switch (o) {
      case VO_ADD: bound = vo_try_add_lb(p, a, b); l = vo_try_add(p, a, b); break;
      case VO_REPLACE: bound = vo_try_replace_lb(p, a, b); l = vo_try_replace(p, a, b); break;
      case VO_REMOVE: l = vo_try_remove(p, a); bound = l; break;
      case VO_MOVE: bound = vo_try_move_lb(p, a, b); l = vo_try_move(p, a, b); break;
      case VO_EXCHANGE: bound = vo_try_exchange_lb(p, a, b); l = vo_try_exchange(p, a, b); break;
      }
      if (len) len[k] = l;
      if (lb) lb[k] = bound;
   }
}

/* ------------------------------------------------------------------ RNG ------------------ */
/* glibc srand()/rand() = srandom_r/random_r TYPE_3 (degree 31, separation 3); the reference
 * draws through randIntMinMax math_common.cpp:21-28 */
void vo_srand(vo_rng *g, unsigned seed)
{
   int32_t word = seed == 0 ? 1 : (int32_t)seed;
   g->r[0] = word;
   for (int i = 1; i < 31; ++i) {
      long hi = word / 127773, lo = word % 127773;
      long w = 16807 * lo - 2836 * hi;
      if (w < 0) w += 2147483647;
      word = (int32_t)w;
      g->r[i] = word;
   }
   g->f = 3;
   g->b = 0;
   for (int i = 0; i < 310; ++i) vo_rand(g);
}

int vo_rand(vo_rng *g)
{
   uint32_t v = (uint32_t)g->r[g->f] + (uint32_t)g->r[g->b];
   g->r[g->f] = (int32_t)v;
   g->f = g->f + 1 == 31 ? 0 : g->f + 1;
   g->b = g->b + 1 == 31 ? 0 : g->b + 1;
   return (int)(v >> 1);
}

int vo_rand_int(vo_rng *g, int lo, int hi)
{
   return lo + (int)(((double)vo_rand(g) / ((double)2147483647 + 1)) * (hi - lo + 1));
}

/* ------------------------------------------------------------------ neighbourhood drivers - */
#define MIN_CHANGE_EPS 0.00001 /* vns_sop_solver.cpp:15 */

static void vec_insert(int *v, int *n, int pos, const int *src, int k)
{
   memmove(v + pos + k, v + pos, sizeof(int) * (*n - pos));
   memcpy(v + pos, src, sizeof(int) * k);
   *n += k;
}
static void vec_erase(int *v, int *n, int a, int b_incl)
{
   memmove(v + a, v + b_incl + 1, sizeof(int) * (*n - b_incl - 1));
   *n -= b_incl - a + 1;
}

/* VnsSopSolver::insertRandom vns_sop_solver.cpp:469-684 */
void vo_insert_random(vo_path *p, int num_changes, double budget, int skip, vo_rng *g, vo_ls_stats *st)
{
   const vo_inst *I = p->I;
   double L = vo_length(p);
   if (p->nvns < 3) return;                                             /* :498 */
   for (int var = 0; var < num_changes; ++var) {
      int from = vo_rand_int(g, 0, p->nvns - 1);                        /* :507 */
      int to = vo_rand_int(g, 0, p->T);                                 /* :508 */
      while (abs(from - to) <= 1) {                                     /* :510-519 */
         from = vo_rand_int(g, 0, p->nvns - 1);
         to = vo_rand_int(g, 0, p->T);
      }
      if (st) st->proposals++;
      int c = p->vns[from];
      double len, dR = 0;
      if (from < p->T) {                                                /* :534-565 */
         double lb = vo_try_move_lb(p, from, to);
         if (skip && lb > budget) { if (st) st->lb_skips++; continue; }
         len = vo_try_move(p, from, to);
      } else {                                                          /* :567-600 */
         double removed = L - L;                                        /* :526,531 */
         double lb = vo_try_add_lb(p, c, to) - removed;
         if (skip && lb > budget) { if (st) st->lb_skips++; continue; }
         double after_add = vo_try_add(p, c, to);
         double added = after_add - L;                                  /* :580 */
         dR = I->reward[c];                                             /* :581 */
         len = L - removed + added;                                     /* :583 */
      }
      if (st) st->evals++;
      if (len <= budget && (dR > 0 || (L - len) > MIN_CHANGE_EPS)) {    /* :619 */
         if (st) st->accepts++;
         if (to > from) {                                               /* :625-635 */
            if (to <= p->T) vo_add_point(p, c, to);
            if (from < p->T) vo_remove_point(p, from);
            vec_insert(p->vns, &p->nvns, to, &c, 1);
            vec_erase(p->vns, &p->nvns, from, from);
         } else {                                                       /* :636-646 */
            if (from < p->T) vo_remove_point(p, from);
            if (to <= p->T) vo_add_point(p, c, to);
            vec_erase(p->vns, &p->nvns, from, from);
            vec_insert(p->vns, &p->nvns, to, &c, 1);
         }
         L = vo_length(p);                                              /* :649 */
      }
   }
}

/* VnsSopSolver::exchangeRandom vns_sop_solver.cpp:686-864 */
void vo_exchange_random(vo_path *p, int num_changes, double budget, int skip, vo_rng *g, vo_ls_stats *st)
{
   const vo_inst *I = p->I;
   double L = vo_length(p);
   if (!(p->nvns >= 4 && p->T > 0)) return;                             /* :711 */
   for (int var = 0; var < num_changes; ++var) {
      int from = vo_rand_int(g, 0, p->nvns - 1);                        /* :723 */
      int to = vo_rand_int(g, 0, p->T - 1);                             /* :724 */
      while (abs(from - to) <= 2) {                                     /* :726-730 */
         from = vo_rand_int(g, 0, p->nvns - 1);
         to = vo_rand_int(g, 0, p->T - 1);
      }
      if (st) st->proposals++;
      int cf = p->vns[from], ct = p->vns[to];
      double len, dR = 0;
      if (from < p->T) {                                                /* :742-772 */
         double lb = vo_try_exchange_lb(p, from, to);
         if (skip && lb > budget) { if (st) st->lb_skips++; continue; }
         len = vo_try_exchange(p, from, to);
      } else {                                                          /* :774-803 */
         double lb = vo_try_replace_lb(p, cf, to);
         if (skip && lb > budget) { if (st) st->lb_skips++; continue; }
         len = vo_try_replace(p, cf, to);
         dR += I->reward[cf] - I->reward[ct];                           /* :784 */
      }
      if (st) st->evals++;
      if (len <= budget && (dR > 0 || (L - len) > MIN_CHANGE_EPS)) {    /* :805-806 */
         if (st) st->accepts++;
         if (to > from) {                                               /* :811-820 */
            vo_remove_point(p, to);
            vo_add_point(p, cf, to);
            if (from < p->T) { vo_remove_point(p, from); vo_add_point(p, ct, from); }
         } else {                                                       /* :821-832 */
            if (from < p->T) { vo_remove_point(p, from); vo_add_point(p, ct, from); }
            vo_remove_point(p, to);
            vo_add_point(p, cf, to);
         }
         p->vns[to] = cf;
         p->vns[from] = ct;
         L = vo_length(p);                                              /* :837 */
      }
   }
}

/* VnsSopSolver::fitDOPtoBudget vns_sop_solver.cpp:364-385 */
void vo_fit(vo_path *p, double budget)
{
   while (vo_length(p) > budget) vo_remove_point(p, p->T - 1);
   for (;;) {
      int id = p->T;
      if (id >= p->nvns) break;
      if (vo_try_add(p, p->vns[id], id) > budget) break;
      vo_add_point(p, p->vns[id], id);
   }
}

static int cmp_int(const void *a, const void *b) { return *(const int *)a - *(const int *)b; }

/* k distinct draws in [0, hi] (unordered_set insert loop :936-939,1007-1010), then sorted
 * (CSort quicksort :945,1015 - only the order statistics matter) */
static void draw_distinct_sorted(vo_rng *g, int hi, int k, int *out)
{
   int n = 0;
   while (n < k) {
      int v = vo_rand_int(g, 0, hi), dup = 0;
      for (int i = 0; i < n; ++i) dup |= out[i] == v;
      if (!dup) out[n++] = v;
   }
   qsort(out, k, sizeof(int), cmp_int);
}
static int imin(int a, int b) { return a < b ? a : b; }

/* VnsSopSolver::pathInsert(.., 1) vns_sop_solver.cpp:931-999 */
void vo_path_insert(vo_path *p, double budget, vo_rng *g)
{
   int sizeV = p->nvns - 1, r[3];
   if (sizeV < 2) return;
   draw_distinct_sorted(g, sizeV, 3, r);
   int fs = r[0], fe = r[1], to = r[2], up = 1;
   if (vo_rand_int(g, 0, 1) == 1) { to = r[0]; fs = r[1]; fe = r[2]; up = 0; }
   int n = fe - fs + 1, seg[n];
   memcpy(seg, p->vns + fs, sizeof(int) * n);
   if (up) {                                                            /* :958-972 */
      vec_insert(p->vns, &p->nvns, to, seg, n);
      vec_erase(p->vns, &p->nvns, fs, fe);
      if (to < p->T) vo_add_points(p, seg, n, to);
      if (fs < p->T) vo_remove_range(p, fs, imin(fe, p->T - 1));
   } else {                                                             /* :974-990 */
      vec_erase(p->vns, &p->nvns, fs, fe);
      vec_insert(p->vns, &p->nvns, to, seg, n);
      if (fs < p->T) vo_remove_range(p, fs, imin(fe, p->T - 1));
      if (to < p->T) vo_add_points(p, seg, n, to);
   }
   vo_fit(p, budget);                                                   /* :992 */
}

/* VnsSopSolver::pathExchange(.., 1) vns_sop_solver.cpp:1001-1089 */
void vo_path_exchange(vo_path *p, double budget, vo_rng *g)
{
   int sizeV = p->nvns - 1, r[4];
   if (sizeV < 3) return;
   draw_distinct_sorted(g, sizeV, 4, r);
   int fs = r[0], fe = r[1], ts = r[2], te = r[3], up = 1;
   if (vo_rand_int(g, 0, 1) == 1) { fs = r[2]; fe = r[3]; ts = r[0]; te = r[1]; up = 0; }
   int nf = fe - fs + 1, nt = te - ts + 1, vf[nf], vt[nt];
   memcpy(vf, p->vns + fs, sizeof(int) * nf);
   memcpy(vt, p->vns + ts, sizeof(int) * nt);
   if (up) {                                                            /* :1030-1051 */
      vec_erase(p->vns, &p->nvns, ts, te);
      vec_insert(p->vns, &p->nvns, ts, vf, nf);
      vec_erase(p->vns, &p->nvns, fs, fe);
      vec_insert(p->vns, &p->nvns, fs, vt, nt);
      if (ts < p->T) { vo_remove_range(p, ts, imin(te, p->T - 1)); vo_add_points(p, vf, nf, ts); }
      if (fs < p->T) { vo_remove_range(p, fs, imin(fe, p->T - 1)); vo_add_points(p, vt, nt, fs); }
   } else {                                                             /* :1053-1080 */
      vec_erase(p->vns, &p->nvns, fs, fe);
      vec_insert(p->vns, &p->nvns, fs, vt, nt);
      vec_erase(p->vns, &p->nvns, ts, te);
      vec_insert(p->vns, &p->nvns, ts, vf, nf);
      if (fs < p->T) { vo_remove_range(p, fs, imin(fe, p->T - 1)); vo_add_points(p, vt, nt, fs); }
      if (ts < p->T) { vo_remove_range(p, ts, imin(te, p->T - 1)); vo_add_points(p, vf, nf, ts); }
   }
   vo_fit(p, budget);                                                   /* :1082 */
}

/* VnsSopSolver::greedy_insertion vns_sop_solver.cpp:1091-1155 */
void vo_greedy(vo_path *p, double budget)
{
   const vo_inst *I = p->I;
   int added = 1;
   while (added) {
      added = 0;
      double best = VO_M, L = vo_length(p), newL = L;
      int bi = -1, bt = -1;
      for (int av = p->T; av < p->nvns; ++av) {
         int c = p->vns[av];
         for (int pos = 0; pos < p->T + 1; ++pos) {
            double d = vo_try_add(p, c, pos);
            double per = (d - L) / I->reward[c];                        /* :1116-1117 */
            if (per < best && d <= budget) { best = per; newL = d; bi = av; bt = pos; }
         }
      }
      if (bi >= 0 && bt >= 0 && newL <= budget) {                       /* :1129-1146 */
         int c = p->vns[bi];
         added = 1;
         vo_add_point(p, c, bt);
         if (bi > bt) { vec_erase(p->vns, &p->nvns, bi, bi); vec_insert(p->vns, &p->nvns, bt, &c, 1); }
         else { vec_insert(p->vns, &p->nvns, bt, &c, 1); vec_erase(p->vns, &p->nvns, bi, bi); }
      }
   }
}

/* VnsSopSolver::iterate vns_sop_solver.cpp:82-260 (wall-clock stop :189-193 left out: the stock
 * limit is 168000 s).  p enters as the empty path with the initial VNS vector (:118-122). */
void vo_solve(vo_path *p, const vo_params *prm, vo_rng *g, vo_result *res)
{
   vo_path *act = vo_path_new(p->I);
   int num_avail = p->nvns, last_impr = 1, it = 0, stop = 0;
   if (prm->greedy) vo_greedy(p, prm->budget);                          /* :127-129 */
   res->n_improvements = 0;
   res->impr_reward[0] = vo_reward(p);                                  /* :150-154 */
   res->impr_length[0] = vo_length(p);
   res->n_improvements = 1;
   while (!stop) {
      it++;
      if (it >= prm->num_iterations) stop = 1;                          /* :178 */
      if (it - last_impr >= prm->num_iterations_unimproved) stop = 1;   /* :182 */
      int k = 1;
      while (k <= 2) {
         vo_path_copy(act, p);                                          /* :195-197 */
         double r0 = vo_reward(act), l0 = vo_length(act);
         if (k == 1) vo_path_insert(act, prm->budget, g); else vo_path_exchange(act, prm->budget, g);
         int iters = num_avail * num_avail;                             /* :346 */
         if (k == 1) vo_insert_random(act, iters, prm->budget, prm->skip, g, 0);
         else vo_exchange_random(act, iters, prm->budget, prm->skip, g, 0);
         double r1 = vo_reward(act), l1 = vo_length(act);
         if (r1 > r0 || (r1 == r0 && l1 < l0)) {                        /* :213 */
            last_impr = it;
            if (res->n_improvements < 4096) {
               res->impr_reward[res->n_improvements] = r1;
               res->impr_length[res->n_improvements] = l1;
            }
            res->n_improvements++;
            vo_path_copy(p, act);
            k = 1;
         } else {
            k++;
         }
      }
   }
   res->num_iters = it;
   res->num_iters_last_impr = last_impr;
   res->reward = vo_reward(p);
   res->length = vo_length(p);
   vo_path_free(act);
}
