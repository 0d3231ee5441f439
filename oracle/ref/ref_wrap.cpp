// TEST INFRASTRUCTURE ONLY: extern "C" access to the UNMODIFIED reference evaluator
// op::VnsSopPath (sources/vns_sop_path.{h,cpp}) and the neighbourhood drivers of
// op::VnsSopSolver (sources/vns_sop_solver.cpp:364-385,469-684,686-864,931-1155) for
// differential tests (tests/), golden-vector generation (tests/golden/make_golden.py) and the
// CPU baseline leg of bench.py.  Built by oracle/ref/build_ref.sh into oracle/_ref/.
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <list>
#include <map>
#include <sstream>
#include <string>
#include <unordered_set>
#include <vector>

// the neighbourhood drivers and the DP arrays are private/protected in the reference
#define private public
#define protected public
#include "ref_common.h"
#include "vns_sop_solver.h"
#undef private
#undef protected

int oracle_ref_verbosity = 0;
namespace crl {
CLoggerPtr logger = 0;
CLoggerPtr getLogger(const char *) { return 0; }
void shutdownLogger(void) {}
}
extern bool randSeeded;  // sources/math_common.cpp:9

using op::VnsSopPath;
using op::VnsSopSolver;

static crl::CConfig *g_cfg = 0;
static VnsSopSolver *g_solver = 0;
static std::vector<VnsSopPath *> g_paths;

static VnsSopPath &P(int h) { return *g_paths.at(h); }

extern "C" {

// Builds the solver exactly as sop.cpp:368 does, then performs the table publication of
// VnsSopSolver::iterate (vns_sop_solver.cpp:86-122) without starting the search.
int oref_load(const char *problem, const char *sop_type, int verbosity)
{
   oracle_ref_verbosity = verbosity;
   for (VnsSopPath *p : g_paths) delete p;
   g_paths.clear();
   delete g_solver;
   delete g_cfg;
   g_cfg = new crl::CConfig();
   oracle_ref::register_config(*g_cfg);
   g_cfg->set<std::string>("sop-type", sop_type);
   g_solver = new VnsSopSolver(*g_cfg, problem);
   VnsSopSolver &s = *g_solver;
   s.calcClusterDistancesBounds();
   VnsSopPath::setSamplesDistances(s.nodesAllClusters, s.nodesAllDistances, s.clustersMinDistances,
                                   s.clustersMaxDistances);
   s.nodesForInsertion = s.nodesAllClusters;
   if (s.startIndex != s.goalIndex) {
      s.nodesForInsertion.erase(s.nodesForInsertion.begin() + std::max(s.startIndex, s.goalIndex));
      s.nodesForInsertion.erase(s.nodesForInsertion.begin() + std::min(s.startIndex, s.goalIndex));
   } else {
      s.nodesForInsertion.erase(s.nodesForInsertion.begin() + s.startIndex);
   }
   s.availableNodes = s.getReachableNodes(s.nodesForInsertion);
   s.numAvailableNodes = s.availableNodes.size();
   s.tourVNSGOPPath = VnsSopPath(s.startIndex, s.goalIndex, s.gopType);
   return 0;
}

void oref_set_skip(int on) { g_solver->lower_bound_above_budget_then_skip = on != 0; }
void oref_set_budget(double b) { g_solver->budget = b; }
int oref_num_clusters() { return (int)g_solver->nodesAllClusters.size(); }
int oref_cluster_size(int c) { return (int)g_solver->nodesAllClusters[c].size(); }
double oref_cluster_reward(int c) { return g_solver->nodesAllClusters[c][0].reward; }
int oref_node_id(int c, int s) { return (int)g_solver->nodesAllClusters[c][s].id; }
void oref_node_xy(int c, int s, double *xy)
{
   xy[0] = g_solver->nodesAllClusters[c][s].x;
   xy[1] = g_solver->nodesAllClusters[c][s].y;
}
double oref_budget() { return g_solver->budget; }
int oref_start() { return g_solver->startIndex; }
int oref_goal() { return g_solver->goalIndex; }
int oref_old_start() { return g_solver->oldstartIndex; }
int oref_old_goal() { return g_solver->oldgoalIndex; }
double oref_max_reward() { return g_solver->maximalRewardAll; }
int oref_num_available() { return g_solver->numAvailableNodes; }
// initial VNS vector of iterate() (vns_sop_solver.cpp:119-122)
int oref_initial_vns(int *out)
{
   int n = 0;
   for (auto &cl : g_solver->availableNodes) out[n++] = cl[0].cluster_id;
   return n;
}
void oref_block(int c1, int c2, double *out)
{
   auto &b = VnsSopPath::allDistances[c1][c2];
   size_t k = 0;
   for (auto &row : b)
      for (double v : row) out[k++] = v;
}
double oref_dmin(int c1, int c2) { return VnsSopPath::clustersMinDistances[c1][c2]; }
double oref_dmax(int c1, int c2) { return VnsSopPath::clustersMaxDistances[c1][c2]; }

// ---- path handles -------------------------------------------------------------------------
int oref_path_new()
{
   VnsSopSolver &s = *g_solver;
   g_paths.push_back(new VnsSopPath(s.startIndex, s.goalIndex, s.gopType));
   return (int)g_paths.size() - 1;
}
void oref_path_copy(int dst, int src) { P(dst) = P(src); }
void oref_path_set_targets(int h, const int *t, int n)
{
   P(h).targets.assign(t, t + n);
   P(h).update();
}
int oref_path_num_targets(int h) { return P(h).getNumTargets(); }
void oref_path_get_targets(int h, int *out)
{
   for (size_t i = 0; i < P(h).targets.size(); ++i) out[i] = P(h).targets[i];
}
double oref_try_add(int h, int c, int idx) { return P(h).tryToAdd(c, idx); }
double oref_try_add_lb(int h, int c, int idx) { return P(h).tryToAddLowerBound(c, idx); }
double oref_try_replace(int h, int c, int idx) { return P(h).tryToReplace(c, idx); }
double oref_try_replace_lb(int h, int c, int idx) { return P(h).tryToReplaceLowerBound(c, idx); }
double oref_try_remove(int h, int idx) { return P(h).tryToRemove(idx); }
double oref_try_move(int h, int a, int b) { return P(h).tryToMove(a, b); }
double oref_try_move_lb(int h, int a, int b) { return P(h).tryToMoveLowerBound(a, b); }
double oref_try_exchange(int h, int a, int b) { return P(h).tryToExchange(a, b); }
double oref_try_exchange_lb(int h, int a, int b) { return P(h).tryToExchangeLowerBound(a, b); }
double oref_length(int h) { return P(h).getPathLength(); }
double oref_length_rear(int h) { return P(h).getPathLengthRear(); }
double oref_length_calc(int h) { return P(h).getPathLengthCalculate(); }
double oref_reward(int h) { return P(h).getReward(); }
void oref_add_point(int h, int c, int idx) { P(h).addPoint(c, idx); }
void oref_remove_point(int h, int idx) { P(h).removePoint(idx); }
void oref_add_points(int h, const int *ids, int n, int idx)
{
   P(h).addPoint(std::vector<int>(ids, ids + n), idx);
}
void oref_remove_range(int h, int a, int b) { P(h).removePoint(a, b); }
// which: 0 = shortest (forward DP), 1 = shortest_back; returns the layer's sample count
int oref_get_layer(int h, int layer, int which, double *dist, int *pred)
{
   auto &mat = which == 0 ? P(h).shortest : P(h).shortest_back;
   auto &row = mat.at(layer);
   for (size_t i = 0; i < row.size(); ++i) {
      dist[i] = row[i].distance;
      pred[i] = row[i].idClusterNode;
   }
   return (int)row.size();
}
int oref_get_path(int h, int *cluster, int *node)
{
   std::vector<IndexSOP> p = P(h).getPath();
   for (size_t i = 0; i < p.size(); ++i) {
      cluster[i] = p[i].clusterIndex;
      node[i] = p[i].nodeIndex;
   }
   return (int)p.size();
}

// ops: 0 ADD(c,idx) 1 REPLACE(c,idx) 2 REMOVE(idx,-) 3 MOVE(what,where) 4 EXCHANGE(i1,i2)
// cands = n x {op,a,b}; len/lb may be null.  This loop is what bench.py's cpu_baseline times.
void oref_eval_batch(int h, int n, const int *cands, double *len, double *lb)
{
   VnsSopPath &p = P(h);
   for (int i = 0; i < n; ++i) {
      int o = cands[3 * i], a = cands[3 * i + 1], b = cands[3 * i + 2];
      double l = 0, bound = 0;
      switch (o) {
      case 0: if (lb) bound = p.tryToAddLowerBound(a, b); l = p.tryToAdd(a, b); break;
      case 1: if (lb) bound = p.tryToReplaceLowerBound(a, b); l = p.tryToReplace(a, b); break;
      case 2: l = p.tryToRemove(a); bound = l; break;
      case 3: if (lb) bound = p.tryToMoveLowerBound(a, b); l = p.tryToMove(a, b); break;
      case 4: if (lb) bound = p.tryToExchangeLowerBound(a, b); l = p.tryToExchange(a, b); break;
      }
      if (len) len[i] = l;
      if (lb) lb[i] = bound;
   }
}

// ---- RNG + neighbourhood drivers -----------------------------------------------------------
void oref_seed(unsigned seed) { randSeeded = true; srand(seed); }
int oref_rand() { return rand(); }
int oref_rand_int(int lo, int hi) { return randIntMinMax(lo, hi); }

static std::vector<int> g_vns;
static void vns_in(const int *v, int n) { g_vns.assign(v, v + n); }
static void vns_out(int *v) { std::copy(g_vns.begin(), g_vns.end(), v); }

void oref_insert_random(int h, int *vns, int n, int num_changes)
{
   vns_in(vns, n); g_solver->insertRandom(P(h), g_vns, num_changes); vns_out(vns);
}
void oref_exchange_random(int h, int *vns, int n, int num_changes)
{
   vns_in(vns, n); g_solver->exchangeRandom(P(h), g_vns, num_changes); vns_out(vns);
}
void oref_path_insert(int h, int *vns, int n)
{
   vns_in(vns, n); g_solver->pathInsert(P(h), g_vns, 1); vns_out(vns);
}
void oref_path_exchange(int h, int *vns, int n)
{
   vns_in(vns, n); g_solver->pathExchange(P(h), g_vns, 1); vns_out(vns);
}
void oref_fit(int h, int *vns, int n)
{
   vns_in(vns, n); g_solver->fitDOPtoBudget(P(h), g_vns); vns_out(vns);
}
void oref_greedy(int h, int *vns, int n)
{
   vns_in(vns, n); g_solver->greedy_insertion(P(h), g_vns); vns_out(vns);
}

} // extern "C"
