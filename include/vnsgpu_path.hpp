// C++ mirror of the reference's path-evaluator interface over the C ABI of vnsgpu.h: the same method
// names, argument meaning and error behaviour as op::VnsSopPath (sources/vns_sop_path.h:22-194) and the
// neighbourhood drivers of op::VnsSopSolver (sources/vns_sop_solver.h), so solver code written against the
// reference reads the same against the GPU engine.  Header-only; link with -lvnsgpu.
//
// Differences a caller sees (all forced by "the state lives in HBM"):
//   * a VnsSopPath is a handle; copy-assignment is a device-to-device copy of targets, VNS vector and both
//     DP tables (the reference deep-copies std::vectors, vns_sop_solver.cpp:195-197,227-228);
//   * the VNS vector travels with the path (the reference passes it alongside, e.g. insertRandom(path, vns, n));
//   * errors the reference reports with ERROR(...) + exit(1) arrive as vnsgpu::Error (what() = the message);
//   * rand() is an explicit vnsgpu_rng value with the same stream as glibc's (math_common.cpp:21-28).
#ifndef VNSGPU_PATH_HPP
#define VNSGPU_PATH_HPP
#include <cmath>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "vnsgpu.h"

namespace vnsgpu {

struct Error : std::runtime_error {
   vnsgpu_status status;
   Error(vnsgpu_status s, const std::string &msg) : std::runtime_error(msg), status(s) {}
};

struct IndexSOP {   // sources/heuristic_types.h: one (cluster, sample) pair of getPath()
   int clusterIndex, nodeIndex;
};

class Engine {
 public:
   explicit Engine(int device = 0)
   {
      vnsgpu_status s = vnsgpu_create(&ctx_, device);
      if (s != VNSGPU_OK) {
         std::string msg = ctx_ ? vnsgpu_last_error(ctx_) : "vnsgpu_create failed (no CUDA device?)";
         if (ctx_) vnsgpu_destroy(ctx_);
         ctx_ = nullptr;
         throw Error(s, msg);
      }
   }
   ~Engine() { if (ctx_) vnsgpu_destroy(ctx_); }
   Engine(const Engine &) = delete;
   Engine &operator=(const Engine &) = delete;

   vnsgpu_ctx *ctx() const { return ctx_; }
   void check(vnsgpu_status s) const
   {
      if (s != VNSGPU_OK) throw Error(s, vnsgpu_last_error(ctx_));
   }
   // VnsSopPath::setSamplesDistances (vns_sop_path.cpp:38-48) fed by getSOPFromNodes / getSOPFromWeightMatrix
   void setClusters(const std::vector<int> &m, const std::vector<double> &reward, int start, int goal)
   {
      check(vnsgpu_set_clusters(ctx_, (int)m.size(), m.data(), reward.data(), start, goal));
   }
   void buildTableEuclid(const std::vector<double> &xy, int rule) { check(vnsgpu_build_table_euclid(ctx_, xy.data(), rule)); }
   void setTableDense(const std::vector<double> &W) { check(vnsgpu_set_table_dense(ctx_, W.data())); }
   void buildTableDubins(const std::vector<double> &centers, int S, double rho, bool do_ceil)
   {
      check(vnsgpu_build_table_dubins(ctx_, centers.data(), S, rho, do_ceil ? 1 : 0));
   }
   // calcClusterDistancesBounds (vns_sop_solver.cpp:262-297)
   void getBounds(std::vector<double> &dmin, std::vector<double> &dmax, int C)
   {
      dmin.resize((size_t)C * C);
      dmax.resize((size_t)C * C);
      check(vnsgpu_get_bounds(ctx_, dmin.data(), dmax.data()));
   }
   int64_t launches() const { return vnsgpu_launch_count(ctx_); }

 private:
   vnsgpu_ctx *ctx_ = nullptr;
};

class VnsSopPath {
 public:
   // VnsSopPath(start, goal, type) + update() (vns_sop_solver.cpp:118, vns_sop_path.cpp:1321-1404): an empty
   // tour start -> goal whose VNS vector is `vns`
   VnsSopPath(Engine &e, const std::vector<int> &vns, int T = 0) : e_(&e)
   {
      e_->check(vnsgpu_path_create(e_->ctx(), &h_));
      set(vns, T);
   }
   explicit VnsSopPath(Engine &e) : e_(&e) { e_->check(vnsgpu_path_create(e_->ctx(), &h_)); }
   VnsSopPath(const VnsSopPath &o) : e_(o.e_)
   {
      e_->check(vnsgpu_path_create(e_->ctx(), &h_));
      e_->check(vnsgpu_path_copy(e_->ctx(), h_, o.h_));
   }
   VnsSopPath &operator=(const VnsSopPath &o)
   {
      if (this != &o) e_->check(vnsgpu_path_copy(e_->ctx(), h_, o.h_));
      return *this;
   }
   ~VnsSopPath() { if (h_ >= 0) vnsgpu_path_destroy(e_->ctx(), h_); }
   int handle() const { return h_; }
   Engine &engine() const { return *e_; }

   void set(const std::vector<int> &vns, int T) { e_->check(vnsgpu_path_set(e_->ctx(), h_, vns.data(), (int)vns.size(), T)); }

   // ---- const queries ---------------------------------------------------------------------------
   int getNumTargets() const
   {
      int T = 0;
      e_->check(vnsgpu_path_get(e_->ctx(), h_, &T, 0, 0, 0, 0, 0, 0));
      return T;
   }
   int getSize() const { return getNumTargets() + 1; }   // vns_sop_path.cpp:58-61
   std::vector<int> getAllTargets() const
   {
      int T = getNumTargets();
      std::vector<int> t((size_t)T + 1);
      e_->check(vnsgpu_path_get(e_->ctx(), h_, &T, t.data(), 0, 0, 0, 0, 0));
      t.resize(T);
      return t;
   }
   std::vector<int> getVnsVector() const
   {
      int n = 0;
      e_->check(vnsgpu_path_get(e_->ctx(), h_, 0, 0, &n, 0, 0, 0, 0));
      std::vector<int> v((size_t)n + 1);
      e_->check(vnsgpu_path_get(e_->ctx(), h_, 0, 0, &n, v.data(), 0, 0, 0));
      v.resize(n);
      return v;
   }
   double getPathLength() const { double v; e_->check(vnsgpu_path_get(e_->ctx(), h_, 0, 0, 0, 0, &v, 0, 0)); return v; }
   double getPathLengthRear() const { double v; e_->check(vnsgpu_path_get(e_->ctx(), h_, 0, 0, 0, 0, 0, &v, 0)); return v; }
   double getReward() const { double v; e_->check(vnsgpu_path_get(e_->ctx(), h_, 0, 0, 0, 0, 0, 0, &v)); return v; }
   // getPath (vns_sop_path.cpp:1542-1599): T + 2 entries, start and goal included
   std::vector<IndexSOP> getPath() const
   {
      const int n = getNumTargets() + 2;
      std::vector<int> c(n), s(n);
      e_->check(vnsgpu_path_get_solution(e_->ctx(), h_, c.data(), s.data(), 0));
      std::vector<IndexSOP> p(n);
      for (int i = 0; i < n; ++i) p[i] = IndexSOP{c[i], s[i]};
      return p;
   }
   double getPathLengthCalculate() const   // vns_sop_path.cpp:1700-1741
   {
      const int n = getNumTargets() + 2;
      std::vector<int> c(n), s(n);
      double v = 0;
      e_->check(vnsgpu_path_get_solution(e_->ctx(), h_, c.data(), s.data(), &v));
      return v;
   }

   // ---- candidate evaluation: one query = a batch of one; solvers should batch ---------------------
   struct Eval { double length, lowerBound; unsigned flags; };
   Eval evaluate(int op, int a, int b, double budget = INFINITY) const
   {
      vnsgpu_cand c{op, a, b, (op == VNSGPU_OP_ADD || op == VNSGPU_OP_REPLACE) ? a : -1};
      double len = 0, lb = 0;
      uint8_t fl = 0;
      e_->check(vnsgpu_eval_batch(e_->ctx(), h_, 1, &c, budget, 0, &len, &lb, &fl, 0));
      if (fl & VNSGPU_F_BAD) throw Error(VNSGPU_E_RANGE, "index out of range in tryTo* (the reference exit(1)s here)");
      return Eval{len, lb, fl};
   }
   double tryToAdd(int cluster, int idx) const { return evaluate(VNSGPU_OP_ADD, cluster, idx).length; }
   double tryToAddLowerBound(int cluster, int idx) const { return evaluate(VNSGPU_OP_ADD, cluster, idx).lowerBound; }
   double tryToReplace(int cluster, int idx) const { return evaluate(VNSGPU_OP_REPLACE, cluster, idx).length; }
   double tryToReplaceLowerBound(int cluster, int idx) const { return evaluate(VNSGPU_OP_REPLACE, cluster, idx).lowerBound; }
   double tryToRemove(int idx) const { return evaluate(VNSGPU_OP_REMOVE, idx, 0).length; }
   double tryToMove(int what, int where) const { return evaluate(VNSGPU_OP_MOVE, what, where).length; }
   double tryToMoveLowerBound(int what, int where) const { return evaluate(VNSGPU_OP_MOVE, what, where).lowerBound; }
   double tryToExchange(int i1, int i2) const { return evaluate(VNSGPU_OP_EXCHANGE, i1, i2).length; }
   double tryToExchangeLowerBound(int i1, int i2) const { return evaluate(VNSGPU_OP_EXCHANGE, i1, i2).lowerBound; }
   // the batched form: len / lb / flags sized to cands.size(); returns the first accepted index or -1
   int64_t evalBatch(const std::vector<vnsgpu_cand> &cands, double budget, bool skip, std::vector<double> *len,
                     std::vector<double> *lb, std::vector<uint8_t> *flags) const
   {
      const int64_t n = (int64_t)cands.size();
      if (len) len->resize(n);
      if (lb) lb->resize(n);
      if (flags) flags->resize(n);
      int64_t first = -1;
      e_->check(vnsgpu_eval_batch(e_->ctx(), h_, n, cands.data(), budget, skip ? 1 : 0, len ? len->data() : 0,
                                  lb ? lb->data() : 0, flags ? flags->data() : 0, &first));
      return first;
   }

   // ---- edits -----------------------------------------------------------------------------------
   void addPoint(int cluster, int idx) { e_->check(vnsgpu_path_add_point(e_->ctx(), h_, cluster, idx)); }
   void removePoint(int idx) { e_->check(vnsgpu_path_remove_point(e_->ctx(), h_, idx)); }

   // ---- neighbourhood drivers of VnsSopSolver, run on the device against this path ------------------
   vnsgpu_ls_stats insertRandom(int num_changes, double budget, bool skip, vnsgpu_rng &rng)   // :469-684
   {
      vnsgpu_ls_stats st;
      e_->check(vnsgpu_local_search(e_->ctx(), h_, 1, num_changes, budget, skip ? 1 : 0, &rng, &st));
      return st;
   }
   vnsgpu_ls_stats exchangeRandom(int num_changes, double budget, bool skip, vnsgpu_rng &rng)   // :686-864
   {
      vnsgpu_ls_stats st;
      e_->check(vnsgpu_local_search(e_->ctx(), h_, 2, num_changes, budget, skip ? 1 : 0, &rng, &st));
      return st;
   }
   void pathInsert(double budget, vnsgpu_rng &rng) { e_->check(vnsgpu_shake(e_->ctx(), h_, 1, budget, &rng)); }     // :931-999
   void pathExchange(double budget, vnsgpu_rng &rng) { e_->check(vnsgpu_shake(e_->ctx(), h_, 2, budget, &rng)); }   // :1001-1089
   void fitDOPtoBudget(double budget) { e_->check(vnsgpu_fit_to_budget(e_->ctx(), h_, budget)); }                   // :364-385
   void greedy_insertion(double budget) { e_->check(vnsgpu_greedy(e_->ctx(), h_, budget)); }                        // :1091-1155
   // the same greedy with every round's tryToAdd grid as one batched evaluation (the choice beyond a few hundred clusters)
   int greedy_insertion_batched(double budget)
   {
      int rounds = 0;
      e_->check(vnsgpu_greedy_batched(e_->ctx(), h_, budget, &rounds, nullptr));
      return rounds;
   }
   // opt-in, not the reference's trajectory: best improving candidate of the whole insert (1) / exchange (2) neighbourhood,
   // its application, and the descent over both neighbourhoods (returns the number of moves applied)
   vnsgpu_best_move_t bestMove(int kind, double budget)
   {
      vnsgpu_best_move_t m;
      e_->check(vnsgpu_best_move(e_->ctx(), h_, kind, budget, &m));
      return m;
   }
   void applyMove(int kind, int from, int to) { e_->check(vnsgpu_apply_move(e_->ctx(), h_, kind, from, to)); }
   int bestImprovement(double budget, int max_steps = 0)
   {
      int steps = 0;
      e_->check(vnsgpu_best_improvement(e_->ctx(), h_, budget, max_steps, &steps, nullptr));
      return steps;
   }

 private:
   Engine *e_;
   int h_ = -1;
};

} // namespace vnsgpu
#endif
