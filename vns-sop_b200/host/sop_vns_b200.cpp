// sop_vns_b200 - the reference's `sop_vns --gui=none --problem=X.sop` front end over the GPU engine.
//
// Same option names, same sop.cfg syntax, same output files (results.log record with the reference's 17
// columns, <iter>/path.txt, gop.txt, targets.txt, settings.txt, info.txt) as sources/sop.cpp:87-406 +
// VnsSopSolver (sources/vns_sop_solver.cpp:25-72,82-260,1443-1601) + crl::CAlgorithm
// (comrob/crl/crl.algorithm/src/algorithm.cc:90-196,331-345).  What runs where:
//   host  : option parsing, .sop parsing + renumbering (sop_instance.hpp), the result files;
//   device: distance table + bounds, and the whole VNS (vnsgpu_vns_run: one persistent chain per solve,
//           or vnsgpu_vns_run_batch: one thread block per seed for --batch=N).
// `--device-loop=0` keeps VnsSopSolver::iterate's control flow on the host, one C-ABI call per shake / local
// search (the binding INTEGRATION.md describes); both modes follow the reference's seeded trajectory.
//
// New keys (not in the reference): --seed=N (srand(N); default: time, like math_common.cpp:9-14),
// --device=N, --gpus=N (--batch only: the seeds are spread over devices N0 .. N0+N-1, one host thread and one context per
// device - replicas, no collective: the honest multi-GPU mode of the search on shipped instances),
// --device-loop=0|1, --config-file=path (default ./sop.cfg when present).
// Not provided: the GUI (gui=none is the only accepted value), log4cxx logging, sop-type=op|opn|dop loaders
// (sop-type=opn throws in the reference itself; its datasets are pre-sampled .sop files), the ILP solver.
#include <sys/stat.h>
#include <sys/utsname.h>

#include <chrono>
#include <cstdio>
#include <cstring>
#include <ctime>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <map>
#include <sstream>
#include <thread>

#include "sop_instance.hpp"
#include "vnsgpu_path.hpp"

namespace {

struct Option { std::string key, value; };

// Typed options with the reference's defaults; print() is what settings.txt holds (crl::CConfig::print)
class Config {
 public:
   Config()
   {
      const char *defaults[][2] = {
          {"output", "./"}, {"results", "results.log"}, {"info", "info.txt"}, {"settings", "settings.txt"},
          {"result-path", "path.txt"}, {"result-canvas-output", "path"}, {"result-canvas-suffixes", ""}, {"name", "no_name"},
          {"iteration", "-1"}, {"batch", "-1"}, {"continue", "0"}, {"save-results", "1"}, {"save-info", "1"},
          {"save-settings", "1"}, {"save-visual", "0"}, {"verbose-result-log", "0"}, {"save-targets", "1"},
          {"save-sampled-path", "1"}, {"targets", "targets.txt"}, {"sampled-path", "sampled-path.txt"},
          {"num-iterations", "2000"}, {"num-iterations-unimproved", "1000"}, {"use-rvns", "0"},
          {"maximal-calculation-time-sec", "168000"}, {"lower-bound-above-budget-then-skip", "1"},
          {"generate-examples-of-operations", "0"}, {"sop-type", "sop"}, {"sop-solver", "vns"},
          {"initial-greedy-solution", "1"}, {"gui", "none"}, {"problem", ""}, {"nowait", "0"},
          {"seed", ""}, {"device", "0"}, {"gpus", "1"}, {"device-loop", "1"}, {"config-file", ""}, {"dump-problem", "0"}};
      for (auto &d : defaults) { order_.push_back(d[0]); v_[d[0]] = d[1]; }
   }
   bool known(const std::string &k) const { return v_.count(k) != 0; }
   // first store wins: CLI is stored before the config file (sop.cpp:224-233)
   void store(const std::string &k, const std::string &val)
   {
      if (!known(k)) order_.push_back(k);
      if (!set_.count(k)) { v_[k] = val; set_[k] = true; }
   }
   std::string str(const std::string &k) const { auto it = v_.find(k); return it == v_.end() ? std::string() : it->second; }
   int num(const std::string &k) const { return std::atoi(str(k).c_str()); }
   bool flag(const std::string &k) const
   {
      const std::string s = str(k);
      return s == "1" || s == "true" || s == "on" || s == "yes";
   }
   bool is_set(const std::string &k) const { return set_.count(k) != 0; }
   void print(std::ostream &out) const
   {
      for (const std::string &k : order_) out << k << "=" << str(k) << std::endl;
   }

 private:
   std::vector<std::string> order_;
   std::map<std::string, std::string> v_;
   std::map<std::string, bool> set_;
};

// gui-* / draw-* / canvas-* and friends belong to the GUI the reference's sop.cfg configures; accepted and ignored
bool gui_key(const std::string &k)
{
   static const char *pre[] = {"gui", "draw-", "canvas-", "pic-", "save-pic", "logger-config", "num-threads", "maximal-memory-MB",
                               "result-canvas", "neighborhood-", "planning-state", "start-", "goal-", "map-", "ilp-", "sampling-"};
   for (const char *p : pre)
      if (k.compare(0, std::strlen(p), p) == 0) return true;
   return false;
}

void load_config_file(Config &cfg, const std::string &file, bool must_exist)
{
   std::ifstream in(file.c_str());
   if (!in) {
      if (must_exist) throw std::runtime_error("Cannot open config file '" + file + "'");
      return;
   }
   for (std::string line; std::getline(in, line);) {
      const size_t hash = line.find('#');
      if (hash != std::string::npos) line.erase(hash);
      const size_t eq = line.find('=');
      if (eq == std::string::npos) continue;
      const std::string k = sopvns::detail::trim(line.substr(0, eq)), v = sopvns::detail::trim(line.substr(eq + 1));
      if (k.empty()) continue;
      if (cfg.known(k) || gui_key(k)) cfg.store(k, v);
      else throw std::runtime_error("unrecognised option '" + k + "' in " + file);   // boost::program_options does the same
   }
}

bool make_dirs(const std::string &dir)   // crl::createDirectory: mkdir -p
{
   std::string cur;
   for (size_t i = 0; i <= dir.size(); ++i) {
      if (i == dir.size() || dir[i] == '/') {
         if (!cur.empty() && cur != "." && cur != "..") {
            struct stat st;
            if (stat(cur.c_str(), &st) != 0 && mkdir(cur.c_str(), 0777) != 0 && errno != EEXIST) return false;
         }
      }
      if (i < dir.size()) cur.push_back(dir[i]);
   }
   return true;
}

// crl::string_format / default ostream formatting: 6 significant digits (stringconversions.h:16-21)
template <class T> std::string fmt(const T &v) { std::ostringstream s; s << v; return s.str(); }

struct RunRecord {
   vnsgpu_vns_result res{};
   std::vector<double> impr_reward, impr_length, impr_ms;
   std::vector<vnsgpu::IndexSOP> path;   // internal (cluster, sample) pairs
   double reward = 0, length = 0;
   long ctime_ms = 0;
};

struct Solver {
   Config &cfg;
   sopvns::Problem pb;
   std::string problem_file, name, output;
   int iter;

   std::string outputPath(const std::string &f, std::string &dir) const   // algorithm.cc:331-340
   {
      if (!output.empty()) { dir = output.back() == '/' ? output : output + "/"; return dir + f; }
      dir = "./";
      return f;
   }
   std::string outputIterPath(const std::string &f, std::string &dir) const   // algorithm.cc:343-356
   {
      const std::string sep = iter >= 0 ? fmt(iter) + "/" : "/";
      if (!output.empty()) dir = (output.back() == '/' && sep[0] == '/') ? output + sep : output + "/" + sep;
      else dir = sep;
      return dir + f;
   }

   std::string record(const RunRecord &r) const   // defineResultLog + fillResultRecord (:1524-1601), CTextLog "COL:value;"
   {
      std::ostringstream nodes, clusters, impr;
      for (size_t i = 0; i < r.path.size(); ++i) {
         if (i) { nodes << ","; clusters << ","; }
         nodes << pb.node_id[pb.off[r.path[i].clusterIndex] + r.path[i].nodeIndex];
         clusters << pb.orig_cluster_id[r.path[i].clusterIndex];
      }
      for (size_t i = 0; i < r.impr_reward.size(); ++i) {
         if (i) impr << ",";
         impr << (long)r.impr_ms[i] << "|" << r.impr_reward[i] << "|" << r.impr_length[i];
      }
      const long max_ms = (long)cfg.num("maximal-calculation-time-sec") * 1000;
      std::ostringstream o;
      o << "NAME:" << name << ";METHOD:vns_sop;PROBLEM_FILE:" << problem_file << ";CTIME:" << r.ctime_ms << ";NUM_ITERS:" << r.res.num_iters
        << ";BUDGET:" << fmt(pb.budget) << ";REWARDS:" << fmt(r.reward) << ";MAX_ACHIEVABLE_REWARDS:" << fmt(pb.max_reward)
        << ";LENGTH:" << fmt(r.length) << ";NUM_ITERS_LAST_IMPR:" << r.res.num_iters_last_impr
        << ";CTIME_LAST_IMPR:" << (long)r.res.time_last_impr_ms << ";MAX_ALLOWED_CALC_TIME_MS:" << max_ms
        << ";GREEDY_INITIAL_SOLUTION:" << (cfg.flag("initial-greedy-solution") ? 1 : 0) << ";RESULT_TARGET_IDS:" << nodes.str()
        << ";RESULT_CLUSTER_IDS:" << clusters.str() << ";SOLUTION_TIME_REWARD_LENGTH_IMPROVEMENTS:" << impr.str()
        << ";LOWER_BOUND_ABOVE_BUDGET_THEN_SKIP:" << (cfg.flag("lower-bound-above-budget-then-skip") ? 1 : 0) << ";";
      return o.str();
   }

   void save(const RunRecord &r, const sopvns::SopFile &sf) const   // VnsSopSolver::save (:1443-1506)
   {
      std::string dir;
      auto open = [&](const std::string &key_or_name, bool is_key) {
         const std::string file = outputIterPath(is_key ? cfg.str(key_or_name) : key_or_name, dir);
         if (!make_dirs(dir)) throw std::runtime_error("Can not create file in path'" + file + "'");
         return file;
      };
      if (cfg.flag("save-settings") && cfg.flag("save-results")) {
         std::ofstream out(open("settings", true).c_str());
         cfg.print(out);
      }
      if (cfg.flag("save-info") && cfg.flag("save-results")) {
         std::ofstream out(open("info", true).c_str());
         out << "# program information " << std::endl << "VERSION: sop_vns_b200 (" << vnsgpu_version() << ")" << std::endl;
         struct utsname n;
         if (uname(&n) == 0)
            out << std::endl << "# host information " << std::endl << "PLATFORM:  " << n.sysname << " " << n.release << std::endl
                << "HOST: " << n.nodename << std::endl << "MACHINE: " << n.machine << std::endl;
         const time_t now = time(nullptr);
         char buf[32];
         out << std::endl << "# time information " << std::endl << "DATE: " << ctime_r(&now, buf) << std::endl;
         out << "# detail performance log" << std::endl << "SOLVE_TIME_CPU: " << r.ctime_ms << std::endl;
         out << "SOLVE_TIME_DEVICE: " << r.res.time_ms << std::endl << std::endl << "# log results" << std::endl;
         std::string rec = record(r);
         for (char &c : rec) if (c == ';') c = '\n';
         out << rec;
      }
      auto write_path = [&](const std::string &file) {
         std::ofstream ofs(file.c_str());
         if (!ofs) throw std::runtime_error("Cannot create path '" + file + "'");
         for (const vnsgpu::IndexSOP &p : r.path) ofs << p.clusterIndex << " " << p.nodeIndex << std::endl;
      };
      if (cfg.flag("save-results")) write_path(open("result-path", true));
      write_path(open("gop.txt", false));
      if (cfg.flag("save-targets")) {
         std::ofstream ofs(open("targets", true).c_str());
         ofs << std::setprecision(14);
         for (int c = 0; c < pb.C; ++c)
            for (int j = 0; j < pb.m[c]; ++j) {
               const int s = pb.off[c] + j, id = pb.node_id[s];
               const double x = sf.coords.empty() ? NAN : sf.coords[2 * (size_t)id], y = sf.coords.empty() ? NAN : sf.coords[2 * (size_t)id + 1];   // GraphNode() of a matrix instance
               ofs << id << " , " << x << " , " << y << " , " << pb.reward[c] << " , " << c << std::endl;
            }
      }
   }

   void append_log(const RunRecord &r) const   // CAlgorithm::appendToLog (algorithm.cc:178-196)
   {
      std::string dir;
      const std::string log = outputPath(cfg.str("results"), dir);
      if (!make_dirs(dir)) throw std::runtime_error("Can not create directory '" + dir + "'");
      std::ofstream out(log.c_str(), std::ios::app | std::ios::out);
      out << record(r) << std::endl;
   }
};

vnsgpu_vns_params params_of(const Config &cfg, const sopvns::Problem &pb, uint32_t seed)
{
   vnsgpu_vns_params p;
   std::memset(&p, 0, sizeof(p));
   p.num_iterations = cfg.num("num-iterations");
   p.num_iterations_unimproved = cfg.num("num-iterations-unimproved");
   p.lower_bound_skip = cfg.flag("lower-bound-above-budget-then-skip");
   p.greedy_initial = cfg.flag("initial-greedy-solution");
   p.budget = pb.budget;
   p.max_time_ms = cfg.num("maximal-calculation-time-sec") * 1000.0;
   p.seed = seed;
   p.max_improvements = 4096;
   return p;
}

void finish_record(RunRecord &r, const vnsgpu::VnsSopPath &tour)
{
   r.path = tour.getPath();
   r.reward = tour.getReward();
   r.length = tour.getPathLength();
}

// VnsSopSolver::iterate (vns_sop_solver.cpp:82-260) with the control flow on the host and every step a C-ABI
// call; `rng` plays the role of the process-global rand() state
RunRecord iterate_host_loop(vnsgpu::Engine &eng, const Config &cfg, const sopvns::Problem &pb, uint32_t seed, vnsgpu::VnsSopPath &tour)
{
   using clock = std::chrono::steady_clock;
   const auto t0 = clock::now();
   auto ms = [&]() { return (long)std::chrono::duration_cast<std::chrono::milliseconds>(clock::now() - t0).count(); };
   const vnsgpu_vns_params prm = params_of(cfg, pb, seed);
   vnsgpu_rng rng;
   vnsgpu_rng_seed(&rng, seed);
   const std::vector<int> vns = pb.initial_vns();
   const int nvns = (int)vns.size();
   tour.set(vns, 0);
   if (prm.greedy_initial) tour.greedy_insertion(pb.budget);                       // :128
   RunRecord r;
   int numItersLastImprovement = 1;
   long timeLastImprovement = ms();
   r.impr_reward.push_back(tour.getReward());
   r.impr_length.push_back(tour.getPathLength());
   r.impr_ms.push_back((double)timeLastImprovement);
   vnsgpu::VnsSopPath actual(eng);
   int act_iter = 0;
   bool stop = false;
   while (!stop) {                                                                // :170
      act_iter++;
      if (act_iter >= prm.num_iterations) stop = true;
      if (act_iter - numItersLastImprovement >= prm.num_iterations_unimproved) stop = true;
      int k = 1;
      while (k <= 2) {
         if (prm.max_time_ms > 0 && ms() >= prm.max_time_ms) stop = true;         // :189-193 (does not leave the k loop)
         actual = tour;                                                           // :195-197
         const double rewardBefore = actual.getReward(), lengthBefore = actual.getPathLength();
         if (k == 1) actual.pathInsert(pb.budget, rng);                           // shake, :328-341
         else actual.pathExchange(pb.budget, rng);
         if (k == 1) r.res.proposals += actual.insertRandom(nvns * nvns, pb.budget, prm.lower_bound_skip, rng).proposals;   // :346
         else r.res.proposals += actual.exchangeRandom(nvns * nvns, pb.budget, prm.lower_bound_skip, rng).proposals;
         const double newReward = actual.getReward(), newLength = actual.getPathLength();
         if (newReward > rewardBefore || (newReward == rewardBefore && newLength < lengthBefore)) {   // :213
            timeLastImprovement = ms();
            numItersLastImprovement = act_iter;
            r.impr_reward.push_back(newReward);
            r.impr_length.push_back(newLength);
            r.impr_ms.push_back((double)timeLastImprovement);
            tour = actual;                                                        // :227-228
            const double calc = tour.getPathLengthCalculate();
            if (calc > pb.budget || newLength > pb.budget || calc - newLength > 0.01)               // :229-237
               throw vnsgpu::Error(VNSGPU_E_INCONSISTENT, "badly calculated length in new solution");
            k = 1;
         } else
            k++;
      }
   }
   r.res.num_iters = act_iter;
   r.res.num_iters_last_impr = numItersLastImprovement;
   r.res.time_last_impr_ms = (double)timeLastImprovement;
   r.res.time_ms = (double)ms();
   r.res.n_improvements = (int)r.impr_reward.size();
   r.ctime_ms = ms();
   finish_record(r, tour);
   return r;
}

RunRecord iterate_device(vnsgpu::Engine &eng, const Config &cfg, const sopvns::Problem &pb, uint32_t seed, vnsgpu::VnsSopPath &tour)
{
   const auto t0 = std::chrono::steady_clock::now();
   vnsgpu_vns_params prm = params_of(cfg, pb, seed);
   tour.set(pb.initial_vns(), 0);
   RunRecord r;
   r.impr_reward.resize(prm.max_improvements);
   r.impr_length.resize(prm.max_improvements);
   r.impr_ms.resize(prm.max_improvements);
   eng.check(vnsgpu_vns_run(eng.ctx(), tour.handle(), &prm, &r.res, r.impr_reward.data(), r.impr_length.data(), r.impr_ms.data()));
   const int n = std::min(r.res.n_improvements, prm.max_improvements);
   r.impr_reward.resize(n);
   r.impr_length.resize(n);
   r.impr_ms.resize(n);
   r.ctime_ms = (long)std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count();
   finish_record(r, tour);
   return r;
}

int run(int argc, char **argv)
{
   Config cfg;
   for (int i = 1; i < argc; ++i) {   // --key=value | --key value | -h
      std::string a(argv[i]);
      if (a == "-h" || a == "--help") {
         std::cout << "sop_vns_b200 --problem=FILE.sop [--seed=N] [--device=N] [--gpus=N] [--device-loop=0|1] [--batch=N] [--output=DIR]\n"
                      "             [--num-iterations=N] [--num-iterations-unimproved=N] [--maximal-calculation-time-sec=N]\n"
                      "             [--lower-bound-above-budget-then-skip=0|1] [--initial-greedy-solution=0|1] [--name=S]\n"
                      "             [--config-file=sop.cfg]   (all keys of the reference's sop.cfg are accepted)\n";
         return 0;
      }
      while (!a.empty() && a[0] == '-') a.erase(0, 1);
      std::string k = a, v;
      const size_t eq = a.find('=');
      if (eq != std::string::npos) { k = a.substr(0, eq); v = a.substr(eq + 1); }
      else if (i + 1 < argc) v = argv[++i];
      if (!cfg.known(k) && !gui_key(k)) { std::cerr << "unrecognised option '--" << k << "'" << std::endl; return 2; }
      cfg.store(k, v);
   }
   if (cfg.is_set("config-file")) load_config_file(cfg, cfg.str("config-file"), true);
   else load_config_file(cfg, "sop.cfg", false);
   if (cfg.str("gui") != "none") { std::cerr << "sop_vns_b200 is headless: only gui=none is available" << std::endl; return 2; }
   if (cfg.str("sop-type") != "sop") { std::cerr << "sop-type=" << cfg.str("sop-type") << " is not available: pre-sampled .sop files only" << std::endl; return 2; }
   if (cfg.str("sop-solver") != "vns") { std::cerr << "sop-solver=" << cfg.str("sop-solver") << " is not available (no ILP solver)" << std::endl; return 2; }
   if (cfg.str("problem").empty()) { std::cerr << "Problem file must be specified (--problem=FILE.sop)" << std::endl; return 2; }

   const sopvns::SopFile sf = sopvns::parse_sop(cfg.str("problem"));
   Solver S{cfg, sopvns::build_problem(sf), cfg.str("problem"), cfg.str("name"), cfg.str("output"), cfg.num("iteration")};
   if (S.name == "no_name") S.name = S.pb.name;   // vns_sop_solver.cpp:33-37
   const sopvns::Problem &pb = S.pb;

   if (cfg.flag("dump-problem")) {   // the loader's result as one JSON line (tests; needs no GPU)
      auto arr = [](const auto &v) { std::ostringstream o; o << "["; for (size_t i = 0; i < v.size(); ++i) o << (i ? "," : "") << std::setprecision(17) << v[i]; o << "]"; return o.str(); };
      std::cout << "{\"name\":\"" << pb.name << "\",\"C\":" << pb.C << ",\"N\":" << pb.N << ",\"start\":" << pb.start << ",\"goal\":" << pb.goal
                << ",\"budget\":" << pb.budget << ",\"rule\":" << pb.rule << ",\"max_reward\":" << pb.max_reward << ",\"m\":" << arr(pb.m)
                << ",\"reward\":" << arr(pb.reward) << ",\"node_id\":" << arr(pb.node_id) << ",\"orig_cluster_id\":" << arr(pb.orig_cluster_id)
                << ",\"xy\":" << arr(pb.xy) << ",\"W_sum\":" << std::setprecision(17) << [&] { double t = 0; for (double w : pb.W) t += w; return t; }()
                << ",\"W_size\":" << pb.W.size() << "}" << std::endl;
      return 0;
   }
   vnsgpu::Engine eng(cfg.num("device"));
   eng.setClusters(pb.m, pb.reward, pb.start, pb.goal);
   if (!pb.xy.empty()) eng.buildTableEuclid(pb.xy, pb.rule);   // getSOPFromNodes on the device
   else eng.setTableDense(pb.W);                               // getSOPFromWeightMatrix

   const uint32_t seed = cfg.is_set("seed") ? (uint32_t)std::strtoul(cfg.str("seed").c_str(), 0, 10) : (uint32_t)time(nullptr);
   const int batch = cfg.num("batch");
   const bool device_loop = cfg.flag("device-loop");
   vnsgpu::VnsSopPath tour(eng, pb.initial_vns(), 0);
   if (batch < 0) {
      RunRecord r = device_loop ? iterate_device(eng, cfg, pb, seed, tour) : iterate_host_loop(eng, cfg, pb, seed, tour);
      S.save(r, sf);
      S.append_log(r);
      std::cout << "found tour with reward " << r.reward << " and length " << r.length << " out of " << pb.budget << " budget" << std::endl;
      return 0;
   }
   // --batch=N: iterations max(iteration,0) .. N-1 (algorithm.cc:147-157).  The reference runs them one after the
   // other on one rand() stream; here run i is seeded seed + i and all runs share one launch.
   const int first = std::max(S.iter, 0), nruns = batch - first;
   if (nruns <= 0) return 0;
   std::vector<RunRecord> recs(nruns);
   if (device_loop) {
      std::vector<uint32_t> seeds(nruns);
      for (int i = 0; i < nruns; ++i) seeds[i] = seed + (uint32_t)(first + i);
      // one launch per device: runs [lo, hi) on engine e (its own context, table and tour)
      auto run_on = [&](vnsgpu::Engine &e, vnsgpu::VnsSopPath &t0path, int lo, int hi) {
         const int cnt = hi - lo;
         if (cnt <= 0) return;
         std::vector<vnsgpu_vns_result> res(cnt);
         std::vector<int> tours(cnt, -1);
         vnsgpu_vns_params prm = params_of(cfg, pb, seed);
         int best = -1;
         const auto t0 = std::chrono::steady_clock::now();
         vnsgpu_status st = vnsgpu_vns_run_batch(e.ctx(), t0path.handle(), cnt, seeds.data() + lo, &prm, res.data(), &best, tours.data());
         if (st == VNSGPU_E_TOO_LARGE) {   // instance beyond the one-block-per-run kernel: run the seeds one by one (any other status is a real failure)
            for (int i = 0; i < cnt; ++i) recs[lo + i] = iterate_device(e, cfg, pb, seeds[lo + i], t0path);
            return;
         }
         e.check(st);
         const long wall = (long)std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count();
         for (int i = 0; i < cnt; ++i) {
            RunRecord &r = recs[lo + i];
            r.res = res[i];
            r.ctime_ms = wall;   // all runs of a device share the launch
            std::vector<int> c, s;
            int T = 0;
            e.check(vnsgpu_path_get(e.ctx(), tours[i], &T, 0, 0, 0, &r.length, 0, &r.reward));
            c.resize(T + 2);
            s.resize(T + 2);
            e.check(vnsgpu_path_get_solution(e.ctx(), tours[i], c.data(), s.data(), 0));
            for (int j = 0; j < T + 2; ++j) r.path.push_back(vnsgpu::IndexSOP{c[j], s[j]});
            vnsgpu_path_destroy(e.ctx(), tours[i]);
         }
      };
      const int gpus = std::max(1, std::min(cfg.num("gpus"), nruns));
      if (gpus == 1) {
         run_on(eng, tour, 0, nruns);
      } else {
         std::vector<std::thread> th;
         std::vector<std::string> errs(gpus);
         for (int g = 0; g < gpus; ++g) {
            const int lo = (int)((long long)nruns * g / gpus), hi = (int)((long long)nruns * (g + 1) / gpus);
            th.emplace_back([&, g, lo, hi] {
               try {
                  if (g == 0) { run_on(eng, tour, lo, hi); return; }
                  vnsgpu::Engine e(cfg.num("device") + g);
                  e.setClusters(pb.m, pb.reward, pb.start, pb.goal);
                  if (!pb.xy.empty()) e.buildTableEuclid(pb.xy, pb.rule);
                  else e.setTableDense(pb.W);
                  vnsgpu::VnsSopPath t(e, pb.initial_vns(), 0);
                  run_on(e, t, lo, hi);
               } catch (const std::exception &ex) { errs[g] = ex.what(); }
            });
         }
         for (auto &t : th) t.join();
         for (int g = 0; g < gpus; ++g)
            if (!errs[g].empty()) throw std::runtime_error("device " + std::to_string(cfg.num("device") + g) + ": " + errs[g]);
      }
   } else {
      for (int i = 0; i < nruns; ++i) recs[i] = iterate_host_loop(eng, cfg, pb, seed + (uint32_t)(first + i), tour);
   }
   for (int i = 0; i < nruns; ++i) {
      S.iter = first + i;
      S.save(recs[i], sf);
      S.append_log(recs[i]);
   }
   return 0;
}

} // namespace

int main(int argc, char **argv)
{
   try {
      return run(argc, argv);
   } catch (const vnsgpu::Error &e) {   // the reference prints and exit(1)s
      std::cerr << "ERROR: vnsgpu: " << e.what() << std::endl;
      return 1;
   } catch (const std::exception &e) {
      std::cerr << "ERROR: " << e.what() << std::endl;
      return 1;
   }
}
