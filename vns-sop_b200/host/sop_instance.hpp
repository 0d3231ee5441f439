// Host side of the drop-in: the reference's `.sop` text format and the loader's start/goal renumbering,
// producing the flat cluster-contiguous model that include/vnsgpu.h uploads.  Runs once per instance.
//
// Restates
//   DatasetLoaderSOP::loadDataset            sources/dataset_loader_sop.cpp:39-287   -> parse_sop
//   SOPLoader::getSOPDefinition (sop branch) sources/sop_loader.cpp:102-173          -> build_problem
//   SOPLoader::fixClusterNodeRenumbering     sources/sop_loader.cpp:185-250          -> Problem::orig_cluster_id
// (the O(N^2) table build of getSOPFromNodes / getSOPFromWeightMatrix, sop_loader.cpp:382-499, is NOT
// done here: it is vnsgpu_build_table_euclid / vnsgpu_set_table_dense on the device).
#ifndef VNSGPU_HOST_SOP_INSTANCE_HPP
#define VNSGPU_HOST_SOP_INSTANCE_HPP
#include <cstdlib>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "vnsgpu.h"

namespace sopvns {

struct SopFile {
   std::string name = "no_name", comment, edge_weight_type = "CEIL_2D";
   int dimension = 0, tmax = 0, start_set = 0, end_set = 0;
   double dubins_radius = 0;
   std::vector<double> coords;                // dimension x 2 (NODE_COORD_SECTION)
   std::vector<double> weights;               // dimension x dimension (EDGE_WEIGHT_SECTION, FULL_MATRIX)
   std::vector<double> set_profit;            // per set
   std::vector<std::vector<int>> set_nodes;   // 0-based node ids per set
   std::vector<double> centers;               // sets x 2 (GTSP_SET_CENTER_COORD_SECTION, DOP files)
};

struct Problem {
   std::string name;
   int C = 0, N = 0, start = 0, goal = 0;
   double budget = 0;
   int rule = -1;                             // VNSGPU_RULE_* for coordinate instances, -1 = EXPLICIT
   std::vector<int> m, off;                   // samples per cluster, prefix sums (C+1)
   std::vector<double> reward;
   std::vector<int> node_id;                  // original 0-based node id of every sample
   std::vector<int> orig_cluster_id;          // set id printed in RESULT_CLUSTER_IDS
   std::vector<double> xy;                    // N x 2, sample order (coordinate instances)
   std::vector<double> W;                     // N x N, sample order (EXPLICIT instances)
   double max_reward = 0;                     // MAX_ACHIEVABLE_REWARDS (vns_sop_solver.cpp:49-53)
   std::vector<int> initial_vns() const
   {  // vnsVector of iterate(): all clusters but start and goal (vns_sop_solver.cpp:91-99,119-122)
      std::vector<int> v;
      for (int c = 0; c < C; ++c)
         if (c != start && c != goal) v.push_back(c);
      return v;
   }
};

namespace detail {
inline std::string trim(const std::string &s)
{
   size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n");
   return a == std::string::npos ? std::string() : s.substr(a, b - a + 1);
}
inline std::vector<std::string> split_ws(const std::string &s)
{
   std::vector<std::string> t;
   std::istringstream is(s);
   for (std::string w; is >> w;) t.push_back(w);
   return t;
}
} // namespace detail

// Keys are split at the first ':'; NODE_COORD_SECTION / EOF are bare lines; the weight section is a token
// stream filling rows of DIMENSION entries (dataset_loader_sop.cpp:73-262).
inline SopFile parse_sop(const std::string &path)
{
   std::ifstream in(path.c_str());
   if (!in) throw std::runtime_error("Cannot open problem file '" + path + "'");
   SopFile sf;
   enum { NONE, COORDS, WEIGHTS, SETS, CENTERS } section = NONE;
   for (std::string raw; std::getline(in, raw);) {
      const std::string line = detail::trim(raw);
      if (line.empty()) continue;
      const size_t colon = line.find(':');
      if (colon != std::string::npos) {
         const std::string key = detail::trim(line.substr(0, colon)), val = detail::trim(line.substr(colon + 1));
         if (key == "NAME") sf.name = val;
         else if (key == "COMMENT") sf.comment = val;
         else if (key == "DIMENSION") sf.dimension = std::atoi(val.c_str());
         else if (key == "TMAX") sf.tmax = std::atoi(val.c_str());   // std::stoi in the reference (:110)
         else if (key == "START_SET") sf.start_set = std::atoi(val.c_str());
         else if (key == "END_SET") sf.end_set = std::atoi(val.c_str());
         else if (key == "EDGE_WEIGHT_TYPE") sf.edge_weight_type = val;
         else if (key == "DUBINS_RADIUS") sf.dubins_radius = std::atof(val.c_str());
         else if (key == "GTSP_SET_SECTION") section = SETS;
         else if (key == "GTSP_SET_CENTER_COORD_SECTION") section = CENTERS;
         else if (key == "EDGE_WEIGHT_SECTION") section = WEIGHTS;
         continue;
      }
      if (line == "NODE_COORD_SECTION") section = COORDS;
      else if (line == "EOF") section = NONE;
      else if (line == "GTSP_SET_SECTION") section = SETS;
      else if (line == "EDGE_WEIGHT_SECTION") section = WEIGHTS;
      else if (section == COORDS) {
         const std::vector<std::string> t = detail::split_ws(line);
         if (t.size() < 3) throw std::runtime_error("bad coordinate line: " + line);
         sf.coords.push_back(std::atof(t[1].c_str()));
         sf.coords.push_back(std::atof(t[2].c_str()));
      } else if (section == WEIGHTS) {
         for (const std::string &w : detail::split_ws(line)) sf.weights.push_back(std::atof(w.c_str()));
      } else if (section == SETS) {
         const std::vector<std::string> t = detail::split_ws(line);
         if (t.size() < 3) throw std::runtime_error("set line needs >= 3 tokens: " + line);
         if (std::atoi(t[0].c_str()) != (int)sf.set_nodes.size()) throw std::runtime_error("wrong cluster_id order in " + path);
         sf.set_profit.push_back(std::atof(t[1].c_str()));
         std::vector<int> nodes;
         for (size_t k = 2; k < t.size(); ++k) {
            const int id = std::atoi(t[k].c_str()) - 1;   // the file numbers nodes from 1; "-1" terminators drop out
            if (id >= 0) nodes.push_back(id);
         }
         sf.set_nodes.push_back(nodes);
      } else if (section == CENTERS) {
         const std::vector<std::string> t = detail::split_ws(line);
         if (t.size() < 3) throw std::runtime_error("bad centre line: " + line);
         sf.centers.push_back(std::atof(t[1].c_str()));
         sf.centers.push_back(std::atof(t[2].c_str()));
      }
   }
   if (!sf.coords.empty()) sf.dimension = (int)sf.coords.size() / 2;
   if (!sf.weights.empty() && (int64_t)sf.weights.size() != (int64_t)sf.dimension * sf.dimension)
      throw std::runtime_error("edge weight matrix size != DIMENSION^2 in " + path);
   if (sf.coords.empty() && sf.weights.empty()) throw std::runtime_error("no coordinates and no weight matrix in " + path);
   if (sf.set_nodes.empty()) throw std::runtime_error("no GTSP_SET_SECTION in " + path);
   return sf;
}

// sop-type=sop branch of getSOPDefinition with change_end_index_to_last = true (sop_loader.cpp:118-156):
// the goal set moves (or, when START_SET == END_SET, is duplicated) to the last index, the start set to 0.
inline Problem build_problem(const SopFile &sf)
{
   struct Cl { double profit; std::vector<int> nodes; int orig; };
   const int S0 = (int)sf.set_nodes.size();
   if (sf.start_set < 0 || sf.start_set >= S0 || sf.end_set < 0 || sf.end_set >= S0)
      throw std::runtime_error("START_SET / END_SET out of range");
   std::vector<Cl> cl;
   for (int c = 0; c < S0; ++c) cl.push_back(Cl{sf.set_profit[c], sf.set_nodes[c], c});
   const int s = sf.start_set, g = sf.end_set;
   const Cl end_cluster = cl[g], last_cluster = cl.back();
   if (g == s) cl.push_back(end_cluster);
   else cl[g] = last_cluster;
   cl.back() = end_cluster;
   if (s != 0) std::swap(cl[s], cl[0]);
   Problem pb;
   pb.name = sf.name;
   pb.C = (int)cl.size();
   pb.start = 0;
   pb.goal = pb.C - 1;
   pb.budget = (double)sf.tmax;
   pb.rule = sf.edge_weight_type == "EUC_2D" ? VNSGPU_RULE_ROUND : sf.edge_weight_type == "CEIL_2D" ? VNSGPU_RULE_CEIL : -1;
   pb.off.push_back(0);
   for (const Cl &c : cl) {
      if (c.nodes.empty()) throw std::runtime_error("empty set in GTSP_SET_SECTION");
      pb.m.push_back((int)c.nodes.size());
      pb.off.push_back(pb.off.back() + (int)c.nodes.size());
      pb.reward.push_back(c.profit);
      pb.orig_cluster_id.push_back(c.orig);
      for (int id : c.nodes) {
         if (id >= sf.dimension) throw std::runtime_error("node id beyond DIMENSION");
         pb.node_id.push_back(id);
      }
   }
   pb.N = pb.off.back();
   if (!sf.coords.empty()) {   // is_with_nodes -> getSOPFromNodes (sop_loader.cpp:163-166); it exits on other types (:483-485)
      if (pb.rule < 0) throw std::runtime_error("unknown edge weight type " + sf.edge_weight_type);
      pb.xy.resize((size_t)pb.N * 2);
      for (int i = 0; i < pb.N; ++i) {
         pb.xy[2 * i] = sf.coords[2 * (size_t)pb.node_id[i]];
         pb.xy[2 * i + 1] = sf.coords[2 * (size_t)pb.node_id[i] + 1];
      }
   } else {
      if (sf.weights.empty()) throw std::runtime_error("EDGE_WEIGHT_TYPE " + sf.edge_weight_type + " needs EDGE_WEIGHT_SECTION");
      pb.W.resize((size_t)pb.N * pb.N);
      for (int i = 0; i < pb.N; ++i)
         for (int j = 0; j < pb.N; ++j) pb.W[(size_t)i * pb.N + j] = sf.weights[(size_t)pb.node_id[i] * sf.dimension + pb.node_id[j]];
      pb.rule = -1;
   }
   // maximalRewardAll sums the first sample's reward of EVERY cluster (vns_sop_solver.cpp:49-53)
   for (int c = 0; c < pb.C; ++c) pb.max_reward += pb.reward[c];
   return pb;
}

} // namespace sopvns
#endif
