// TEST INFRASTRUCTURE ONLY: headless, seedable entry point for the UNMODIFIED reference solver
// (replaces sources/sop.cpp:319-406, which needs boost::program_options + log4cxx + a GUI).
//   sop_vns_ref --problem=X.sop [--seed=N] [--verbose=0|1|2] [--key=value ...]
// Output: the reference's own results.log / path.txt / ... under --output (default ./).
#include <cstring>
#include <iostream>

#include "ref_common.h"
#include "vns_sop_solver.h"

int oracle_ref_verbosity = 0;
namespace crl {
CLoggerPtr logger = 0;
CLoggerPtr getLogger(const char *) { return 0; }
void shutdownLogger(void) {}
}
extern bool randSeeded;  // sources/math_common.cpp:9 (non-static global)

int main(int argc, char **argv)
{
   crl::CConfig cfg;
   oracle_ref::register_config(cfg);
   std::string problem;
   for (int i = 1; i < argc; ++i) {
      std::string a(argv[i]);
      while (!a.empty() && a[0] == '-') a.erase(0, 1);
      size_t eq = a.find('=');
      if (eq == std::string::npos) continue;
      std::string k = a.substr(0, eq), v = a.substr(eq + 1);
      if (k == "problem") problem = v;
      else if (k == "seed") { randSeeded = true; srand((unsigned)std::atol(v.c_str())); }
      else if (k == "verbose") oracle_ref_verbosity = std::atoi(v.c_str());
      else if (k == "gui") {}
      else if (!oracle_ref::set_option(cfg, k, v)) {
         std::cerr << "unknown option " << k << std::endl;
         return 2;
      }
   }
   if (problem.empty()) { std::cerr << "--problem=... required" << std::endl; return 2; }
   op::VnsSopSolver solver(cfg, problem);
   solver.setCanvas(0);
   solver.solve();
   return 0;
}
