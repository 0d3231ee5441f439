// TEST INFRASTRUCTURE ONLY (oracle/_ref): glue compiled together with the UNMODIFIED reference
// sources in /root/reference/sources.  Nothing here is on the product path.
//
// Provides what sources/sop.cpp (not compilable here: boost::program_options, log4cxx appenders)
// would provide: the global crl::logger and the option registry of addCommonConfig
// (sources/sop.cpp:87-188), restated with the stock sop.cfg values (sources/sop.cfg:28-42).
#pragma once
#include <cstdlib>
#include <fstream>
#include <string>

#include <crl/config.h>
#include <crl/logging.h>

namespace oracle_ref {

// keys the VNS solver / CAlgorithm read (vns_sop_solver.cpp:25-71, algorithm.cc:72-84,
// sop_loader.cpp:43,121) with defaults = stock sop.cfg, gui=none
inline void register_config(crl::CConfig &c)
{
   c.add<std::string>("output", "", "./");
   c.add<std::string>("results", "", "results.log");
   c.add<std::string>("info", "", "info.txt");
   c.add<std::string>("settings", "", "settings.txt");
   c.add<std::string>("result-path", "", "path.txt");
   c.add<std::string>("result-canvas-output", "", "path");
   c.add<std::string>("result-canvas-suffixes", "", "");
   c.add<std::string>("name", "", "no_name");
   c.add<int>("iteration", "", -1);
   c.add<int>("batch", "", -1);
   c.add<bool>("continue", "", false);
   c.add<bool>("save-results", "", true);
   c.add<bool>("save-info", "", true);
   c.add<bool>("save-settings", "", true);
   c.add<bool>("save-visual", "", false);
   c.add<bool>("verbose-result-log", "", false);
   c.add<double>("canvas-border", "", 1);
   c.add<bool>("save-targets", "", true);
   c.add<bool>("save-sampled-path", "", true);
   c.add<std::string>("targets", "", "targets.txt");
   c.add<std::string>("sampled-path", "", "sampled-path.txt");
   c.add<int>("num-iterations", "", 2000);
   c.add<int>("num-iterations-unimproved", "", 1000);
   c.add<bool>("use-rvns", "", false);
   c.add<bool>("draw-neighborhood-points", "", false);
   c.add<bool>("draw-path", "", false);
   c.add<bool>("draw-cluster-points", "", false);
   c.add<bool>("draw-targets-reward", "", false);
   c.add<int>("maximal-calculation-time-sec", "", 168000);
   c.add<bool>("lower-bound-above-budget-then-skip", "", true);
   c.add<bool>("generate-examples-of-operations", "", false);
   c.add<std::string>("sop-type", "", "sop");
   c.add<std::string>("sop-solver", "", "vns");
   c.add<bool>("initial-greedy-solution", "", true);
   c.add<std::string>("pic-dir", "", "detail");
   c.add<std::string>("pic-ext", "", "png");
   c.add<bool>("save-pic", "", false);
}

// "--key=value" / "key=value" with the type taken from the registered default
inline bool set_option(crl::CConfig &c, const std::string &key, const std::string &val)
{
   try { c.get<std::string>(key); c.set<std::string>(key, val); return true; } catch (...) {}
   try { c.get<int>(key); c.set<int>(key, std::atoi(val.c_str())); return true; } catch (...) {}
   try { c.get<bool>(key); c.set<bool>(key, val == "1" || val == "true"); return true; } catch (...) {}
   try { c.get<double>(key); c.set<double>(key, std::atof(val.c_str())); return true; } catch (...) {}
   return false;
}

} // namespace oracle_ref
