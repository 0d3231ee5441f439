// Build-time stand-in for log4cxx (absent in this image) used ONLY when compiling the
// unmodified reference sources from /root/reference into oracle/_ref/ (test infrastructure).
// The reference logs through LOG4CXX_* macros on a global logger (crl/logging.h:17-20).
#pragma once
#include <algorithm>
#include <unistd.h>
#include <iostream>
#include <sstream>
#include <string>
namespace log4cxx {
struct Logger {};
typedef Logger* LoggerPtr;
}
extern int oracle_ref_verbosity;  // 0 = silent, 1 = errors, 2 = everything
#define ORACLE_REF_LOG(lvl, tag, expr)                                   \
   do {                                                                  \
      if (oracle_ref_verbosity >= (lvl)) {                               \
         std::ostringstream oss__;                                       \
         oss__ << expr;                                                  \
         std::cerr << tag << oss__.str() << std::endl;                   \
      }                                                                  \
   } while (0)
#define LOG4CXX_INFO(logger, expr) ORACLE_REF_LOG(2, "INFO ", expr)
#define LOG4CXX_DEBUG(logger, expr) ORACLE_REF_LOG(2, "DEBUG ", expr)
#define LOG4CXX_WARN(logger, expr) ORACLE_REF_LOG(1, "WARN ", expr)
#define LOG4CXX_ERROR(logger, expr) ORACLE_REF_LOG(1, "ERROR ", expr)
#define LOG4CXX_FATAL(logger, expr) ORACLE_REF_LOG(1, "FATAL ", expr)
