#pragma once
/* Levelled stderr logging: LEVEL[file:line]{rank} message.
   Parity: reference include/stencil/logging.hpp:8-53 — compile-time level STENCIL_OUTPUT_LEVEL
   (0 FATAL .. 5 SPEW, default 3 = INFO). Additionally the level can be LOWERED at run time with the environment
   variable STENCIL_LOG_LEVEL (benchmarks use it to silence setup chatter without a rebuild). */
#include <cstdlib>
#include <iostream>
#include <sstream>

#include "stencil/mpi.hpp"

#ifndef STENCIL_OUTPUT_LEVEL
#define STENCIL_OUTPUT_LEVEL 3
#endif

namespace stencil {
namespace log {
inline int runtime_level() {
  static const int lvl = [] {
    const char *e = std::getenv("STENCIL_LOG_LEVEL");
    return e ? std::atoi(e) : STENCIL_OUTPUT_LEVEL;
  }();
  return lvl;
}
} // namespace log
} // namespace stencil

#define STENCIL_LOG_AT(lvl, name, x)                                                                                   \
  do {                                                                                                                 \
    if ((lvl) <= STENCIL_OUTPUT_LEVEL && (lvl) <= ::stencil::log::runtime_level()) {                                   \
      std::ostringstream stencil_log_ss_;                                                                              \
      stencil_log_ss_ << name "[" << __FILE__ << ":" << __LINE__ << "]{" << mpi::world_rank() << "} " << x << "\n";    \
      std::cerr << stencil_log_ss_.str();                                                                              \
    }                                                                                                                  \
  } while (0)

#define LOG_SPEW(x) STENCIL_LOG_AT(5, "SPEW", x)
#define LOG_DEBUG(x) STENCIL_LOG_AT(4, "DEBUG", x)
#define LOG_INFO(x) STENCIL_LOG_AT(3, "INFO", x)
#define LOG_WARN(x) STENCIL_LOG_AT(2, "WARN", x)
#define LOG_ERROR(x) STENCIL_LOG_AT(1, "ERROR", x)
#define LOG_FATAL(x)                                                                                                   \
  do {                                                                                                                 \
    std::ostringstream stencil_log_ss_;                                                                                \
    stencil_log_ss_ << "FATAL[" << __FILE__ << ":" << __LINE__ << "]{" << mpi::world_rank() << "} " << x << "\n";      \
    std::cerr << stencil_log_ss_.str();                                                                                \
    std::exit(1);                                                                                                      \
  } while (0)
