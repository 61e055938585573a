// baspacho/baspacho/DebugMacros.h — facade of the assertion macros the reference uses from BaSpaCho
// (lib/small_thing/SoftLoss.h:10,67; viba/common/FindOrDie.h:10,14): throwing checks, std::runtime_error.
#pragma once
#include <sstream>
#include <stdexcept>

#define BASPACHO_CHECK(cond)                                                                       \
  do {                                                                                             \
    if (!(cond)) {                                                                                 \
      std::ostringstream baspacho_os_;                                                             \
      baspacho_os_ << "[" << __FILE__ << ":" << __LINE__ << "] Check failed: " #cond;              \
      throw std::runtime_error(baspacho_os_.str());                                                \
    }                                                                                              \
  } while (0)
#define BASPACHO_CHECK_OP(a, op, b)                                                                \
  do {                                                                                             \
    if (!((a)op(b))) {                                                                             \
      std::ostringstream baspacho_os_;                                                             \
      baspacho_os_ << "[" << __FILE__ << ":" << __LINE__ << "] Check failed: " #a " " #op " " #b " (" << (a) << " vs " << (b) << ")"; \
      throw std::runtime_error(baspacho_os_.str());                                                \
    }                                                                                              \
  } while (0)
#define BASPACHO_CHECK_EQ(a, b) BASPACHO_CHECK_OP(a, ==, b)
#define BASPACHO_CHECK_NE(a, b) BASPACHO_CHECK_OP(a, !=, b)
#define BASPACHO_CHECK_LT(a, b) BASPACHO_CHECK_OP(a, <, b)
#define BASPACHO_CHECK_LE(a, b) BASPACHO_CHECK_OP(a, <=, b)
#define BASPACHO_CHECK_GT(a, b) BASPACHO_CHECK_OP(a, >, b)
#define BASPACHO_CHECK_GE(a, b) BASPACHO_CHECK_OP(a, >=, b)
#define BASPACHO_CHECK_NOTNULL(p) BASPACHO_CHECK((p) != nullptr)
