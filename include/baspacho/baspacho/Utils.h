// baspacho/baspacho/Utils.h — facade: BaSpaCho::timeStamp() (used by the reference's timing printouts)
#pragma once
#include <chrono>
#include <iomanip>
#include <sstream>
#include <string>

namespace BaSpaCho {

inline std::string timeStamp() {
  using namespace std::chrono;
  const auto now = system_clock::now();
  const std::time_t t = system_clock::to_time_t(now);
  const auto ms = duration_cast<milliseconds>(now.time_since_epoch()).count() % 1000;
  std::tm tm{};
  localtime_r(&t, &tm);
  std::ostringstream os;
  os << std::put_time(&tm, "%H:%M:%S") << "." << std::setw(3) << std::setfill('0') << ms;
  return os.str();
}

}  // namespace BaSpaCho
