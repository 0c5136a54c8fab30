/* wall clock in seconds (reference include/dmlc/timer.h:27-47) */
#ifndef DMLC_TIMER_H_
#define DMLC_TIMER_H_
#include <chrono>
namespace dmlc {
inline double GetTime(void) {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
}  // namespace dmlc
#endif  // DMLC_TIMER_H_
