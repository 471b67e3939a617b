// Out-of-line pieces of the headless LMS shim (shim/lms/module.h).
#include <chrono>

#include "lms/module.h"

namespace lms {
const Time Time::ZERO(0);
Time Time::now()
{
    using namespace std::chrono;
    return Time(duration_cast<microseconds>(steady_clock::now().time_since_epoch()).count());
}
LogLine::~LogLine()
{
    if (live_) std::cerr << "[" << level_ << "] " << os_.str() << "\n";
}
Blackboard& Blackboard::instance()
{
    static Blackboard b;
    return b;
}
}  // namespace lms

// Stub of the missing andromeda solver (shim/andromeda.h): no motion.
extern "C" void call_andromeda(double*, double*, double*, double*, double*, double*, double, double*, double*, double,
                               double, double, double, double, double, double, double, double, double* v_star,
                               double* u_1_star, double* u_2_star)
{
    for (int i = 0; i < 10; ++i) { v_star[i] = 0; u_1_star[i] = 0; u_2_star[i] = 0; }
}
