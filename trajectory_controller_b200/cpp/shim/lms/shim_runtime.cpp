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
