// Plugin entry: the LMS loader dlopen()s the MODULE library and asks this factory for the module object
// (same registration as the reference's src/interface.cpp:1-3).
#include "trajectory_point_follower.h"

LMS_MODULE_INTERFACE(TrajectoryPointController)
