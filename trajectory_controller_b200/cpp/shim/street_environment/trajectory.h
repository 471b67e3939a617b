// HEADLESS SHIM of street_environment::Trajectory / TrajectoryPoint (un-vendored) — the
// payload of the TRAJECTORY / TRAJECTORY_POINT / TRAJECTORY_DEBUG channels (src/trajectory_point_follower.cpp:9-13).
#pragma once
#include <vector>
#include "lms/math/vertex.h"
namespace street_environment {
struct TrajectoryPoint {
    lms::math::vertex2f position;     // vehicle frame: x forward, y left
    lms::math::vertex2f directory;    // unit tangent
    float velocity;
    float distanceToMiddleLane;
    TrajectoryPoint() : velocity(0), distanceToMiddleLane(0) {}
    TrajectoryPoint(lms::math::vertex2f p, lms::math::vertex2f d, float v, float dm = 0)
        : position(p), directory(d), velocity(v), distanceToMiddleLane(dm) {}
    bool isRight() const { return distanceToMiddleLane < 0; }
};
class Trajectory : public std::vector<TrajectoryPoint> {
public:
    using std::vector<TrajectoryPoint>::vector;
};
}  // namespace street_environment
