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
    // Resample at a fixed arc-length spacing (used only by the out-of-scope mikMPC branch, reference .cpp:136).
    Trajectory getWithDistanceBetweenPoints(float spacing) const
    {
        Trajectory out;
        if (empty() || !(spacing > 0)) return out;
        out.push_back(front());
        float need = spacing;
        for (size_t i = 1; i < size(); ++i) {
            lms::math::vertex2f a = (*this)[i - 1].position;
            const lms::math::vertex2f b = (*this)[i].position;
            float seg = a.distance(b);
            while (seg >= need && seg > 0) {
                const lms::math::vertex2f dir = (b - a).normalize();
                a = a + dir * need;
                TrajectoryPoint p = (*this)[i];
                p.position = a;
                out.push_back(p);
                seg -= need;
                need = spacing;
            }
            need -= seg;
        }
        return out;
    }
};
}  // namespace street_environment
