// HEADLESS SHIM of sensor_utils::PID (un-vendored): set(Kp,Ki,Kd,dt), pid(error), reset().
// Textbook discrete PID; PARITY UNPINNED.  Only the host-side glue uses it (crossing slow-down, PID fallback).
#pragma once
namespace sensor_utils {
class PID {
    float kp_ = 1, ki_ = 0, kd_ = 0, dt_ = 0.01f, integral_ = 0, prev_ = 0;
    bool first_ = true;
public:
    void set(float kp, float ki, float kd, float dt) { kp_ = kp; ki_ = ki; kd_ = kd; dt_ = dt; }
    void reset() { integral_ = 0; prev_ = 0; first_ = true; }
    float pid(float error)
    {
        integral_ += error * dt_;
        const float deriv = first_ ? 0.0f : (error - prev_) / dt_;
        prev_ = error; first_ = false;
        return kp_ * error + ki_ * integral_ + kd_ * deriv;
    }
};
}  // namespace sensor_utils
