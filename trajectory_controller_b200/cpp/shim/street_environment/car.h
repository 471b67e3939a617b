// HEADLESS SHIM of street_environment::CarCommand (un-vendored) — payload of the CAR channel:
// a priority list of named states (src/trajectory_point_follower.cpp:37-61,114-125,286).
#pragma once
#include <string>
#include <vector>
namespace street_environment {
class CarCommand {
public:
    enum class StateType { NOT_DEFINED, IDLE, DRIVING, PARKING, RACE };
    struct State {
        StateType state;
        int priority;
        std::string name;
        float steering_front, steering_rear, targetSpeed, targetDistance;
        bool indicatorLeft, indicatorRight;
        State() : state(StateType::NOT_DEFINED), priority(0), steering_front(0), steering_rear(0), targetSpeed(0),
                  targetDistance(0), indicatorLeft(false), indicatorRight(false) {}
    };
    State* getState(const std::string& name)
    {
        for (State& s : states_) if (s.name == name) return &s;
        return nullptr;
    }
    void putState(const State& s)
    {
        if (State* old = getState(s.name)) *old = s; else states_.push_back(s);
    }
    void removeState(const std::string& name)
    {
        for (size_t i = 0; i < states_.size(); ++i) if (states_[i].name == name) { states_.erase(states_.begin() + i); return; }
    }
    // measured vehicle state (set by the sensor side; the controller only reads it)
    float velocity() const { return velocity_; }
    float steeringFront() const { return steeringFront_; }
    float steeringRear() const { return steeringRear_; }
    void setMeasured(float v, float sf, float sr) { velocity_ = v; steeringFront_ = sf; steeringRear_ = sr; }
    const std::vector<State>& states() const { return states_; }
private:
    std::vector<State> states_;
    float velocity_ = 0, steeringFront_ = 0, steeringRear_ = 0;
};
typedef CarCommand Car;
}  // namespace street_environment
