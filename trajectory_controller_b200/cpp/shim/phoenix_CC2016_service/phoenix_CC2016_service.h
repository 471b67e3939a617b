// HEADLESS SHIM of the Phoenix CC2016 service (un-vendored): only driveMode() is read
// (src/trajectory_point_follower.cpp:35-36,70).
#pragma once
namespace phoenix_CC2016_service {
enum class CCDriveMode { IDLE, FOH, FMH, PARKING };
class Phoenix_CC2016Service {
    CCDriveMode mode_ = CCDriveMode::FMH;
public:
    CCDriveMode driveMode() const { return mode_; }
    void setDriveMode(CCDriveMode m) { mode_ = m; }
};
}  // namespace phoenix_CC2016_service
