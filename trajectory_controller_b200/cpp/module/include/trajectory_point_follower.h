// TrajectoryPointController — the LMS plugin module of lms-org/trajectory_controller, re-hosted on the B200 engine.
//
// Drop-in surface (reference include/trajectory_point_follower.h:26-70): same class name and base class, the same four
// life-cycle overrides, the same data channels ("TRAJECTORY" in; "TRAJECTORY_POINT", "TRAJECTORY_DEBUG", "CAR" out),
// service ("PHOENIX_SERVICE") and config keys.  What changed underneath:
//   * getTrajectoryPoint()'s arc-length scan + interpolation runs in bmpc_extract_points (CUDA);
//   * mpcControllerTobi() builds the same A/B/C/Q/R and solves through dlib::mpc<2,2,4> — which, with
//     cpp/dlib_compat on the include path, is the GPU-backed facade (bmpc/mpc.h) instead of dlib's CPU solver.
// Builds against the real LMS headers when they are on the include path, else against cpp/shim (headless).
#pragma once
#include <vector>

#include "lms/module.h"
#include "lms/math/vertex.h"
#include "lms/math/lookup_table.h"
#include "lms/math/interpolation.h"
#include "street_environment/car.h"
#include "street_environment/trajectory.h"
#include "sensor_utils/pid_controller.h"
#include "phoenix_CC2016_service/phoenix_CC2016_service.h"

#include <dlib/control.h>

#include "bmpc.h"

class TrajectoryPointController : public lms::Module {
public:
    bool initialize() override;
    bool deinitialize() override;
    bool cycle() override;

    ~TrajectoryPointController() override;

private:
    typedef lms::math::LookupTable<float, lms::math::LookupTableOrder::ASC> Lookup;
    static constexpr size_t MPC_HORIZON = 4;       // reference header :48
    static constexpr int STATES = 2, CONTROLS = 2; // y/phi and front/rear steering (reference .cpp:303-304)

    struct MpcParameters {
        double weight_y, weight_phi, weight_steeringFront, weight_steeringRear, stepSize;
    } mpcParameters;

    void configsChanged() override;
    street_environment::TrajectoryPoint getTrajectoryPoint(const float distanceToPoint);
    void mpcControllerTobi(double v, double delta_y, double delta_phi, double* steering_front, double* steering_rear);

    // cycle() is split by controller type; each fills the steering / speed fields of `state`
    bool stepTobiMpc(phoenix_CC2016_service::CCDriveMode mode, street_environment::CarCommand::State& state);
    bool stepPid(street_environment::CarCommand::State& state);
    void applyCrossingSlowDown(street_environment::TrajectoryPoint& point);
    void setIndicators(street_environment::CarCommand::State& state);

    Lookup m_mpcLookupVelocity;
    Lookup m_trajectoryPointDistanceLookup;
    sensor_utils::PID slowDownCar;          // slows the car down when the trajectory ends at a stop point
    sensor_utils::PID pidControllerFront;
    sensor_utils::PID pidControllerRear;

    double l = 0.21;                        // wheelbase [m] (reference header :47)
    dlib::matrix<double, CONTROLS, 1> lower, upper;

    lms::ReadDataChannel<street_environment::Trajectory> trajectory;
    lms::WriteDataChannel<street_environment::CarCommand> car;
    lms::WriteDataChannel<street_environment::TrajectoryPoint> debugging_trajectoryPoint;
    lms::WriteDataChannel<street_environment::Trajectory> trajectoryDebug;

    lms::Time turnStart;
    bool isTurn = false;

    // engine slot used for trajectory-point extraction (the MPC facade owns its own slot)
    bmpc_t* engine_ = nullptr;
    std::vector<float> px_, py_, dx_, dy_, vel_;   // SoA staging of the polyline
};
