// TrajectoryPointController on the B200 engine.  Behavioural contract: the reference's
// src/trajectory_point_follower.cpp (line numbers cited per function); the arithmetic of the hot path runs in
// libbmpc.so (CUDA), everything else is O(1) host glue.
#include "trajectory_point_follower.h"

#include <cmath>
#include <string>

namespace {
typedef street_environment::CarCommand::State CarState;
typedef street_environment::CarCommand::StateType CarStateType;
typedef phoenix_CC2016_service::CCDriveMode DriveMode;

// Start from the stored state of that name (if any) so fields this module does not write survive (:37-43, :55-61).
CarState inheritedState(street_environment::CarCommand& cmd, const char* name, int priority)
{
    CarState s;
    if (CarState* old = cmd.getState(name)) s = *old;
    s.name = name;
    s.priority = priority;
    return s;
}
}  // namespace

TrajectoryPointController::~TrajectoryPointController()
{
    if (engine_) bmpc_destroy(engine_);
}

// reference :8-28
bool TrajectoryPointController::initialize()
{
    trajectory = readChannel<street_environment::Trajectory>("TRAJECTORY");
    debugging_trajectoryPoint = writeChannel<street_environment::TrajectoryPoint>("TRAJECTORY_POINT");
    trajectoryDebug = writeChannel<street_environment::Trajectory>("TRAJECTORY_DEBUG");
    car = writeChannel<street_environment::CarCommand>("CAR");

    const double alpha_max = 22 * M_PI / 180;      // steering box (:16-18)
    lower = -alpha_max, -alpha_max;
    upper = alpha_max, alpha_max;

    configsChanged();
    isTurn = false;
    turnStart = lms::Time::ZERO;

    // One engine slot for the point extraction.  No GPU => the module refuses to start: there is no CPU fallback.
    if (!engine_) {
        const int rc = bmpc_create(&engine_, STATES, CONTROLS, (int)MPC_HORIZON, 1, BMPC_FP64, BMPC_MODE_EXACT, -1);
        if (rc != BMPC_OK) {
            logger.error("trajectory_point_controller") << "GPU engine unavailable: " << bmpc_last_error();
            engine_ = nullptr;
            return false;
        }
    }
    return true;
}

bool TrajectoryPointController::deinitialize()
{
    if (engine_) { bmpc_destroy(engine_); engine_ = nullptr; }
    return true;
}

// reference :291-299
void TrajectoryPointController::configsChanged()
{
    m_mpcLookupVelocity.vx = config().getArray<float>("mpcLookupVelocityX");
    m_mpcLookupVelocity.vy = config().getArray<float>("mpcLookupVelocityY");
    m_trajectoryPointDistanceLookup.vx = config().getArray<float>("trajectoryPointDistanceLookupX");
    m_trajectoryPointDistanceLookup.vy = config().getArray<float>("trajectoryPointDistanceLookupY");
    const float dt = config().get<float>("dt", 0.01);
    slowDownCar.set(config().get<float>("PID_Kp", 1), config().get<float>("PID_Ki", 0), config().get<float>("PID_Kd", 0), dt);
    pidControllerFront.set(config().get<float>("PID_front_Kp", 1), config().get<float>("PID_front_Ki", 1),
                           config().get<float>("PID_front_Kd", 0), dt);
    pidControllerRear.set(config().get<float>("PID_rear_Kp", 1), config().get<float>("PID_rear_Ki", 1),
                          config().get<float>("PID_rear_Kd", 0), dt);
}

// reference :34-289
bool TrajectoryPointController::cycle()
{
    auto phoenix = getService<phoenix_CC2016_service::Phoenix_CC2016Service>("PHOENIX_SERVICE");
    const DriveMode mode = phoenix->driveMode();
    if (mode == DriveMode::IDLE) {                                   // :36-49 park the car
        CarState idle = inheritedState(*car, "IDLE", 100);
        idle.state = CarStateType::IDLE;
        idle.steering_front = 0;
        idle.steering_rear = 0;
        idle.targetSpeed = 0;
        car->putState(idle);
        return true;
    }
    car->removeState("IDLE");                                         // :51

    CarState state = inheritedState(*car, "DEFAULT", 10);            // :54-61
    const std::string type = config().get<std::string>("type", "tobiMPC");
    bool ok;
    if (type == "tobiMPC") ok = stepTobiMpc(mode, state);
    else if (type == "mikMPC") {
        // :127-213 calls call_andromeda() from the `control-systems` submodule, whose source is not part of the
        // reference tree (.SUBMODULES.json: missing).  Not available in this build.
        logger.error("mikMPC") << "controller type mikMPC is not available in this build (andromeda solver source missing)";
        ok = false;
    } else ok = stepPid(state);
    if (!ok) return false;

    setIndicators(state);                                             // :227-240
    if (state.targetSpeed < 0.5) {                                    // :276-283 "we probably are in a crossing"
        state.indicatorLeft = false;
        state.indicatorRight = false;
        state.steering_front = 0;
        state.steering_rear = 0;
    }
    car->putState(state);                                             // :286
    logger.debug("car state, v, sf, sr") << state.targetSpeed << " " << state.steering_front << " " << state.steering_rear;
    return true;
}

// reference :64-126
bool TrajectoryPointController::stepTobiMpc(DriveMode mode, CarState& state)
{
    float lookahead = m_trajectoryPointDistanceLookup.linearSearch(car->velocity());                 // :66
    if (mode == DriveMode::FOH)                                                                       // :70-73
        lookahead = config().get<float>("regelpunktMin", 0.6) + car->velocity() * config().get<float>("regelpunktSlope", 0.1);

    street_environment::TrajectoryPoint point = getTrajectoryPoint(lookahead);                       // :76
    double v = car->velocity();
    if (std::fabs(v) < 0.1) {                                                                         // :78-82
        logger.debug("cycle") << "car is slow: " << car->velocity();
        v = 0.1;
    }
    // the float components are widened before atan2
    const double phi_soll = std::atan2((double)point.directory.y, (double)point.directory.x);        // :84
    const double y_soll = point.position.y;                                                           // :85

    mpcParameters.weight_y = config().get<double>("weight_y", 20);                                    // :92-96 live tuning
    mpcParameters.weight_phi = config().get<double>("weight_phi", 7);
    mpcParameters.weight_steeringFront = config().get<double>("weight_steering_front", 0.0005);
    mpcParameters.weight_steeringRear = config().get<double>("weight_steering_rear", 10);
    mpcParameters.stepSize = 0.1;

    double steering_front = 0, steering_rear = 0;
    mpcControllerTobi(v, y_soll, phi_soll, &steering_front, &steering_rear);                         // :97
    if (std::isnan(steering_front) || std::isnan(steering_rear) || std::isnan(point.velocity))       // :101-103
        logger.error("trajectory_point_controller: ") << "invalid vals: " << steering_front << " " << steering_rear;

    state.steering_front = steering_front;                                                            // :114-125
    state.steering_rear = steering_rear;
    state.targetSpeed = point.velocity;
    state.targetDistance = point.position.length();
    *debugging_trajectoryPoint = point;
    state.state = point.velocity == 0 ? CarStateType::IDLE : CarStateType::DRIVING;
    return true;
}

// reference :214-225
bool TrajectoryPointController::stepPid(CarState& state)
{
    const float lookahead = m_trajectoryPointDistanceLookup.linearSearch(car->velocity());
    street_environment::TrajectoryPoint point = getTrajectoryPoint(lookahead);
    state.steering_front = pidControllerFront.pid(point.position.y);
    state.steering_rear = pidControllerRear.pid(point.directory.angle());
    state.targetSpeed = point.velocity;
    state.targetDistance = point.position.length();
    return true;
}

// reference :227-240: indicate towards the lane the trajectory changes to
void TrajectoryPointController::setIndicators(CarState& state)
{
    state.indicatorLeft = false;
    state.indicatorRight = false;
    if (trajectory->size() == 0) return;
    const bool startsRight = trajectory->at(0).isRight();
    for (const street_environment::TrajectoryPoint& tp : *trajectory) {
        if (tp.isRight() == startsRight) continue;
        state.indicatorLeft = startsRight;
        state.indicatorRight = !startsRight;
        break;
    }
}

// reference :301-389 — model of the two-axle-steered car in path coordinates, solved over MPC_HORIZON steps
void TrajectoryPointController::mpcControllerTobi(double v, double delta_y, double delta_phi, double* steering_front,
                                                  double* steering_rear)
{
    const double T = mpcParameters.stepSize;
    v = m_mpcLookupVelocity.linearSearch(v);                          // :323

    dlib::matrix<double, STATES, STATES> A;
    A = 1, T * v,
        0, 1;                                                         // :326-327
    dlib::matrix<double, STATES, CONTROLS> B;
    B = 0, T * v,
        T * v / l, -T * v / l;                                        // :329-330
    dlib::matrix<double, STATES, 1> C;
    C = 0, 0;                                                         // :332-333 no constant disturbance
    dlib::matrix<double, STATES, 1> Q;
    Q = mpcParameters.weight_y, mpcParameters.weight_phi;             // :359-360
    dlib::matrix<double, CONTROLS, 1> R;
    R = mpcParameters.weight_steeringFront, mpcParameters.weight_steeringRear;   // :362-363

    dlib::mpc<STATES, CONTROLS, MPC_HORIZON> controller(A, B, C, Q, R, lower, upper);   // :366 (fresh every cycle)
    dlib::matrix<double, STATES, 1> target;
    target = delta_y, delta_phi;
    controller.set_target(target);                                    // :368-371
    dlib::matrix<double, STATES, 1> current_state;
    current_state = 0, 0;
    const dlib::matrix<double, CONTROLS, 1> action = controller(current_state);   // :377-380
    *steering_front = action(0, 0);
    *steering_rear = action(1, 0);
}

// reference :392-476
street_environment::TrajectoryPoint TrajectoryPointController::getTrajectoryPoint(const float distanceToPoint)
{
    street_environment::TrajectoryPoint point;                        // default (:394-403) — also what the engine
    point.position.x = distanceToPoint;                               // returns for an empty polyline
    point.position.y = 0;
    point.directory.x = 1;
    point.directory.y = 0;
    point.velocity = 0;
    const size_t n = trajectory->size();
    if (n == 0) {                                                     // :404-407
        logger.warn("cycle") << "Can't follow anything";
        return point;
    }
    // AoS channel payload -> SoA staging -> arc-length scan + interpolation on the device (:423-442)
    px_.resize(n); py_.resize(n); dx_.resize(n); dy_.resize(n); vel_.resize(n);
    for (size_t i = 0; i < n; ++i) {
        const street_environment::TrajectoryPoint& tp = trajectory->at(i);
        px_[i] = tp.position.x; py_[i] = tp.position.y;
        dx_[i] = tp.directory.x; dy_[i] = tp.directory.y;
        vel_[i] = tp.velocity;
    }
    int32_t idx = -1;
    uint8_t flag = BMPC_PT_EMPTY;
    const int rc = bmpc_extract_points(engine_, 1, (int)n, nullptr, px_.data(), py_.data(), dx_.data(), dy_.data(),
                                       vel_.data(), 0, 1, &distanceToPoint, &idx, &flag, &point.position.x,
                                       &point.position.y, &point.directory.x, &point.directory.y, &point.velocity);
    if (rc != BMPC_OK) {
        logger.error("getTrajectoryPoint") << "engine error: " << bmpc_last_error();
        return point;
    }
    if (idx >= 0) point.distanceToMiddleLane = trajectory->at((size_t)idx).distanceToMiddleLane;   // `= top` copies it
    if (flag == BMPC_PT_LAST)                                         // :439-442
        logger.warn("No trajectoryPoint found, returning the last point of the trajectory")
            << "trajPointCount" << n << " distanceToTrajectoryPoint: " << distanceToPoint;
    applyCrossingSlowDown(point);                                     // :445-473 (host: stateful PID)
    return point;
}

// reference :445-473: approach stop points (velocity == 0 on the trajectory) with a PID on the remaining distance
void TrajectoryPointController::applyCrossingSlowDown(street_environment::TrajectoryPoint& point)
{
    float minVelocity = config().get<float>("maxVelocityCrossing", 1.0);
    for (const street_environment::TrajectoryPoint& tp : *trajectory) {
        if (tp.velocity != 0) continue;
        const float distanceToStop = lms::math::sgn(tp.position.x) * tp.position.length() - config().get<float>("stoppingDistance", 0.35);
        if (!(distanceToStop < config().get<float>("distanceToStop", 1))) {
            slowDownCar.reset();
            continue;
        }
        logger.debug("distanceToStop") << distanceToStop;
        const float cap = config().get<float>("maxVelocityCrossing", 1.0);
        float velocity = slowDownCar.pid(distanceToStop);
        if (std::isnan(velocity) || velocity >= cap) velocity = cap;
        if (distanceToStop <= config().get<float>("crossingSaftyZone", 0.05) || velocity < 0) velocity = 0.0;
        if (velocity < minVelocity) minVelocity = velocity;
        point.velocity = minVelocity;
        if (minVelocity == 0) break;
    }
}
