// The re-hosted LMS module driven headless through the shim blackboard: TRAJECTORY in, CAR / TRAJECTORY_POINT out.
// Expected values come from the CPU oracle (oracle/liboracle.so — test infrastructure): the extraction restatement
// and the dlib::mpc restatement fed with exactly the glue arithmetic of src/trajectory_point_follower.cpp:64-126.
// Needs a GPU.
#include <cmath>
#include <memory>

#include "trajectory_point_follower.h"
#include "check.h"

extern "C" {
#include "extract_oracle.h"
#include "mpc_oracle.h"
void* getInstance();
}

using street_environment::CarCommand;
using street_environment::Trajectory;
using street_environment::TrajectoryPoint;
typedef phoenix_CC2016_service::Phoenix_CC2016Service Phoenix;
typedef phoenix_CC2016_service::CCDriveMode Mode;

namespace {

Trajectory sine_path(int n, float ds, float amp, float vel)
{
    Trajectory t;
    for (int i = 0; i < n; ++i) {
        const float x = ds * i, y = amp * std::sin(1.3f * x) + 0.05f;
        const float tx = 1.0f, ty = amp * 1.3f * std::cos(1.3f * x);
        const float len = std::sqrt(tx * tx + ty * ty);
        t.push_back(TrajectoryPoint(lms::math::vertex2f(x, y), lms::math::vertex2f(tx / len, ty / len), vel, -0.2f));
    }
    return t;
}

// What the reference glue computes on the CPU for one cycle (tobiMPC branch), via the oracle restatements.
void expected_cycle(const Trajectory& t, float lookahead, double v, double* sf, double* sr, orc_point* pt)
{
    std::vector<float> px, py, dx, dy, vel;
    for (const TrajectoryPoint& p : t) { px.push_back(p.position.x); py.push_back(p.position.y); dx.push_back(p.directory.x); dy.push_back(p.directory.y); vel.push_back(p.velocity); }
    orc_extract_point(px.data(), py.data(), dx.data(), dy.data(), vel.data(), 1, (int)t.size(), lookahead, pt);
    if (std::fabs(v) < 0.1) v = 0.1;
    const double phi = std::atan2((double)pt->dir_y, (double)pt->dir_x), y = pt->pos_y;
    const double T = 0.1, l = 0.21, amax = 22 * M_PI / 180;
    const double A[4] = {1, T * v, 0, 1}, B[4] = {0, T * v, T * v / l, -T * v / l}, C[2] = {0, 0}, Q[2] = {20, 7};
    const double R[2] = {0.0005, 10}, lo[2] = {-amax, -amax}, hi[2] = {amax, amax}, tgt[2] = {y, phi}, x0[2] = {0, 0};
    static orc_mpc m;
    orc_mpc_init(&m, 2, 2, 4, A, B, C, Q, R, lo, hi);
    orc_mpc_set_target(&m, tgt, -1);
    double u[2];
    orc_mpc_step(&m, x0, u);
    *sf = u[0]; *sr = u[1];
}

}  // namespace

int main()
{
    lms::Blackboard& bb = lms::Blackboard::instance();
    auto phoenix = std::make_shared<Phoenix>();
    bb.provideService<Phoenix>("PHOENIX_SERVICE", phoenix);
    auto traj = bb.channel<Trajectory>("TRAJECTORY");
    auto car = bb.channel<CarCommand>("CAR");
    auto dbg = bb.channel<TrajectoryPoint>("TRAJECTORY_POINT");

    std::unique_ptr<lms::Module> mod(static_cast<lms::Module*>(getInstance()));   // the LMS loader's path
    mod->logger.verbosity = 0;
    mod->config().setArray<float>("trajectoryPointDistanceLookupX", {0.0f, 4.0f});
    mod->config().setArray<float>("trajectoryPointDistanceLookupY", {0.4f, 1.2f});
    if (!mod->initialize()) { std::fprintf(stderr, "initialize() failed (no GPU?)\n"); return 2; }

    // ---- tobiMPC over a range of speeds: steering and trajectory point equal the CPU path bit for bit ----------
    *traj = sine_path(48, 0.07f, 0.25f, 1.5f);
    for (int k = 0; k < 12; ++k) {
        const float v = 0.02f + 0.3f * k;
        car->setMeasured(v, 0, 0);
        phoenix->setDriveMode(k % 3 == 0 ? Mode::FOH : Mode::FMH);
        CHECK(mod->cycle());
        const float lookahead = (k % 3 == 0) ? 0.6f + v * 0.1f : 0.4f + (1.2f - 0.4f) * (v - 0.0f) / (4.0f - 0.0f);
        double sf, sr; orc_point pt;
        expected_cycle(*traj, lookahead, v, &sf, &sr, &pt);
        const CarCommand::State* st = car->getState("DEFAULT");
        CHECK(st != nullptr);
        if (!st) continue;
        CHECK_MSG(st->steering_front == (float)sf && st->steering_rear == (float)sr, "k=%d steering (%.9g,%.9g) want (%.9g,%.9g)",
                  k, st->steering_front, st->steering_rear, sf, sr);
        CHECK(dbg->position.x == pt.pos_x && dbg->position.y == pt.pos_y && dbg->directory.x == pt.dir_x && dbg->velocity == pt.vel);
        CHECK(st->targetSpeed == 1.5f && st->state == CarCommand::StateType::DRIVING && st->priority == 10);
        CHECK(st->targetDistance == std::sqrt(pt.pos_x * pt.pos_x + pt.pos_y * pt.pos_y));
        CHECK(dbg->distanceToMiddleLane == -0.2f);
    }

    // ---- lookahead beyond the polyline: last point; lane change ahead: indicators ---------------------------------
    (*traj)[40].distanceToMiddleLane = 0.2f;
    mod->config().setArray<float>("trajectoryPointDistanceLookupY", {50.0f, 50.0f});
    mod->configsChanged();
    car->setMeasured(1.0f, 0, 0);
    phoenix->setDriveMode(Mode::FMH);
    CHECK(mod->cycle());
    CHECK(dbg->position.x == (*traj)[47].position.x && dbg->position.y == (*traj)[47].position.y);
    CHECK(car->getState("DEFAULT")->indicatorLeft == true && car->getState("DEFAULT")->indicatorRight == false);

    // ---- empty trajectory: default point, velocity 0 -> IDLE state type, steering zeroed (:276-283, :394-407) ------
    traj->clear();
    CHECK(mod->cycle());
    CHECK(dbg->position.x == 50.0f && dbg->position.y == 0.0f && dbg->directory.x == 1.0f && dbg->velocity == 0.0f);
    CHECK(car->getState("DEFAULT")->state == CarCommand::StateType::IDLE);
    CHECK(car->getState("DEFAULT")->steering_front == 0 && car->getState("DEFAULT")->steering_rear == 0);

    // ---- drive mode IDLE parks the car with priority 100; leaving IDLE removes that state (:36-52) -------------------
    phoenix->setDriveMode(Mode::IDLE);
    CHECK(mod->cycle());
    CHECK(car->getState("IDLE") != nullptr && car->getState("IDLE")->priority == 100 && car->getState("IDLE")->targetSpeed == 0);
    phoenix->setDriveMode(Mode::FMH);
    *traj = sine_path(20, 0.1f, 0.1f, 1.0f);
    CHECK(mod->cycle());
    CHECK(car->getState("IDLE") == nullptr);

    // ---- stop point close ahead: crossing slow-down clamps the target speed (:445-473) --------------------------------
    (*traj)[6].velocity = 0.0f;                                     // 0.6 m ahead -> distanceToStop = 0.25 < 1
    mod->config().setArray<float>("trajectoryPointDistanceLookupY", {0.3f, 0.3f});
    mod->configsChanged();
    CHECK(mod->cycle());
    CHECK(car->getState("DEFAULT")->targetSpeed <= 1.0f && car->getState("DEFAULT")->targetSpeed >= 0.0f);

    // ---- controller types -----------------------------------------------------------------------------------------------
    mod->config().set<std::string>("type", "mikMPC");
    CHECK(mod->cycle() == false);                                   // solver source unavailable: loud failure
    mod->config().set<std::string>("type", "pid");
    CHECK(mod->cycle());
    mod->config().set<std::string>("type", "tobiMPC");
    CHECK(mod->cycle());
    CHECK(mod->deinitialize());
    return finish("test_module");
}
