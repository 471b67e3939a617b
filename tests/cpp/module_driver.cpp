// Drives ANY build of the trajectory_point_controller MODULE library through the LMS loader's path
// (dlopen + getInstance, reference src/interface.cpp:3) over a fixed scenario and prints what the module wrote to
// the CAR and TRAJECTORY_POINT channels, as hex floats.  Used to compare three builds cycle by cycle:
//   oracle/_ref/libref_module_dlib.so      the reference's OWN module source + the real dlib::mpc   (CPU; golden)
//   tests/cpp/build/libref_module_bmpc.so  the reference's OWN module source + the GPU facade       (drop-in proof)
//   cpp/build/libtrajectory_point_controller.so   the re-hosted module (extraction on the GPU too)
#include <dlfcn.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>

#include "lms/module.h"
#include "phoenix_CC2016_service/phoenix_CC2016_service.h"
#include "street_environment/car.h"
#include "street_environment/trajectory.h"

using street_environment::CarCommand;
using street_environment::Trajectory;
using street_environment::TrajectoryPoint;
typedef phoenix_CC2016_service::Phoenix_CC2016Service Phoenix;
typedef phoenix_CC2016_service::CCDriveMode Mode;

static Trajectory path(int n, float ds, float amp, float freq, float off, float vel, int lane_change_at)
{
    Trajectory t;
    for (int i = 0; i < n; ++i) {
        const float x = ds * i, y = amp * std::sin(freq * x) + off;
        const float tx = 1.0f, ty = amp * freq * std::cos(freq * x);
        const float len = std::sqrt(tx * tx + ty * ty);
        t.push_back(TrajectoryPoint(lms::math::vertex2f(x, y), lms::math::vertex2f(tx / len, ty / len), vel,
                                    (lane_change_at >= 0 && i >= lane_change_at) ? 0.2f : -0.2f));
    }
    return t;
}

static void dump(int k, bool ok, const CarCommand& car, const TrajectoryPoint& p)
{
    std::printf("cycle %d ok=%d", k, (int)ok);
    for (const char* name : {"DEFAULT", "IDLE"}) {
        const CarCommand::State* s = const_cast<CarCommand&>(car).getState(name);
        if (!s) { std::printf(" %s:-", name); continue; }
        std::printf(" %s: sf=%a sr=%a v=%a d=%a st=%d prio=%d il=%d ir=%d", name, s->steering_front, s->steering_rear,
                    s->targetSpeed, s->targetDistance, (int)s->state, s->priority, (int)s->indicatorLeft, (int)s->indicatorRight);
    }
    std::printf(" pt: %a %a %a %a %a\n", p.position.x, p.position.y, p.directory.x, p.directory.y, p.velocity);
}

int main(int argc, char** argv)
{
    if (argc < 2) { std::fprintf(stderr, "usage: module_driver <module.so> [--time CYCLES]\n"); return 64; }
    // --time N: instead of dumping the 48-cycle scenario, time N tobiMPC cycles (the same scenario, repeated) and print
    // the per-cycle latency of cycle() as one JSON line
    const int timed = (argc >= 4 && std::strcmp(argv[2], "--time") == 0) ? std::atoi(argv[3]) : 0;
    void* so = dlopen(argv[1], RTLD_NOW | RTLD_GLOBAL);
    if (!so) { std::fprintf(stderr, "dlopen: %s\n", dlerror()); return 65; }
    typedef void* (*factory_t)();
    factory_t factory = (factory_t)dlsym(so, "getInstance");
    if (!factory) { std::fprintf(stderr, "no getInstance in %s\n", argv[1]); return 66; }

    lms::Blackboard& bb = lms::Blackboard::instance();
    auto phoenix = std::make_shared<Phoenix>();
    bb.provideService<Phoenix>("PHOENIX_SERVICE", phoenix);
    auto traj = bb.channel<Trajectory>("TRAJECTORY");
    auto car = bb.channel<CarCommand>("CAR");
    auto dbg = bb.channel<TrajectoryPoint>("TRAJECTORY_POINT");

    std::unique_ptr<lms::Module> mod(static_cast<lms::Module*>(factory()));
    mod->logger.verbosity = 0;
    mod->config().setArray<float>("trajectoryPointDistanceLookupX", {0.0f, 1.0f, 4.0f});
    mod->config().setArray<float>("trajectoryPointDistanceLookupY", {0.4f, 0.7f, 1.6f});
    mod->config().setArray<float>("mpcLookupVelocityX", {0.0f, 2.0f, 5.0f});
    mod->config().setArray<float>("mpcLookupVelocityY", {0.3f, 2.0f, 3.5f});
    if (!mod->initialize()) { std::fprintf(stderr, "initialize() failed\n"); return 2; }

    std::vector<double> lat;
    const int cycles = timed > 0 ? timed + 50 : 48;
    for (int kk = 0; kk < cycles; ++kk) {
        const int k = timed > 0 ? (kk % 40 == 20 ? 21 : kk % 40) : kk;            // timing: tobiMPC cycles only (no IDLE / pid)
        const float v = 0.03f + 0.11f * k;
        car->setMeasured(v, 0.01f * k, -0.005f * k);
        *traj = path(24 + (k % 5) * 9, 0.05f + 0.01f * (k % 4), 0.05f + 0.02f * (k % 7), 0.7f + 0.1f * (k % 3),
                     0.04f * ((k % 9) - 4), 0.6f + 0.2f * (k % 6), (k % 4 == 1) ? 10 + k % 8 : -1);
        if (k % 11 == 7) (*traj)[5 + k % 6].velocity = 0.0f;                        // stop point ahead (crossing logic)
        phoenix->setDriveMode(k == 20 ? Mode::IDLE : (k % 3 == 0 ? Mode::FOH : Mode::FMH));
        mod->config().set<double>("weight_y", 20.0 + (k % 5));                      // live tuning (reference :92-95)
        mod->config().set<double>("weight_steering_rear", 10.0 - 0.5 * (k % 4));
        mod->config().set<std::string>("type", (k >= 40 && k < 44) ? "pid" : "tobiMPC");
        const auto t0 = std::chrono::steady_clock::now();
        const bool ok = mod->cycle();
        const auto t1 = std::chrono::steady_clock::now();
        if (timed > 0) { if (kk >= 50) lat.push_back(std::chrono::duration<double, std::micro>(t1 - t0).count()); }
        else dump(k, ok, *car, *dbg);
    }
    if (timed > 0) {
        std::sort(lat.begin(), lat.end());
        double sum = 0; for (double x : lat) sum += x;
        std::printf("{\"module\": \"%s\", \"cycles\": %d, \"cycle_us\": {\"p50\": %.2f, \"p99\": %.2f, \"mean\": %.2f, \"min\": %.2f}}\n",
                    argv[1], (int)lat.size(), lat[lat.size() / 2], lat[(size_t)(lat.size() * 0.99)], sum / lat.size(), lat.front());
    }
    mod->deinitialize();
    return 0;
}
