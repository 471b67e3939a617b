// Developer probe: host-side cost of every C-ABI call the facade makes in one module cycle (capacity-1 handle, the tobi
// shape (2,2,4)), one call at a time and as the whole sequence.  g++ -O2 -Iinclude scripts/probes/call_latency.cpp -Ltrajectory_controller_b200 -lbmpc
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <vector>

#include "bmpc.h"

static double now_us() { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
static double p50(std::vector<double>& v) { std::sort(v.begin(), v.end()); return v[v.size() / 2]; }

int main()
{
    bmpc_t* h = nullptr;
    if (bmpc_create(&h, 2, 2, 4, 1, BMPC_FP64, BMPC_MODE_EXACT, -1)) { std::printf("create: %s\n", bmpc_last_error()); return 1; }
    double A[4] = {1, 0.1, 0, 1}, B[4] = {0.005, 0.005, 0.1, -0.1}, C[2] = {0, 0}, Q[2] = {20, 7}, R[2] = {0.05, 10};
    double lo[2] = {-0.38, -0.38}, hi[2] = {0.38, 0.38}, tgt[2] = {0.02, 0.01}, x0[2] = {0.0, 0.0}, u[2];
    uint32_t it;
    const int reps = 2000;
    std::vector<double> t_reset, t_opt, t_model, t_target, t_solve, t_all, t_solve_only, t_sync;
    for (int k = 0; k < reps + 50; ++k) {
        A[1] = 0.1 + 1e-4 * (k % 7); tgt[0] = 0.02 + 1e-3 * (k % 5);
        const double a = now_us();
        bmpc_reset(h);
        const double b = now_us();
        bmpc_set_options(h, 0.01, 10000);
        const double c = now_us();
        if (bmpc_set_model(h, A, B, C, Q, R, lo, hi, 0, 1)) { std::printf("model: %s\n", bmpc_last_error()); return 1; }
        const double d = now_us();
        bmpc_set_target(h, tgt, -1, 0, 0);
        const double e = now_us();
        if (bmpc_solve(h, 1, x0, u, &it, nullptr, BMPC_WARM)) { std::printf("solve: %s\n", bmpc_last_error()); return 1; }
        const double f = now_us();
        if (k >= 50) { t_reset.push_back(b - a); t_opt.push_back(c - b); t_model.push_back(d - c); t_target.push_back(e - d); t_solve.push_back(f - e); t_all.push_back(f - a); }
    }
    for (int k = 0; k < reps; ++k) {       // the solve alone, queue empty before it
        bmpc_sync(h);
        const double a = now_us();
        bmpc_solve(h, 1, x0, u, &it, nullptr, 0);
        const double b = now_us();
        bmpc_sync(h);
        const double c = now_us();
        t_solve_only.push_back(b - a); t_sync.push_back(c - b);
    }
    float kms = 0; bmpc_last_kernel_ms(h, &kms);
    std::printf("{\"reset\": %.2f, \"set_options\": %.2f, \"set_model\": %.2f, \"set_target\": %.2f, \"solve_after_them\": %.2f, \"sequence\": %.2f, "
                "\"solve_alone_cold\": %.2f, \"empty_sync\": %.2f, \"iters\": %u, \"kernel_us\": %.1f, \"launches\": %lld}\n",
                p50(t_reset), p50(t_opt), p50(t_model), p50(t_target), p50(t_solve), p50(t_all), p50(t_solve_only), p50(t_sync), it, kms * 1e3,
                (long long)bmpc_launch_count(h));
    bmpc_destroy(h);
    return 0;
}
