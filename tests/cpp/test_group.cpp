// Device groups through the C ABI from host C++ (the north star's "host C++ calling the GPU through a thin extern C ABI",
// sharded over the GPUs of one box): bmpc_group_solve over >= 2 members must equal, bit for bit, one handle solving the
// same instances.  With one GPU visible the members share device 0 (same sharding / packing / gather code).
//   usage: test_group [ndev]      ndev = members (default: all visible devices, at least 2)
#include <cmath>
#include <cstdint>
#include <cstring>
#include <random>
#include <vector>

#include <cuda_runtime_api.h>

#include "bmpc.h"
#include "check.h"

namespace {

// the synthetic (4,1) lateral model (forward Euler, T = 0.1, l = 0.21, actuator w = 5 rad/s, zeta = 1)
struct Bike { double A[16], B[4], C[4], Q[4], R[1], lo[1], hi[1]; };
Bike bike(double v, double wy, double wpsi, double r)
{
    const double T = 0.1, l = 0.21, w = 5.0, z = 1.0;
    Bike m{};
    const double A[16] = {1, T * v, 0, 0, 0, 1, T * v / l, 0, 0, 0, 1, T, 0, 0, -T * w * w, 1 - 2 * z * w * T};
    std::memcpy(m.A, A, sizeof A);
    m.B[3] = T * w * w;
    m.Q[0] = wy; m.Q[1] = wpsi;
    m.R[0] = r;
    m.hi[0] = 22.0 * M_PI / 180.0; m.lo[0] = -m.hi[0];
    return m;
}

bool same_bits(const void* a, const void* b, size_t n) { return std::memcmp(a, b, n) == 0; }

}  // namespace

int main(int argc, char** argv)
{
    int visible = 0;
    if (cudaGetDeviceCount(&visible) != cudaSuccess || visible == 0) {
        std::fprintf(stderr, "test_group: no CUDA device (the engine has no CPU path)\n");
        return 2;
    }
    int ndev = argc > 1 ? std::atoi(argv[1]) : (visible >= 2 ? visible : 3);
    std::vector<int> devs(ndev);
    for (int k = 0; k < ndev; ++k) devs[k] = k % visible;
    std::printf("test_group: %d member(s) over %d visible device(s)\n", ndev, visible);

    const int S = 4, I = 1, N = 30;
    const int64_t B = 50021;
    std::mt19937_64 gen(42);
    std::uniform_real_distribution<double> U(-1.0, 1.0);
    std::vector<double> x0(B * S, 0.0);
    for (int64_t b = 0; b < B; ++b) { x0[b * S] = 0.3 * U(gen); x0[b * S + 1] = 0.4 * U(gen); }
    const Bike m = bike(1.0, 20.0, 7.0, 0.05);

    // ---- one handle on device 0: the reference for the group
    bmpc_t* one = nullptr;
    CHECK(bmpc_create(&one, S, I, N, B, BMPC_FP64, BMPC_MODE_EXACT, 0) == BMPC_OK);
    if (!one) { std::fprintf(stderr, "%s\n", bmpc_last_error()); return 2; }
    CHECK(bmpc_set_model(one, m.A, m.B, m.C, m.Q, m.R, m.lo, m.hi, 0, 1) == BMPC_OK);
    std::vector<double> u1(B * I), c1(B * N * I);
    std::vector<uint32_t> it1(B);
    std::vector<uint8_t> st1(B);
    CHECK(bmpc_solve(one, B, x0.data(), u1.data(), it1.data(), st1.data(), BMPC_KEEP_CONTROLS) == BMPC_OK);
    CHECK(bmpc_get_controls(one, 0, B, c1.data()) == BMPC_OK);

    for (int assignment : {BMPC_GROUP_INTERLEAVED, BMPC_GROUP_CONTIGUOUS}) {
        bmpc_group_t* g = nullptr;
        CHECK_MSG(bmpc_group_create(&g, S, I, N, B, BMPC_FP64, BMPC_MODE_EXACT, devs.data(), ndev, assignment) == BMPC_OK, "%s", bmpc_last_error());
        if (!g) continue;
        CHECK(bmpc_group_size(g) == ndev);
        int64_t total = 0;
        for (int k = 0; k < ndev; ++k) { int64_t f = 0, n = 0; CHECK(bmpc_group_shard(g, k, B, &f, &n) == BMPC_OK); total += n; }
        CHECK(total == B);
        CHECK(bmpc_group_set_model(g, m.A, m.B, m.C, m.Q, m.R, m.lo, m.hi, 0, 1) == BMPC_OK);
        std::vector<double> ug(B * I, -1.0), cg(B * N * I);
        std::vector<uint32_t> itg(B, 0xdeadbeefu);
        std::vector<uint8_t> stg(B, 0xff);
        CHECK_MSG(bmpc_group_solve(g, B, x0.data(), ug.data(), itg.data(), stg.data(), BMPC_KEEP_CONTROLS) == BMPC_OK, "%s", bmpc_last_error());
        CHECK(same_bits(ug.data(), u1.data(), B * I * 8));
        CHECK(same_bits(itg.data(), it1.data(), B * 4));
        CHECK(same_bits(stg.data(), st1.data(), B));
        CHECK(bmpc_group_get_controls(g, 0, B, cg.data()) == BMPC_OK);
        CHECK(same_bits(cg.data(), c1.data(), B * N * I * 8));
        float ms = 0;
        CHECK(bmpc_group_last_kernel_ms(g, &ms) == BMPC_OK && ms > 0);
        std::printf("  assignment %d: %lld instances, slowest member %.3f ms of kernels\n", assignment, (long long)B, ms);
        // per-member handles stay reachable (options, kernel info)
        CHECK(bmpc_group_member(g, 0) != nullptr && bmpc_group_member(g, ndev) == nullptr);
        // error path: a count beyond the capacity
        CHECK(bmpc_group_solve(g, B + 1, x0.data(), ug.data(), nullptr, nullptr, 0) == BMPC_E_INVALID);
        bmpc_group_destroy(g);
    }

    // ---- persistent controllers with per-instance models and targets, a few warm ticks
    {
        const int64_t Bw = 4099;
        std::vector<double> A(Bw * 16), Bm(Bw * 4), C(Bw * 4, 0.0), Q(Bw * 4), R(Bw), lo(Bw), hi(Bw), tg(Bw * S, 0.0);
        for (int64_t b = 0; b < Bw; ++b) {
            const Bike mb = bike(0.5 + 0.001 * (b % 1000), 5.0 * (1 + b % 4), 1.75 * (1 + (b / 4) % 4), 0.05);
            std::memcpy(&A[b * 16], mb.A, sizeof mb.A); std::memcpy(&Bm[b * 4], mb.B, sizeof mb.B);
            std::memcpy(&Q[b * 4], mb.Q, sizeof mb.Q);
            R[b] = mb.R[0]; lo[b] = mb.lo[0]; hi[b] = mb.hi[0];
            tg[b * S] = 0.02 * U(gen);
        }
        bmpc_t* h = nullptr; bmpc_group_t* g = nullptr;
        CHECK(bmpc_create(&h, S, I, N, Bw, BMPC_FP64, BMPC_MODE_EXACT, 0) == BMPC_OK);
        CHECK(bmpc_group_create(&g, S, I, N, Bw, BMPC_FP64, BMPC_MODE_EXACT, devs.data(), ndev, BMPC_GROUP_INTERLEAVED) == BMPC_OK);
        if (h && g) {
            CHECK(bmpc_set_model(h, A.data(), Bm.data(), C.data(), Q.data(), R.data(), lo.data(), hi.data(), 1, Bw) == BMPC_OK);
            CHECK(bmpc_group_set_model(g, A.data(), Bm.data(), C.data(), Q.data(), R.data(), lo.data(), hi.data(), 1, Bw) == BMPC_OK);
            CHECK(bmpc_set_target(h, tg.data(), -1, 1, Bw) == BMPC_OK);
            CHECK(bmpc_group_set_target(g, tg.data(), -1, 1, Bw) == BMPC_OK);
            std::vector<double> x(x0.begin(), x0.begin() + Bw * S), ua(Bw), ub(Bw);
            std::vector<uint32_t> ia(Bw), ib(Bw);
            for (int tick = 0; tick < 3; ++tick) {
                CHECK(bmpc_solve(h, Bw, x.data(), ua.data(), ia.data(), nullptr, BMPC_WARM) == BMPC_OK);
                CHECK(bmpc_group_solve(g, Bw, x.data(), ub.data(), ib.data(), nullptr, BMPC_WARM) == BMPC_OK);
                CHECK(same_bits(ua.data(), ub.data(), Bw * 8) && same_bits(ia.data(), ib.data(), Bw * 4));
                for (int64_t b = 0; b < Bw; ++b) { x[b * S] *= 0.9; x[b * S + 2] += 0.1 * ua[b]; }
            }
        }
        bmpc_group_destroy(g);
        bmpc_destroy(h);
    }
    bmpc_destroy(one);
    return finish("test_group");
}
