// Problem generation shared by the host entry point (mpcb_generate_problems, csrc/mpcb_planner_host.cpp) and the device
// kernel (gen_problems_kernel, csrc/mpcb_kernels.cu; one problem per thread):
//   MotionPlanner::sample_random_state (/root/reference/mpc_solver/motionPlanner.cpp:102-114), the target generation of
//   examples/benchmark.cpp:19-42 (kind 0) and examples/offline_trajectory.cpp:18-41 (kind 1), PandaWrapper::inverse_velocities /
//   forward_velocities (robot_utils/pandaWrapper.cpp:60-107).
// Problem i draws from its own counter-based stream (seed, i), so any slice of a batch can be generated anywhere.
#pragma once
#include <math.h>
#include <stdint.h>

#include "mpcb.h"
#include "mpcb_common.h"
#include "mpcb_rbd.h"

namespace mpcb {
namespace gen {

// counter-based RNG: splitmix64 of (seed, stream, counter) -> U(-1,1)   (replaces Eigen::Random / std::rand, Q13)
struct Rng {
    uint64_t state;
    MPCB_HD Rng(uint64_t seed, uint64_t stream) : state(seed * 0x9E3779B97F4A7C15ull + stream * 0xD1B54A32D192ED03ull + 0x632BE59BD9B4E019ull) {}
    MPCB_HD uint64_t next() {
        uint64_t z = (state += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    MPCB_HD double sym() { return 2.0 * ((next() >> 11) * (1.0 / 9007199254740992.0)) - 1.0; }
};

#define MPCB_MAX_LINEAR_VELOCITY 1.7      // pandaWrapper.hpp:37-38
#define MPCB_MAX_ANGULAR_VELOCITY 2.5

// world-aligned tool Jacobian (6x7)
MPCB_HD void tool_jacobian(const double* q, double* J) {
    double oR[7][9], op[7][3];
    rbd::fk_world(q, oR, op);
    rbd::frame_jacobian(oR, op, PANDA(PANDA_TOOL_OFFSET), J, nullptr);
}
MPCB_HD void forward_velocities(const double* q, const double* qd, double* tv) {    // pandaWrapper.cpp:90-107
    double J[42];
    tool_jacobian(q, J);
    for (int r = 0; r < 6; r++) { double s = 0; for (int j = 0; j < 7; j++) s += J[r * 7 + j] * qd[j]; tv[r] = s; }
}
// J^T (J J^T + 1e-5 I)^-1 v   with an LDL^T solve of the 6x6 system                  // pandaWrapper.cpp:60-88
MPCB_HD void inverse_velocities(const double* q, const double* task, double* qd) {
    double J[42], A[36], L[36], D[6], y[6];
    tool_jacobian(q, J);
    for (int r = 0; r < 6; r++)
        for (int c = 0; c < 6; c++) { double s = 0; for (int j = 0; j < 7; j++) s += J[r * 7 + j] * J[c * 7 + j]; A[r * 6 + c] = s + (r == c ? 1e-5 : 0.0); }
    for (int k = 0; k < 36; k++) L[k] = 0;
    for (int j = 0; j < 6; j++) {
        double d = A[j * 6 + j];
        for (int k = 0; k < j; k++) d -= L[j * 6 + k] * L[j * 6 + k] * D[k];
        D[j] = d; L[j * 6 + j] = 1;
        for (int i = j + 1; i < 6; i++) {
            double s = A[i * 6 + j];
            for (int k = 0; k < j; k++) s -= L[i * 6 + k] * L[j * 6 + k] * D[k];
            L[i * 6 + j] = s / d;
        }
    }
    for (int i = 0; i < 6; i++) { double s = task[i]; for (int k = 0; k < i; k++) s -= L[i * 6 + k] * y[k]; y[i] = s; }
    for (int i = 0; i < 6; i++) y[i] /= D[i];
    for (int i = 5; i >= 0; i--) { double s = y[i]; for (int k = i + 1; k < 6; k++) s -= L[k * 6 + i] * y[k]; y[i] = s; }
    for (int j = 0; j < 7; j++) { double s = 0; for (int r = 0; r < 6; r++) s += J[r * 7 + j] * y[r]; qd[j] = s; }
}
#define MPCB_GEN_MAX_HEIGHT 1.0     /* joint 7 of the Panda reaches about 1.19 m; above 1.0 m the rejection loop hardly ever accepts */
MPCB_HD double norm3(const double* v) { return sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); }

// sample_random_state (motionPlanner.cpp:102-114): rejection on the height of JOINT 7 (not the tool, Q12)
MPCB_HD void sample_random_state(const mpcb_limits* lim, Rng& rng, double* q, double* v) {
    const double mp = lim->margins[0], mv = lim->margins[1];
    const double zero[3] = {0, 0, 0};
    double pos[3];
    int tries = 0;        // capped: an unreachable min_height must not hang a CPU thread or a GPU (the entry points refuse min_height > MPCB_GEN_MAX_HEIGHT)
    do {
        for (int j = 0; j < 7; j++) {
            const double range = lim->qmax[j] - lim->qmin[j];
            const double safety = (1 - mp) * range / 2;
            q[j] = 0.5 * (rng.sym() * (range - 2 * safety) + (lim->qmax[j] + lim->qmin[j]));
        }
        rbd::fk_point7(q, zero, pos);
    } while (pos[2] < lim->min_height && ++tries < (1 << 16));
    for (int j = 0; j < 7; j++) v[j] = mv * rng.sym() * lim->vmax[j];
}
// Cartesian scale-downs shared by both examples (benchmark.cpp:25-36, offline_trajectory.cpp:26-39)
MPCB_HD void scale_task_velocity(const double* q, double* v) {
    double tv[6];
    forward_velocities(q, v, tv);
    if (norm3(tv) > MPCB_MAX_LINEAR_VELOCITY) {
        const double f = 0.9 * MPCB_MAX_LINEAR_VELOCITY / norm3(tv);
        for (int j = 0; j < 7; j++) v[j] *= f;
        forward_velocities(q, v, tv);
    }
    if (norm3(tv + 3) > MPCB_MAX_ANGULAR_VELOCITY) {
        const double f = 0.9 * MPCB_MAX_ANGULAR_VELOCITY / norm3(tv + 3);
        for (int j = 0; j < 7; j++) v[j] *= f;
    }
}

// problem i of a batch: start x0 = a[14], target xT = b[14]
MPCB_HD void generate_one(const mpcb_limits* lim, int kind, uint64_t seed, uint64_t i, double* a, double* b) {
    Rng rng(seed, i);
    if (kind == 0) {
        // start: mid-range pose at rest (MotionPlanner constructor, motionPlanner.cpp:5-8; never reset in benchmark.cpp)
        for (int j = 0; j < 7; j++) { a[j] = (lim->qmax[j] + lim->qmin[j]) / 2; a[7 + j] = 0; }
        sample_random_state(lim, rng, b, b + 7);
        double task[6] = {0, 0, 0, 0, 0, 0};
        for (int k = 0; k < 3; k++) task[k] = rng.sym() * MPCB_MAX_LINEAR_VELOCITY;
        inverse_velocities(b, task, b + 7);                                     // benchmark.cpp:20
        scale_task_velocity(b, b + 7);                                          // benchmark.cpp:22-36
        double worst = 0;                                                       // benchmark.cpp:38-42
        bool any = false;
        for (int j = 0; j < 7; j++) {
            const double r = fabs(b[7 + j]) / (lim->margins[1] * lim->vmax[j]);
            if (r > 1) any = true;
            if (r > worst) worst = r;
        }
        if (any) for (int j = 0; j < 7; j++) b[7 + j] /= 1.1 * worst;
    } else {
        sample_random_state(lim, rng, a, a + 7);                                // offline_trajectory.cpp:18-19
        sample_random_state(lim, rng, b, b + 7);                                // :22
        scale_task_velocity(b, b + 7);                                          // :26-39
    }
}

}  // namespace gen
}  // namespace mpcb
