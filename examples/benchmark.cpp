// mpc_benchmark on the GPU: n (default 1000) random Panda targets solved as ONE batch, then the per-trajectory checks of
// the reference's examples/benchmark.cpp (:82-160: finite-difference jerk > 10 x max_jerk, Cartesian linear / angular speed,
// tool below the table; :164-194: 162-column row appended per trajectory) for the warm-start profile (host) and the MPC
// trajectory (sample kernel).  Row format = what analysis/benchmark_analysis.ipynb parses:
//   min x28 rk | max x28 rk | min x28 mpc | max x28 mpc | final error 14 rk | 14 mpc | flags [jerk lin ang collision] rk, mpc | target q, v
// usage: mpc_benchmark [n] [out.txt] [seed]
#include <chrono>
#include <cstdlib>
#include <fstream>
#include <iomanip>
#include <iostream>

#include "motion_planner.hpp"

using mpcb::Traj;
using mpcb::Vec6;
using mpcb::Vec7;

int main(int argc, char** argv) {
    const int n = argc > 1 ? std::atoi(argv[1]) : 1000;
    const char* out = argc > 2 ? argv[2] : "benchmark_data.txt";
    const unsigned long long seed = argc > 3 ? std::strtoull(argv[3], nullptr, 10) : 1;
    const int nPoints = 200;

    MotionPlanner planner("robot_utils/panda-model/panda_arm.urdf");
    planner.set_constraint_margins(0.8, 0.8, 0.6, 0.9, 0.1);                     // benchmark.cpp:9
    const int N = planner.num_nodes();

    // targets: benchmark.cpp:19-42 (start = mid-range pose at rest, never reset in the reference's loop)
    std::vector<double> x0((size_t)14 * n), xT((size_t)14 * n), sx((size_t)14 * N * n), su((size_t)7 * N * n), sT(n);
    std::vector<mpcb_stats> stats(n);
    mpcb_generate_problems(&planner.limits(), 0, seed, n, x0.data(), xT.data());

    const auto t0 = std::chrono::steady_clock::now();
    planner.solve_batch(n, x0.data(), xT.data(), sx.data(), su.data(), sT.data(), stats.data());
    std::vector<mpcb_account> acc(n);
    if (mpcb_sample_batch(planner.handle(), n, sx.data(), su.data(), sT.data(), xT.data(), nPoints, nullptr, acc.data(), MPCB_MEM_HOST, nullptr) != MPCB_OK)
        throw std::runtime_error(mpcb_last_error(planner.handle()));
    // the warm-start ("Ruckig") half of the accounting, benchmark.cpp:72-114, on the device too (sample_profile_kernel)
    std::vector<mpcb_account> acc_rk(n);
    if (mpcb_sample_profile_batch(planner.handle(), n, x0.data(), xT.data(), nPoints, nullptr, acc_rk.data(), nullptr, MPCB_MEM_HOST, nullptr) != MPCB_OK)
        throw std::runtime_error(mpcb_last_error(planner.handle()));
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    std::ofstream f(out, std::ios_base::app);
    if (!f.is_open()) { std::cerr << "could not open " << out << "\n"; return 1; }
    f << std::setprecision(6);
    int fails[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    double meanT = 0;
    for (int i = 0; i < n; i++) {
        // row format of benchmark.cpp:169-193: min / max of the 28 signals (warm start, MPC), final-state errors, 8 flags, target
        for (int r = 0; r < 28; r++) f << acc_rk[i].mins[r] << " ";
        for (int r = 0; r < 28; r++) f << acc_rk[i].maxs[r] << " ";
        for (int r = 0; r < 28; r++) f << acc[i].mins[r] << " ";
        for (int r = 0; r < 28; r++) f << acc[i].maxs[r] << " ";
        for (int j = 0; j < 14; j++) f << acc_rk[i].final_err[j] << " ";
        for (int j = 0; j < 14; j++) f << acc[i].final_err[j] << " ";
        for (int k = 0; k < 4; k++) { f << acc_rk[i].flags[k] << " "; fails[k] += !acc_rk[i].flags[k]; }
        for (int k = 0; k < 4; k++) { f << acc[i].flags[k] << " "; fails[4 + k] += !acc[i].flags[k]; }
        for (int j = 0; j < 7; j++) f << xT[14 * i + j] << " ";
        for (int j = 0; j < 7; j++) f << xT[14 * i + 7 + j] << (j < 6 ? " " : "\n");
        meanT += sT[i];
    }
    std::cout << n << " trajectories solved + sampled in " << secs << " s (" << n / secs << " OCP/s end to end); mean final time " << meanT / n << " s\n"
              << "failures warm start [jerk lin ang collision] = " << fails[0] << " " << fails[1] << " " << fails[2] << " " << fails[3]
              << "   MPC = " << fails[4] << " " << fails[5] << " " << fails[6] << " " << fails[7] << "\n";
    return 0;
}
