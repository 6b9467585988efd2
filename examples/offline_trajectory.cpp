// One random Panda start -> target minimum-time OCP solved on the GPU through the MotionPlanner mirror; writes the
// 403 x 29 text file analysis/data_analysis.ipynb of the reference reads (row 0: target, rows 1-201: warm-start profile,
// rows 202-402: MPC; columns [t, q(7), v(7), a(7), tau(7)]).  Same call sequence as the reference's
// examples/offline_trajectory.cpp (:9 margins, :18-22 sampling, :26-41 Cartesian scale-down, :46-48 bounds check,
// :58 solve, :69-105 file).  usage: offline_trajectory [out.txt] [seed]
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iomanip>
#include <iostream>

#include "motion_planner.hpp"

using mpcb::Traj;
using mpcb::Vec6;
using mpcb::Vec7;

static void write_rows(std::ofstream& f, const Traj<1>& t, const Traj<7>& q, const Traj<7>& v, const Traj<7>& a, const Traj<7>& tau) {
    for (int i = 0; i < t.cols; i++) {
        f << t(0, i);
        for (int j = 0; j < 7; j++) f << " " << q(j, i);
        for (int j = 0; j < 7; j++) f << " " << v(j, i);
        for (int j = 0; j < 7; j++) f << " " << a(j, i);
        for (int j = 0; j < 7; j++) f << " " << tau(j, i);
        f << "\n";
    }
}

int main(int argc, char** argv) {
    const char* out = argc > 1 ? argv[1] : "optimal_solution.txt";
    const unsigned long long seed = argc > 2 ? std::strtoull(argv[2], nullptr, 10) : 0;
    MotionPlanner planner("robot_utils/panda-model/panda_arm.urdf", 0, seed);
    planner.set_constraint_margins(0.9, 0.9, 0.5, 0.9, 0.1);

    Vec7 target_position, target_velocity, current_position, current_velocity;
    planner.sample_random_state(current_position, current_velocity);
    planner.set_current_state(current_position, current_velocity);
    planner.sample_random_state(target_position, target_velocity);

    Vec6 task_velocity = planner.robot.forward_velocities(target_position, target_velocity);
    if (task_velocity.head<3>().norm() > planner.robot.max_linear_velocity) {
        target_velocity *= 0.9 * planner.robot.max_linear_velocity / task_velocity.head<3>().norm();
        task_velocity = planner.robot.forward_velocities(target_position, target_velocity);
    }
    if (task_velocity.tail<3>().norm() > planner.robot.max_angular_velocity) {
        target_velocity *= 0.9 * planner.robot.max_angular_velocity / task_velocity.tail<3>().norm();
        task_velocity = planner.robot.forward_velocities(target_position, target_velocity);
    }
    planner.set_target_state(target_position, target_velocity);

    if (planner.check_state_in_bounds(current_position, current_velocity) != 0 || planner.check_state_in_bounds(target_position, target_velocity) != 0)
        throw std::runtime_error("Initial or final state outside bounds");

    planner.solve_trajectory(true);
    std::cout << "status " << planner.last_stats.status << "  sqp iterations " << planner.last_stats.sqp_iters << "  qp iterations "
              << planner.last_stats.qp_iters_total << "  final time " << planner.solution_p() << " s\n";

    const int nPoints = 200;
    std::ofstream f(out);
    if (!f.is_open()) { std::cerr << "could not open " << out << "\n"; return 1; }
    f << std::setprecision(6);                       // Eigen's default stream precision: what fixes the reference's file format
    f << 0.0;
    for (int j = 0; j < 7; j++) f << " " << target_position[j];
    for (int j = 0; j < 7; j++) f << " " << target_velocity[j];
    for (int j = 0; j < 14; j++) f << " " << 0;
    f << "\n";
    Traj<1> time;
    Traj<7> q, v, a, tau;
    planner.get_ruckig_trajectory<nPoints>(time, q, v, a, tau);
    write_rows(f, time, q, v, a, tau);
    planner.get_MPC_trajectory<nPoints>(time, q, v, a, tau);
    write_rows(f, time, q, v, a, tau);
    return 0;
}
