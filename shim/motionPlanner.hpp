// shim/motionPlanner.hpp -- lets the reference's own example mains compile UNCHANGED against the B200 planner:
//     g++ -Ishim -Iinclude /root/reference/examples/offline_trajectory.cpp -lmpcb        (make examples-unchanged)
// The mains include "motionPlanner.hpp" and use Eigen types plus three raw pinocchio calls
// (/root/reference/examples/benchmark.cpp:108-110,154-156).  This header provides exactly that surface:
//   * class MotionPlanner with the reference's public members and signatures (/root/reference/mpc_solver/motionPlanner.hpp:16-176),
//     Eigen-typed, forwarding to the mirror of include/motion_planner.hpp (which talks to libmpcb through the C ABI);
//   * planner.robot with forward_velocities / inverse_velocities / the limit vectors / model / data / frame_id
//     (/root/reference/robot_utils/pandaWrapper.hpp);
//   * namespace pinocchio with forwardKinematics / updateFramePlacement and data.oMf[frame].translation().
// Eigen and pinocchio themselves are not available in this environment; shim/Eigen/Dense holds the minimal matrix type.
#pragma once
#include <cstdlib>
#include <ctime>
#include <fstream>
#include <iostream>
#include <limits>

#include "Eigen/Dense"

#define MPCB_PLANNER_CLASS MpcbPlannerCore
#include "motion_planner.hpp"

using namespace Eigen;

namespace pinocchio {
struct Model {};
struct SE3 {
    double t[3] = {0, 0, 0};
    const double* translation() const { return t; }
};
struct Data {
    std::vector<SE3> oMf;
    double q[7] = {0, 0, 0, 0, 0, 0, 0};
    Data() : oMf(2) {}
};
// the frame placement is computed when it is asked for (updateFramePlacement); forwardKinematics only records q
inline void forwardKinematics(const Model&, Data& data, const Eigen::Mat& q) { for (int j = 0; j < 7; j++) data.q[j] = q.d[j]; }
inline void updateFramePlacement(const Model&, Data& data, int frame_id) {
    double R[9];
    mpcb_tool_pose(data.q, R, data.oMf[frame_id].t);
}
}  // namespace pinocchio

namespace shim_detail {
template <int N>
inline mpcb::Vec<N> to_vec(const Eigen::Mat& m) {
    if (m.size() != N) throw std::runtime_error("shim: vector size mismatch");
    mpcb::Vec<N> v;
    for (int i = 0; i < N; i++) v[i] = m.d[i];
    return v;
}
template <int N>
inline Eigen::Matrix<double, N, 1> to_eigen(const mpcb::Vec<N>& v) {
    Eigen::Matrix<double, N, 1> m;
    for (int i = 0; i < N; i++) m.d[i] = v[i];
    return m;
}
}  // namespace shim_detail

// robot_utils/pandaWrapper.hpp with Eigen-typed members
class PandaWrapper {
  public:
    Matrix<double, NDOF, 1> min_position, max_position, max_velocity, max_acceleration, max_jerk, max_torque;
    double max_torqueDot = 1000;
    double max_linear_velocity = 1.7, max_angular_velocity = 2.5;
    double min_height = 0.05;
    pinocchio::Model model;
    pinocchio::Data data;
    int frame_id = 1;
    explicit PandaWrapper(const std::string& urdf_path) : core_(urdf_path) {
        min_position = shim_detail::to_eigen<7>(core_.min_position); max_position = shim_detail::to_eigen<7>(core_.max_position);
        max_velocity = shim_detail::to_eigen<7>(core_.max_velocity); max_acceleration = shim_detail::to_eigen<7>(core_.max_acceleration);
        max_jerk = shim_detail::to_eigen<7>(core_.max_jerk); max_torque = shim_detail::to_eigen<7>(core_.max_torque);
    }
    Matrix<double, 7, 1> inverse_velocities(const Mat& q, const Mat& linear_velocity, const Mat& angular_velocity) {
        return shim_detail::to_eigen<7>(core_.inverse_velocities(shim_detail::to_vec<7>(q), shim_detail::to_vec<3>(linear_velocity), shim_detail::to_vec<3>(angular_velocity)));
    }
    Matrix<double, 6, 1> forward_velocities(const Mat& q, const Mat& qdot) {
        return shim_detail::to_eigen<6>(core_.forward_velocities(shim_detail::to_vec<7>(q), shim_detail::to_vec<7>(qdot)));
    }
  private:
    mpcb::PandaWrapper core_;
};

class MotionPlanner {
  public:
    explicit MotionPlanner(std::string urdf_path) : robot(urdf_path), core_(urdf_path) { sync_public(); }

    PandaWrapper robot;
    Matrix<double, 2 * NDOF, 1> current_state, target_state;
    const double eps = 1e-2;
    const double inf = std::numeric_limits<double>::infinity();
    double margin_position_ = 1, margin_velocity_ = 1, margin_acceleration_ = 1, margin_torque_ = 1, margin_jerk_ = 1;

    void set_target_state(Matrix<double, NDOF, 1> q, Matrix<double, NDOF, 1> v, Matrix<double, NDOF, 1> a = Matrix<double, NDOF, 1>::Zero()) {
        core_.set_target_state(shim_detail::to_vec<7>(q), shim_detail::to_vec<7>(v), shim_detail::to_vec<7>(a));
        sync_public();
    }
    void set_current_state(Matrix<double, NDOF, 1> q, Matrix<double, NDOF, 1> v, Matrix<double, NDOF, 1> a = Matrix<double, NDOF, 1>::Zero()) {
        core_.set_current_state(shim_detail::to_vec<7>(q), shim_detail::to_vec<7>(v), shim_detail::to_vec<7>(a));
        sync_public();
    }
    void set_constraint_margins(double mp, double mv, double ma, double mt, double mj) {
        core_.set_constraint_margins(mp, mv, ma, mt, mj);
        margin_position_ = mp; margin_velocity_ = mv; margin_acceleration_ = ma; margin_torque_ = mt; margin_jerk_ = mj;
    }
    void set_min_height(double h) { core_.set_min_height(h); }
    void sample_random_state(Matrix<double, NDOF, 1>& q, Matrix<double, NDOF, 1>& v) {
        mpcb::Vec7 a, b;
        core_.sample_random_state(a, b);
        q = shim_detail::to_eigen<7>(a); v = shim_detail::to_eigen<7>(b);
    }
    void solve_trajectory(bool use_ruckig_as_warm_start) { core_.solve_trajectory(use_ruckig_as_warm_start); }
    int check_state_in_bounds(Matrix<double, NDOF, 1>& q, Matrix<double, NDOF, 1>& v, Matrix<double, NDOF, 1> a = Matrix<double, NDOF, 1>::Zero()) {
        return core_.check_state_in_bounds(shim_detail::to_vec<7>(q), shim_detail::to_vec<7>(v), shim_detail::to_vec<7>(a));
    }
    template <int N>
    void get_ruckig_trajectory(Matrix<double, 1, N + 1>& time, Matrix<double, 7, N + 1>& pos, Matrix<double, 7, N + 1>& vel, Matrix<double, 7, N + 1>& acc,
                               Matrix<double, 7, N + 1>& torque) {
        mpcb::Traj<1> t; mpcb::Traj<7> p, v, a, tq;
        core_.template get_ruckig_trajectory<N>(t, p, v, a, tq);
        unpack<N>(t, p, v, a, tq, time, pos, vel, acc, torque);
    }
    template <int N>
    void get_MPC_trajectory(Matrix<double, 1, N + 1>& time, Matrix<double, 7, N + 1>& pos, Matrix<double, 7, N + 1>& vel, Matrix<double, 7, N + 1>& acc,
                            Matrix<double, 7, N + 1>& torque) {
        mpcb::Traj<1> t; mpcb::Traj<7> p, v, a, tq;
        core_.template get_MPC_trajectory<N>(t, p, v, a, tq);
        unpack<N>(t, p, v, a, tq, time, pos, vel, acc, torque);
    }
    void get_MPC_point(double t, Matrix<double, 7, 1>& q, Matrix<double, 7, 1>& v, Matrix<double, 7, 1>& a, Matrix<double, 7, 1>& tau) {
        mpcb::Vec7 x, y, z, w;
        core_.get_MPC_point(t, x, y, z, w);
        q = shim_detail::to_eigen<7>(x); v = shim_detail::to_eigen<7>(y); a = shim_detail::to_eigen<7>(z); tau = shim_detail::to_eigen<7>(w);
    }
    void get_RK_point(double t, Matrix<double, 7, 1>& q, Matrix<double, 7, 1>& v, Matrix<double, 7, 1>& a, Matrix<double, 7, 1>& tau) {
        mpcb::Vec7 x, y, z, w;
        core_.get_RK_point(t, x, y, z, w);
        q = shim_detail::to_eigen<7>(x); v = shim_detail::to_eigen<7>(y); a = shim_detail::to_eigen<7>(z); tau = shim_detail::to_eigen<7>(w);
    }
    // motionPlanner.hpp:145-172: guess from regularly time-spaced samples (7 x nPoint each)
    void warm_start(double final_time, MatrixXd pos, MatrixXd vel, MatrixXd acc) {
        mpcb::Traj<7> p(pos.cols()), v(pos.cols()), a(pos.cols());
        for (int i = 0; i < pos.cols(); i++)
            for (int j = 0; j < 7; j++) { p(j, i) = pos(j, i); v(j, i) = vel(j, i); a(j, i) = acc(j, i); }
        core_.warm_start(final_time, p, v, a);
    }
    MpcbPlannerCore& core() { return core_; }

  private:
    MpcbPlannerCore core_;
    void sync_public() {
        for (int i = 0; i < 14; i++) { current_state.d[i] = core_.current_state[i]; target_state.d[i] = core_.target_state[i]; }
    }
    template <int N>
    static void unpack(const mpcb::Traj<1>& t, const mpcb::Traj<7>& p, const mpcb::Traj<7>& v, const mpcb::Traj<7>& a, const mpcb::Traj<7>& tq,
                       Matrix<double, 1, N + 1>& time, Matrix<double, 7, N + 1>& pos, Matrix<double, 7, N + 1>& vel, Matrix<double, 7, N + 1>& acc,
                       Matrix<double, 7, N + 1>& torque) {
        for (int i = 0; i <= N; i++) {
            time(0, i) = t(0, i);
            for (int j = 0; j < 7; j++) { pos(j, i) = p(j, i); vel(j, i) = v(j, i); acc(j, i) = a(j, i); torque(j, i) = tq(j, i); }
        }
    }
};
