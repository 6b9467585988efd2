// motion_planner.hpp -- C++ mirror of the reference's `class MotionPlanner`
// (/root/reference/mpc_solver/motionPlanner.hpp:16-176, motionPlanner.cpp) on top of the C ABI of libmpcb (mpcb.h).
//
// Same public names, argument meaning and call order as the reference, so examples/offline_trajectory.cpp and
// examples/benchmark.cpp read like the reference's mains.  Differences, all deliberate:
//   * fixed-size vectors are mpcb::Vec<N> (std::array based) instead of Eigen::Matrix<double,N,1>; trajectories are
//     mpcb::Traj<ROWS> (column = sample) -- Eigen is not available in this environment;
//   * the solver runs on the GPU (no CPU path): the constructor throws std::runtime_error if no CUDA device is found;
//   * the Ruckig warm start is our own jerk-limited synchronised generator (csrc/mpcb_otg.h; ruckig itself is absent,
//     SURVEY.md 8f-N1; pinned on the Ruckig trajectory the reference embeds): warm_start_RK() /
//     get_ruckig_trajectory() / get_RK_point() keep their names and operate on that profile;
//   * bounds: initial pin and final box are always applied last; get_MPC_point clamps to t = T (Q9);
//     set_min_height does not print (Q10); random numbers come from a seeded counter RNG (Q13);
//   * solve_batch() is the batched entry the reference does not have.
#pragma once
#include <array>
#include <cmath>
#include <stdexcept>
#include <string>
#include <vector>

#include "mpcb.h"

// The class is called MotionPlanner, as in the reference.  shim/motionPlanner.hpp (the Eigen-typed adapter that lets the
// reference's own mains compile unchanged) includes this header under another class name and wraps it.
#ifndef MPCB_PLANNER_CLASS
#define MPCB_PLANNER_CLASS MotionPlanner
#endif

#define NDOF 7

namespace mpcb {

template <int N>
struct Vec : std::array<double, N> {
    static Vec Zero() { Vec v; v.fill(0.0); return v; }
    Vec operator*(double s) const { Vec r; for (int i = 0; i < N; i++) r[i] = (*this)[i] * s; return r; }
    Vec operator/(double s) const { Vec r; for (int i = 0; i < N; i++) r[i] = (*this)[i] / s; return r; }
    Vec operator-(const Vec& o) const { Vec r; for (int i = 0; i < N; i++) r[i] = (*this)[i] - o[i]; return r; }
    Vec& operator*=(double s) { for (auto& x : *this) x *= s; return *this; }
    Vec& operator/=(double s) { for (auto& x : *this) x /= s; return *this; }
    double norm() const { double s = 0; for (double x : *this) s += x * x; return std::sqrt(s); }
    double maxCoeff() const { double m = (*this)[0]; for (double x : *this) m = x > m ? x : m; return m; }
    double minCoeff() const { double m = (*this)[0]; for (double x : *this) m = x < m ? x : m; return m; }
    template <int M> Vec<M> head() const { Vec<M> r; for (int i = 0; i < M; i++) r[i] = (*this)[i]; return r; }
    template <int M> Vec<M> tail() const { Vec<M> r; for (int i = 0; i < M; i++) r[i] = (*this)[N - M + i]; return r; }
};
typedef Vec<7> Vec7;
typedef Vec<3> Vec3;
typedef Vec<6> Vec6;

// ROWS x cols matrix, column-major like Eigen's fixed-size types (column = one time sample)
template <int ROWS>
struct Traj {
    int cols = 0;
    std::vector<double> d;
    Traj() {}
    explicit Traj(int c) : cols(c), d((size_t)ROWS * c, 0.0) {}
    void resize(int c) { cols = c; d.assign((size_t)ROWS * c, 0.0); }
    double& operator()(int r, int c) { return d[(size_t)c * ROWS + r]; }
    double operator()(int r, int c) const { return d[(size_t)c * ROWS + r]; }
    Vec<ROWS> col(int c) const { Vec<ROWS> v; for (int r = 0; r < ROWS; r++) v[r] = (*this)(r, c); return v; }
    double rowMin(int r) const { double m = (*this)(r, 0); for (int c = 1; c < cols; c++) m = std::fmin(m, (*this)(r, c)); return m; }
    double rowMax(int r) const { double m = (*this)(r, 0); for (int c = 1; c < cols; c++) m = std::fmax(m, (*this)(r, c)); return m; }
};

// robot_utils/pandaWrapper.hpp: limits table + the kinematic helpers the example mains call through `planner.robot`
class PandaWrapper {
  public:
    Vec7 min_position, max_position, max_velocity, max_acceleration, max_jerk, max_torque;
    double max_torqueDot = 1000;
    double max_linear_velocity = 1.7, max_angular_velocity = 2.5;      // pandaWrapper.hpp:37-38
    double min_height = 0.05;                                          // pandaWrapper.hpp:40
    explicit PandaWrapper(const std::string& urdf_path) {
        // the model is baked from robot_utils/panda-model/panda_arm.urdf at build time (tools/bake_panda_model.py): any other
        // model file is refused instead of being ignored
        const std::string::size_type slash = urdf_path.find_last_of("/\\");
        const std::string base = slash == std::string::npos ? urdf_path : urdf_path.substr(slash + 1);
        if (base != "panda_arm.urdf") throw std::invalid_argument("PandaWrapper: only panda_arm.urdf is supported (the model is baked at build time), got '" + urdf_path + "'");
        mpcb_limits l;
        mpcb_default_limits(&l);
        for (int j = 0; j < 7; j++) {
            min_position[j] = l.qmin[j]; max_position[j] = l.qmax[j]; max_velocity[j] = l.vmax[j]; max_acceleration[j] = l.amax[j];
            max_jerk[j] = l.jmax[j]; max_torque[j] = l.taumax[j];
        }
    }
    Vec7 inverse_velocities(const Vec7& q, const Vec3& linear_velocity, const Vec3& angular_velocity) const {
        Vec7 qd;
        mpcb_inverse_velocities(q.data(), linear_velocity.data(), angular_velocity.data(), qd.data());
        return qd;
    }
    Vec6 forward_velocities(const Vec7& q, const Vec7& qdot) const {
        Vec6 t;
        mpcb_forward_velocities(q.data(), qdot.data(), t.data());
        return t;
    }
    // pandaWrapper.cpp:14-58 (orientation column-major 3 x 3 as Eigen::Matrix3d::data(); random start seeded by `seed`)
    Vec7 inverse_kinematic(const std::array<double, 9>& orientation, const Vec3& position, unsigned long long seed = 0) const {
        mpcb_limits l;
        mpcb_default_limits(&l);
        Vec7 q;
        mpcb_inverse_kinematic(&l, orientation.data(), position.data(), seed, q.data(), nullptr);
        return q;
    }
    Vec7 rnea(const Vec7& q, const Vec7& v, const Vec7& a) const { Vec7 tau; mpcb_rnea(q.data(), v.data(), a.data(), tau.data()); return tau; }
    double tool_height(const Vec7& q) const { double z; mpcb_frame_height(q.data(), 1, &z); return z; }      // oMf[panda_tool].z
    double joint7_height(const Vec7& q) const { double z; mpcb_frame_height(q.data(), 0, &z); return z; }    // oMi[7].z
};

}  // namespace mpcb

class MPCB_PLANNER_CLASS {
  public:
    typedef mpcb::Vec7 Vec7;
    typedef mpcb::Vec<14> State;

    mpcb::PandaWrapper robot;                        // motionPlanner.hpp:32
    State current_state, target_state;               // motionPlanner.hpp:40-41
    const double eps = 1e-2;                         // motionPlanner.hpp:44
    double margin_position_, margin_velocity_, margin_acceleration_, margin_torque_, margin_jerk_;
    mpcb_stats last_stats;                           // polympc's info() of the last solve

    explicit MPCB_PLANNER_CLASS(const std::string& urdf_path, int device = 0, unsigned long long seed = 0) : robot(urdf_path), seed_(seed) {
        mpcb_default_settings(&settings_);           // motionPlanner.cpp:15-20
        mpcb_default_limits(&limits_);
        const int rc = mpcb_create(&settings_, device, &h_);
        if (rc != MPCB_OK) throw std::runtime_error(std::string("MotionPlanner: mpcb_create failed: ") + mpcb_last_error(nullptr));
        N_ = mpcb_num_nodes(&settings_);
        nodes_.resize(N_);
        mpcb_time_nodes(&settings_, nodes_.data());
        x_guess_.assign(14 * N_, 0.0); u_guess_.assign(7 * N_, 0.0); sol_x_ = x_guess_; sol_u_ = u_guess_;
        lam_.assign(mpcb_num_dual(&settings_), 0.0);
        Vec7 mid;
        for (int j = 0; j < 7; j++) mid[j] = (robot.max_position[j] + robot.min_position[j]) / 2;    // motionPlanner.cpp:5
        set_target_state(mid, Vec7::Zero());
        set_current_state(mid, Vec7::Zero());
        set_constraint_margins(1.0, 1.0, 1.0, 1.0, 1.0);                                             // motionPlanner.cpp:24
    }
    ~MPCB_PLANNER_CLASS() { mpcb_destroy(h_); }
    MPCB_PLANNER_CLASS(const MPCB_PLANNER_CLASS&) = delete;
    MPCB_PLANNER_CLASS& operator=(const MPCB_PLANNER_CLASS&) = delete;

    // motionPlanner.cpp:27-39: final state box = target +- eps
    // The acceleration argument only feeds Ruckig in the reference (motionPlanner.cpp:36-38); the profile generator here starts and
    // ends at rest in acceleration, so a nonzero value is refused instead of being dropped silently.
    void set_target_state(const Vec7& q, const Vec7& v, const Vec7& a = Vec7::Zero()) {
        require_zero_acceleration(a, "set_target_state");
        for (int j = 0; j < 7; j++) { target_state[j] = q[j]; target_state[7 + j] = v[j]; }
    }
    // motionPlanner.cpp:41-54: initial state pinned
    void set_current_state(const Vec7& q, const Vec7& v, const Vec7& a = Vec7::Zero()) {
        require_zero_acceleration(a, "set_current_state");
        for (int j = 0; j < 7; j++) { current_state[j] = q[j]; current_state[7 + j] = v[j]; }
    }
    // motionPlanner.cpp:56-90
    void set_constraint_margins(double mp, double mv, double ma, double mt, double mj) {
        margin_position_ = mp; margin_velocity_ = mv; margin_acceleration_ = ma; margin_torque_ = mt; margin_jerk_ = mj;
        limits_.margins[0] = mp; limits_.margins[1] = mv; limits_.margins[2] = ma; limits_.margins[3] = mt; limits_.margins[4] = mj;
        limits_.min_height = robot.min_height;
        push_limits();
    }
    void set_min_height(double min_height) { limits_.min_height = min_height; push_limits(); }    // motionPlanner.cpp:92-100
    // motionPlanner.cpp:102-114 (rejection on the height of joint 7)
    void sample_random_state(Vec7& q, Vec7& v) {
        if (mpcb_sample_random_state(&limits_, seed_, stream_++, q.data(), v.data()) != MPCB_OK)
            throw std::runtime_error("MotionPlanner::sample_random_state: min_height is out of the robot's reach");
    }
    // motionPlanner.cpp:116-144: 0 ok, 1 position, 2 velocity, 3 both, +10 acceleration
    int check_state_in_bounds(const Vec7& q, const Vec7& v, const Vec7& a = Vec7::Zero()) const {
        bool pc = false, vc = false, ac = false;
        for (int j = 0; j < 7; j++) {
            const double safety = (1 - margin_position_) * (robot.max_position[j] - robot.min_position[j]) / 2;
            pc = pc || q[j] > robot.max_position[j] - safety || q[j] < robot.min_position[j] + safety;
            vc = vc || std::fabs(v[j]) > margin_velocity_ * robot.max_velocity[j];
            ac = ac || std::fabs(a[j]) > margin_acceleration_ * robot.max_acceleration[j];
        }
        int flag = 0;
        if (pc && !vc) flag = 1;
        if (!pc && vc) flag = 2;
        if (pc && vc) flag = 3;
        if (ac) flag += 10;
        return flag;
    }
    // motionPlanner.cpp:177-208
    void solve_trajectory(bool use_ruckig_as_warm_start) {
        if (use_ruckig_as_warm_start) warm_start_RK();
        double T = 0;
        const int rc = mpcb_solve_batch(h_, 1, current_state.data(), target_state.data(), x_guess_.data(), u_guess_.data(), &p_guess_,
                                        carry_duals_ ? lam_.data() : nullptr, sol_x_.data(), sol_u_.data(), &T, lam_.data(), &last_stats,
                                        MPCB_MEM_HOST, nullptr);
        if (rc != MPCB_OK) throw std::runtime_error(std::string("mpcb_solve_batch: ") + mpcb_last_error(h_));
        sol_T_ = T;
        // re-seed the next guess from the solution with the end points put back (motionPlanner.cpp:199-207)
        x_guess_ = sol_x_; u_guess_ = sol_u_; p_guess_ = T;
        for (int c = 0; c < 14; c++) { x_guess_[c] = current_state[c]; x_guess_[14 * (N_ - 1) + c] = target_state[c]; }
    }
    // polympc keeps the multipliers of the previous solve; off by default, on for receding-horizon use
    void carry_duals(bool on) { carry_duals_ = on; }
    double solution_p() const { return sol_T_; }
    const std::vector<double>& solution_x() const { return sol_x_; }
    const std::vector<double>& solution_u() const { return sol_u_; }
    const std::vector<double>& time_nodes() const { return nodes_; }

    // motionPlanner.hpp:99-116; N samples + 1
    template <int N>
    void get_MPC_trajectory(mpcb::Traj<1>& time, mpcb::Traj<7>& pos, mpcb::Traj<7>& vel, mpcb::Traj<7>& acc, mpcb::Traj<7>& torque) {
        std::vector<double> rows((size_t)29 * (N + 1));
        const int rc = mpcb_sample_batch(h_, 1, sol_x_.data(), sol_u_.data(), &sol_T_, nullptr, N, rows.data(), nullptr, MPCB_MEM_HOST, nullptr);
        if (rc != MPCB_OK) throw std::runtime_error(std::string("mpcb_sample_batch: ") + mpcb_last_error(h_));
        unpack_rows(rows, N + 1, time, pos, vel, acc, torque);
    }
    // motionPlanner.hpp:73-96 on the warm-start profile
    template <int N>
    void get_ruckig_trajectory(mpcb::Traj<1>& time, mpcb::Traj<7>& pos, mpcb::Traj<7>& vel, mpcb::Traj<7>& acc, mpcb::Traj<7>& torque) {
        time.resize(N + 1); pos.resize(N + 1); vel.resize(N + 1); acc.resize(N + 1); torque.resize(N + 1);
        for (int i = 0; i <= N; i++) {
            Vec7 q, v, a, tau;
            get_RK_point(rk_T_ / N * i, q, v, a, tau);
            time(0, i) = rk_T_ / N * i;
            for (int j = 0; j < 7; j++) { pos(j, i) = q[j]; vel(j, i) = v[j]; acc(j, i) = a[j]; torque(j, i) = tau[j]; }
        }
    }
    // motionPlanner.hpp:118-128 (t in seconds; clamped to [0, T])
    void get_MPC_point(double t, Vec7& q, Vec7& v, Vec7& a, Vec7& tau) const {
        const double s = sol_T_ > 0 ? std::fmin(std::fmax(t / sol_T_, 0.0), 1.0) : 0.0;
        interpolate(s, q, v, a);
        tau = robot.rnea(q, v, a);
    }
    // motionPlanner.hpp:130-142: the synchronised jerk-limited profile at time t (Trajectory::at_time); before the first
    // warm start, or for a problem the generator rejected, the cubic-Hermite fallback of mpcb_guess_batch
    void get_RK_point(double t, Vec7& q, Vec7& v, Vec7& a, Vec7& tau) const {
        if (rk_ok_) {
            mpcb_otg_at_time(&rk_traj_, std::fmin(std::fmax(t, 0.0), rk_T_), q.data(), v.data(), a.data());
            tau = robot.rnea(q, v, a);
            return;
        }
        const double T = rk_T_, s = T > 0 ? std::fmin(std::fmax(t / T, 0.0), 1.0) : 0.0;
        const double h00 = 2 * s * s * s - 3 * s * s + 1, h10 = s * s * s - 2 * s * s + s, h01 = -2 * s * s * s + 3 * s * s, h11 = s * s * s - s * s;
        const double d00 = 6 * s * s - 6 * s, d10 = 3 * s * s - 4 * s + 1, d01 = -6 * s * s + 6 * s, d11 = 3 * s * s - 2 * s;
        const double e00 = 12 * s - 6, e10 = 6 * s - 4, e01 = -12 * s + 6, e11 = 6 * s - 2;
        for (int j = 0; j < 7; j++) {
            const double q0 = rk_x0_[j], v0 = rk_x0_[7 + j], q1 = rk_xT_[j], v1 = rk_xT_[7 + j];
            q[j] = h00 * q0 + h10 * T * v0 + h01 * q1 + h11 * T * v1;
            v[j] = (d00 * q0 + d10 * T * v0 + d01 * q1 + d11 * T * v1) / T;
            a[j] = (e00 * q0 + e10 * T * v0 + e01 * q1 + e11 * T * v1) / (T * T);
        }
        tau = robot.rnea(q, v, a);
    }
    // motionPlanner.hpp:145-172: guess from a regularly time-spaced trajectory (7 x nPoint each, column = sample)
    void warm_start(double final_time, const mpcb::Traj<7>& pos, const mpcb::Traj<7>& vel, const mpcb::Traj<7>& acc) {
        const int nPoint = pos.cols;
        for (int k = 0; k < N_; k++) {
            const int idx = (int)std::lround(nodes_[k] * (nPoint - 1));
            for (int j = 0; j < 7; j++) { x_guess_[14 * k + j] = pos(j, idx); x_guess_[14 * k + 7 + j] = vel(j, idx); u_guess_[7 * k + j] = acc(j, idx); }
        }
        p_guess_ = final_time;
    }

    // Batched entry (not in the reference): n independent problems with this planner's limits and margins; the guess is the
    // same warm start solve_trajectory(true) uses.  x0, xT: n x 14; outputs sol_x n x 14N, sol_u n x 7N, sol_T n.
    void solve_batch(int n, const double* x0, const double* xT, double* sol_x, double* sol_u, double* sol_T, mpcb_stats* stats = nullptr) {
        std::vector<double> gx((size_t)n * 14 * N_), gu((size_t)n * 7 * N_), gT(n);
        mpcb_guess_batch(&settings_, &limits_, n, x0, xT, gx.data(), gu.data(), gT.data());
        const int rc = mpcb_solve_batch(h_, n, x0, xT, gx.data(), gu.data(), gT.data(), nullptr, sol_x, sol_u, sol_T, nullptr, stats, MPCB_MEM_HOST, nullptr);
        if (rc != MPCB_OK) throw std::runtime_error(std::string("mpcb_solve_batch: ") + mpcb_last_error(h_));
    }
    // compute the warm-start profile for the current / target state without solving (benchmark accounting of that profile)
    void warm_start_profile() { warm_start_RK(); }
    mpcb_handle* handle() { return h_; }
    const mpcb_limits& limits() const { return limits_; }
    const mpcb_settings& settings() const { return settings_; }
    int num_nodes() const { return N_; }

  private:
    // motionPlanner.cpp:146-175: otg.calculate + sampling at the node times (mpcb_guess_batch does both)
    void warm_start_RK() {
        mpcb_guess_batch(&settings_, &limits_, 1, current_state.data(), target_state.data(), x_guess_.data(), u_guess_.data(), &p_guess_);
        rk_T_ = p_guess_; rk_x0_ = current_state; rk_xT_ = target_state;
        rk_ok_ = mpcb_otg_calculate(&limits_, current_state.data(), target_state.data(), &rk_traj_) >= 0 && rk_traj_.T == p_guess_;
    }
    // polympc solution_x_at / solution_u_at: Lagrange interpolation on the segment containing s in [0,1]
    void interpolate(double s, Vec7& q, Vec7& v, Vec7& a) const {
        const int PO = settings_.poly_order, NS = settings_.num_seg;
        int sg = (int)std::floor(s * NS);
        if (sg > NS - 1) sg = NS - 1;
        const double t0 = nodes_[sg * PO], t1 = nodes_[sg * PO + PO];
        for (int j = 0; j < 7; j++) q[j] = v[j] = a[j] = 0;
        for (int jn = 0; jn <= PO; jn++) {
            double w = 1;
            for (int l = 0; l <= PO; l++)
                if (l != jn) w *= (s - nodes_[sg * PO + l]) / (nodes_[sg * PO + jn] - nodes_[sg * PO + l]);
            (void)t0; (void)t1;
            const int k = sg * PO + jn;
            for (int j = 0; j < 7; j++) { q[j] += w * sol_x_[14 * k + j]; v[j] += w * sol_x_[14 * k + 7 + j]; a[j] += w * sol_u_[7 * k + j]; }
        }
    }
    static void unpack_rows(const std::vector<double>& rows, int n, mpcb::Traj<1>& time, mpcb::Traj<7>& pos, mpcb::Traj<7>& vel,
                            mpcb::Traj<7>& acc, mpcb::Traj<7>& torque) {
        time.resize(n); pos.resize(n); vel.resize(n); acc.resize(n); torque.resize(n);
        for (int i = 0; i < n; i++) {
            const double* r = rows.data() + (size_t)29 * i;
            time(0, i) = r[0];
            for (int j = 0; j < 7; j++) { pos(j, i) = r[1 + j]; vel(j, i) = r[8 + j]; acc(j, i) = r[15 + j]; torque(j, i) = r[22 + j]; }
        }
    }

    void push_limits() {
        const int rc = mpcb_set_limits(h_, &limits_);
        if (rc != MPCB_OK) throw std::runtime_error(std::string("MotionPlanner: mpcb_set_limits failed: ") + mpcb_last_error(h_));
    }
    static void require_zero_acceleration(const Vec7& a, const char* who) {
        for (int j = 0; j < 7; j++)
            if (a[j] != 0.0) throw std::invalid_argument(std::string("MotionPlanner::") + who + ": nonzero boundary accelerations are not supported by the profile generator");
    }
    mpcb_handle* h_ = nullptr;
    mpcb_settings settings_;
    mpcb_limits limits_;
    int N_ = 0;
    std::vector<double> nodes_, x_guess_, u_guess_, sol_x_, sol_u_, lam_;
    double p_guess_ = 1.0, sol_T_ = 0.0, rk_T_ = 1.0;
    mpcb_otg_traj rk_traj_{};
    bool rk_ok_ = false;
    State rk_x0_{}, rk_xT_{};
    bool carry_duals_ = false;
    unsigned long long seed_ = 0, stream_ = 0;
};
