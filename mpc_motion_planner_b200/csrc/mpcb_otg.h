// Jerk-limited, time-synchronised point-to-point profiles for the 7 joints: the warm start of the OCP and the baseline
// trajectory of the benchmark accounting.  This is the role Ruckig<7> plays in the reference
// (/root/reference/mpc_solver/motionPlanner.cpp:36-38,50-52,86-88,146-175; motionPlanner.hpp:73-96,130-142): state
// (q0, v0, a = 0) -> (qT, vT, a = 0) under |v| <= m_v v_max, |a| <= m_a a_max, |j| <= m_j j_max, every joint ending at the
// duration of the slowest one.  Ruckig's source is not in this environment; the generator below is written from the
// problem statement and pinned on the one Ruckig trajectory the reference embeds
// (tests/golden/embedded_trajectory.npz, tests/test_otg.py: duration to 1e-6, 201 x 7 samples to the 6 digits printed).
//
// Per joint the profile is  [pulse A: v0 -> vc] [cruise at vc] [pulse B: vc -> vT]  where a pulse is the fastest
// jerk-limited velocity change with zero acceleration at both ends (jerk +-j, optional hold at +-a_max, jerk -+j).
//   step 1 (per joint): minimum duration = best of { cruise at +-v_max,  no cruise with the peak velocity vc that makes the
//           two pulses cover the distance exactly }
//   step 2: T = max over joints; every other joint lowers |vc| until its duration is T (the form the embedded
//           trajectory shows for all six non-limiting joints); if a joint cannot meet T (a blocked interval), T grows.
// Host code only (plain C++), no GPU work.
#pragma once
#include <math.h>
#include "mpcb.h"

namespace mpcb {
namespace otg {

struct Pulse { double tj, th, sgn; };      // jerk time, hold time, sign of the acceleration

inline Pulse pulse(double va, double vb, double amax, double jmax) {
    Pulse p;
    const double dv = fabs(vb - va);
    p.sgn = vb >= va ? 1.0 : -1.0;
    if (dv * jmax >= amax * amax) { p.tj = amax / jmax; p.th = dv / amax - p.tj; }
    else { p.tj = sqrt(dv / jmax); p.th = 0.0; }
    return p;
}
inline double ptime(const Pulse& p) { return 2.0 * p.tj + p.th; }

// duration of the profile with cruise velocity vc, or -1 when the distance cannot be met with a cruise time >= 0
struct Eval { double T, tc, dist; bool ok; };
inline Eval eval_vc(double dp, double v0, double vT, double vc, double amax, double jmax) {
    const Pulse A = pulse(v0, vc, amax, jmax), B = pulse(vc, vT, amax, jmax);
    const double tA = ptime(A), tB = ptime(B);
    Eval e;
    e.dist = 0.5 * (v0 + vc) * tA + 0.5 * (vc + vT) * tB;      // a symmetric pulse moves at the mean velocity
    const double rest = dp - e.dist;
    e.tc = 0.0; e.ok = false;
    if (fabs(vc) > 1e-14) { e.tc = rest / vc; e.ok = e.tc >= -1e-12; if (e.tc < 0) e.tc = 0; }
    else e.ok = fabs(rest) < 1e-14;
    e.T = tA + tB + e.tc;
    return e;
}

constexpr int GRID = 512;

// step 1: minimum duration and its cruise velocity
inline bool min_time(double dp, double v0, double vT, double vmax, double amax, double jmax, double* Tmin, double* vcbest) {
    double best = INFINITY, bvc = 0;
    for (int s = -1; s <= 1; s += 2) {
        const Eval e = eval_vc(dp, v0, vT, s * vmax, amax, jmax);
        if (e.ok && e.T < best) { best = e.T; bvc = s * vmax; }
    }
    // no-cruise profiles: roots of h(vc) = dist(vc) - dp
    double vprev = -vmax, hprev = eval_vc(dp, v0, vT, vprev, amax, jmax).dist - dp;
    for (int i = 1; i <= GRID; i++) {
        const double vcur = -vmax + 2.0 * vmax * i / GRID, hcur = eval_vc(dp, v0, vT, vcur, amax, jmax).dist - dp;
        if (hprev == 0.0 || (hprev < 0) != (hcur < 0)) {
            double lo = vprev, hi = vcur, hlo = hprev;
            for (int it = 0; it < 80; it++) {
                const double mid = 0.5 * (lo + hi), hm = eval_vc(dp, v0, vT, mid, amax, jmax).dist - dp;
                if ((hm < 0) == (hlo < 0)) { lo = mid; hlo = hm; } else hi = mid;
            }
            const double vc = 0.5 * (lo + hi);
            const Pulse A = pulse(v0, vc, amax, jmax), B = pulse(vc, vT, amax, jmax);
            const double T = ptime(A) + ptime(B);
            if (T < best) { best = T; bvc = vc; }
        }
        vprev = vcur; hprev = hcur;
    }
    *Tmin = best; *vcbest = bvc;
    return best < INFINITY;
}

// step 2: cruise velocity whose profile lasts exactly T (searched from the time-optimal vc towards 0)
inline bool sync_vc(double dp, double v0, double vT, double vmax, double amax, double jmax, double T, double vcopt, double* vc) {
    const double dir = vcopt >= 0 ? 1.0 : -1.0;
    double vprev = vcopt;
    Eval eprev = eval_vc(dp, v0, vT, vprev, amax, jmax);
    if (eprev.ok && fabs(eprev.T - T) <= 1e-13 * fmax(1.0, T)) { *vc = vprev; return true; }
    // scan |vc| downwards through 0 to the other limit
    for (int i = 1; i <= 2 * GRID; i++) {
        const double vcur = vcopt - dir * (fabs(vcopt) + vmax) * i / (2 * GRID);
        if (fabs(vcur) > vmax) break;
        const Eval ecur = eval_vc(dp, v0, vT, vcur, amax, jmax);
        if (eprev.ok && ecur.ok && (eprev.T < T) != (ecur.T < T) && (vprev > 0) == (vcur > 0)) {
            double lo = vprev, hi = vcur;
            const bool lolt = eprev.T < T;
            for (int it = 0; it < 80; it++) {
                const double mid = 0.5 * (lo + hi);
                const Eval em = eval_vc(dp, v0, vT, mid, amax, jmax);
                if (em.ok && (em.T < T) == lolt) lo = mid; else hi = mid;
            }
            *vc = 0.5 * (lo + hi);
            return true;
        }
        vprev = vcur; eprev = ecur;
    }
    return false;
}

inline void fill_dof(mpcb_otg_traj* tr, int d, double q0, double v0, double vT, double vc, double tc, double amax, double jmax) {
    const Pulse A = pulse(v0, vc, amax, jmax), B = pulse(vc, vT, amax, jmax);
    double* t = tr->t[d];
    double* j = tr->j[d];
    t[0] = A.tj; j[0] = A.sgn * jmax; t[1] = A.th; j[1] = 0; t[2] = A.tj; j[2] = -A.sgn * jmax;
    t[3] = tc; j[3] = 0;
    t[4] = B.tj; j[4] = B.sgn * jmax; t[5] = B.th; j[5] = 0; t[6] = B.tj; j[6] = -B.sgn * jmax;
    tr->q0[d] = q0; tr->v0[d] = v0; tr->vc[d] = vc;
}

// 0 = ok, 1 = ok after growing T past a blocked interval, < 0 = no profile (input outside the limits)
inline int calculate(const mpcb_limits* lim, const double* x0, const double* xT, mpcb_otg_traj* tr) {
    double Tmin[7], vco[7], vl[7], al[7], jl[7];
    double T = 0;
    for (int d = 0; d < 7; d++) {
        vl[d] = lim->margins[1] * lim->vmax[d]; al[d] = lim->margins[2] * lim->amax[d]; jl[d] = lim->margins[4] * lim->jmax[d];
        if (!(vl[d] > 0 && al[d] > 0 && jl[d] > 0)) return -2;
        if (fabs(x0[7 + d]) > vl[d] * (1 + 1e-12) || fabs(xT[7 + d]) > vl[d] * (1 + 1e-12)) return -1;   // Ruckig rejects such input too
        if (!min_time(xT[d] - x0[d], x0[7 + d], xT[7 + d], vl[d], al[d], jl[d], &Tmin[d], &vco[d])) return -3;
        if (Tmin[d] > T) T = Tmin[d];
    }
    int status = 0;
    for (int attempt = 0; attempt < 400; attempt++) {
        bool all = true;
        for (int d = 0; d < 7 && all; d++) {
            double vc = vco[d];
            if (Tmin[d] < T * (1 - 1e-15)) all = sync_vc(xT[d] - x0[d], x0[7 + d], xT[7 + d], vl[d], al[d], jl[d], T, vco[d], &vc);
            if (all) {
                const Eval e = eval_vc(xT[d] - x0[d], x0[7 + d], xT[7 + d], vc, al[d], jl[d]);
                fill_dof(tr, d, x0[d], x0[7 + d], xT[7 + d], vc, e.tc, al[d], jl[d]);
            }
        }
        if (all) { tr->T = T; tr->status = status; return status; }
        T *= 1.005; status = 1;
    }
    return -4;
}

inline void at_time(const mpcb_otg_traj* tr, double t, double* q, double* v, double* a) {
    if (t < 0) t = 0;
    for (int d = 0; d < 7; d++) {
        double p = tr->q0[d], vel = tr->v0[d], acc = 0, left = t;
        for (int ph = 0; ph < 7; ph++) {
            const double dt = left < tr->t[d][ph] ? left : tr->t[d][ph], j = tr->j[d][ph];
            p += dt * (vel + dt * (0.5 * acc + dt * j / 6.0));
            vel += dt * (acc + 0.5 * dt * j);
            acc += dt * j;
            left -= dt;
            if (left <= 0) break;
        }
        if (left > 0) { p += vel * left; }       // past the end: constant velocity, as Ruckig's at_time does
        q[d] = p; v[d] = vel; a[d] = acc;
    }
}

}  // namespace otg
}  // namespace mpcb
