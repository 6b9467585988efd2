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
//           trajectory shows for all six non-limiting joints); if a joint cannot meet T (a blocked interval), T moves to
//           the end of that interval and all joints are synchronised again.
// The per-joint routines compile for host and device: mpcb_guess_batch (host) and otg_guess_kernel (csrc/mpcb_kernels.cu, one
// joint per lane) share them.
#pragma once
#include <math.h>
#include "mpcb.h"
#include "mpcb_common.h"

namespace mpcb {
namespace otg {

struct Pulse { double tj, th, sgn; };      // jerk time, hold time, sign of the acceleration

MPCB_HD Pulse pulse(double va, double vb, double amax, double jmax) {
    Pulse p;
    const double dv = fabs(vb - va);
    p.sgn = vb >= va ? 1.0 : -1.0;
    if (dv * jmax >= amax * amax) { p.tj = amax / jmax; p.th = dv / amax - p.tj; }
    else { p.tj = sqrt(dv / jmax); p.th = 0.0; }
    return p;
}
MPCB_HD double ptime(const Pulse& p) { return 2.0 * p.tj + p.th; }

// duration of the profile with cruise velocity vc, or -1 when the distance cannot be met with a cruise time >= 0
struct Eval { double T, tc, dist; bool ok; };
MPCB_HD Eval eval_vc(double dp, double v0, double vT, double vc, double amax, double jmax) {
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

#define MPCB_OTG_GRID 512

// a joint that does not move at all (same position, at rest at both ends): its profile is "stay", for any duration.  Ruckig
// accepts such input (moving a subset of the joints is the common case); the cruise-velocity search below cannot represent it
// (vc = 0 with an arbitrary cruise time), so it is handled apart: Tmin = 0, never synchronised, all phases empty.
MPCB_HD bool is_hold(double dp, double v0, double vT) { return fabs(dp) <= 1e-12 && fabs(v0) <= 1e-12 && fabs(vT) <= 1e-12; }

// step 1: minimum duration and its cruise velocity
MPCB_HD bool min_time(double dp, double v0, double vT, double vmax, double amax, double jmax, double* Tmin, double* vcbest) {
    if (is_hold(dp, v0, vT)) { *Tmin = 0.0; *vcbest = 0.0; return true; }
    double best = INFINITY, bvc = 0;
    for (int s = -1; s <= 1; s += 2) {
        const Eval e = eval_vc(dp, v0, vT, s * vmax, amax, jmax);
        if (e.ok && e.T < best) { best = e.T; bvc = s * vmax; }
    }
    // no-cruise profiles: roots of h(vc) = dist(vc) - dp
    double vprev = -vmax, hprev = eval_vc(dp, v0, vT, vprev, amax, jmax).dist - dp;
    for (int i = 1; i <= MPCB_OTG_GRID; i++) {
        const double vcur = -vmax + 2.0 * vmax * i / MPCB_OTG_GRID, hcur = eval_vc(dp, v0, vT, vcur, amax, jmax).dist - dp;
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

// step 2: cruise velocity whose profile lasts exactly T.  The candidates are scanned from the time-optimal vc towards 0 and on
// to the other velocity limit over a fixed grid (MPCB_OTG_SYNC cells over [-vmax, vmax]; vc = 0 is a grid point).  Between two
// neighbouring candidates of the same sign the duration is continuous, so the pair reaches every duration between theirs; next
// to vc = 0 the cruise time grows without bound when the remaining distance has the sign of vc.
//   returns true  and the bracket [*lo, *hi] (duration(*lo) < T iff *lolt) when some pair reaches T;
//   returns false and *Tnext = the smallest reachable duration above T (INFINITY if none: T lies past every profile of this
//   joint), which is where a blocked interval of this joint ends.
#define MPCB_OTG_SYNC 1024
MPCB_HD double sync_grid(double vmax, int k) { return k == MPCB_OTG_SYNC / 2 ? 0.0 : -vmax + 2.0 * vmax * k / MPCB_OTG_SYNC; }
MPCB_HD bool sync_scan(double dp, double v0, double vT, double vmax, double amax, double jmax, double T, double vcopt, double* lo, double* hi,
                       double* Tnext) {
    const int dir = vcopt >= 0 ? 1 : -1;
    // first grid point strictly past vcopt in the scan direction
    int k = (int)floor((vcopt + vmax) / (2.0 * vmax) * MPCB_OTG_SYNC);
    k = k < 0 ? 0 : (k > MPCB_OTG_SYNC ? MPCB_OTG_SYNC : k);
    if (dir > 0) { while (k >= 0 && sync_grid(vmax, k) >= vcopt) k--; }
    else { k++; while (k <= MPCB_OTG_SYNC && sync_grid(vmax, k) <= vcopt) k++; }
    const double rest0 = dp - eval_vc(dp, v0, vT, 0.0, amax, jmax).dist;     // distance left to a cruise at vc -> 0
    double vprev = vcopt, tnext = INFINITY;
    Eval eprev = eval_vc(dp, v0, vT, vprev, amax, jmax);
    for (; k >= 0 && k <= MPCB_OTG_SYNC; k -= dir) {
        const double vcur = sync_grid(vmax, k);
        const Eval ecur = eval_vc(dp, v0, vT, vcur, amax, jmax);
        // the pair reaches the durations (plo, phi] (empty when phi < plo); blo is the end with the shorter duration
        double plo = 0, phi = -1, blo = vprev, bhi = vcur;
        if (eprev.ok && ecur.ok && ((vprev > 0) == (vcur > 0) || vprev == 0.0 || vcur == 0.0)) {
            plo = fmin(eprev.T, ecur.T); phi = fmax(eprev.T, ecur.T);
            if (ecur.T < eprev.T) { blo = vcur; bhi = vprev; }
        } else if (vcur == 0.0 && eprev.ok && rest0 * vprev > 0) { plo = eprev.T; phi = INFINITY; }
        else if (vprev == 0.0 && ecur.ok && rest0 * vcur > 0) { plo = ecur.T; phi = INFINITY; blo = vcur; bhi = vprev; }
        if (phi >= plo) {
            if (T > plo && T <= phi) { *lo = blo; *hi = bhi; *Tnext = T; return true; }
            if (plo >= T) tnext = fmin(tnext, plo * (1.0 + 1e-12) + 1e-300);
        }
        vprev = vcur; eprev = ecur;
    }
    *Tnext = tnext;
    return false;
}
// bisection inside the bracket of sync_scan
MPCB_HD double sync_bisect(double dp, double v0, double vT, double amax, double jmax, double T, double lo, double hi) {
    for (int it = 0; it < 80; it++) {
        const double mid = 0.5 * (lo + hi);
        const Eval em = eval_vc(dp, v0, vT, mid, amax, jmax);
        if (em.ok && em.T < T) lo = mid; else hi = mid;
    }
    return 0.5 * (lo + hi);
}
#define MPCB_OTG_ROUNDS 16       // blocked intervals skipped before the input is given up

// the 7 constant-jerk phases of one joint: durations t[7] and jerks j[7]
MPCB_HD void dof_phases(double v0, double vT, double vc, double tc, double amax, double jmax, double* t, double* j) {
    const Pulse A = pulse(v0, vc, amax, jmax), B = pulse(vc, vT, amax, jmax);
    t[0] = A.tj; j[0] = A.sgn * jmax; t[1] = A.th; j[1] = 0; t[2] = A.tj; j[2] = -A.sgn * jmax;
    t[3] = tc; j[3] = 0;
    t[4] = B.tj; j[4] = B.sgn * jmax; t[5] = B.th; j[5] = 0; t[6] = B.tj; j[6] = -B.sgn * jmax;
}
// state of one joint at time `time` (clamped below at 0; constant velocity past the end, as Ruckig's at_time)
MPCB_HD void dof_at_time(const double* t, const double* j, double q0, double v0, double time, double* q, double* v, double* a) {
    double p = q0, vel = v0, acc = 0, left = time < 0 ? 0 : time;
    for (int ph = 0; ph < 7; ph++) {
        const double dt = left < t[ph] ? left : t[ph], jj = j[ph];
        p += dt * (vel + dt * (0.5 * acc + dt * jj / 6.0));
        vel += dt * (acc + 0.5 * dt * jj);
        acc += dt * jj;
        left -= dt;
        if (left <= 0) break;
    }
    if (left > 0) p += vel * left;
    *q = p; *v = vel; *a = acc;
}
inline void fill_dof(mpcb_otg_traj* tr, int d, double q0, double v0, double vT, double vc, double tc, double amax, double jmax) {
    dof_phases(v0, vT, vc, tc, amax, jmax, tr->t[d], tr->j[d]);
    tr->q0[d] = q0; tr->v0[d] = v0; tr->vc[d] = vc;
}

// ---- fallback for inputs the generator rejects: time-scaled cubic Hermite profile q(s), s = t / T ---------------------------
// one joint respects |v| <= vl over [0, T] and |a| <= al (the acceleration is linear in s: extrema at the ends; the velocity
// is quadratic: extrema at the ends or at the vertex)
MPCB_HD bool hermite_ok(double dq, double v0, double v1, double T, double vl, double al) {
    const double a0 = (6 * dq / T - 4 * v0 - 2 * v1) / T, a1 = (-6 * dq / T + 2 * v0 + 4 * v1) / T;
    if (fabs(a0) > al || fabs(a1) > al) return false;
    if (a1 != a0) {
        const double sv = -a0 / (a1 - a0);
        if (sv > 0 && sv < 1) {
            const double vv = v0 + a0 * T * sv + 0.5 * (a1 - a0) * T * sv * sv;
            if (fabs(vv) > vl) return false;
        }
    }
    return true;
}
MPCB_HD void hermite_eval(double q0, double v0, double q1, double v1, double T, double sg, double* q, double* v, double* a) {
    const double h00 = 2 * sg * sg * sg - 3 * sg * sg + 1, h10 = sg * sg * sg - 2 * sg * sg + sg, h01 = -2 * sg * sg * sg + 3 * sg * sg,
                 h11 = sg * sg * sg - sg * sg;
    const double d00 = 6 * sg * sg - 6 * sg, d10 = 3 * sg * sg - 4 * sg + 1, d01 = -6 * sg * sg + 6 * sg, d11 = 3 * sg * sg - 2 * sg;
    const double e00 = 12 * sg - 6, e10 = 6 * sg - 4, e01 = -12 * sg + 6, e11 = 6 * sg - 2;
    *q = h00 * q0 + h10 * T * v0 + h01 * q1 + h11 * T * v1;
    *v = (d00 * q0 + d10 * T * v0 + d01 * q1 + d11 * T * v1) / T;
    *a = (e00 * q0 + e10 * T * v0 + e01 * q1 + e11 * T * v1) / (T * T);
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
    double lo[7], hi[7];
    bool need[7], all = false;
    for (int round = 0; round < MPCB_OTG_ROUNDS && !all; round++) {
        all = true;
        double Tn = T;
        for (int d = 0; d < 7; d++) {
            need[d] = Tmin[d] < T * (1 - 1e-15) && !is_hold(xT[d] - x0[d], x0[7 + d], xT[7 + d]);
            if (!need[d]) continue;
            double tn;
            if (!sync_scan(xT[d] - x0[d], x0[7 + d], xT[7 + d], vl[d], al[d], jl[d], T, vco[d], &lo[d], &hi[d], &tn)) { all = false; Tn = fmax(Tn, tn); }
        }
        if (all) break;
        if (!(Tn < INFINITY)) return -4;          // some joint has no profile that lasts T or longer
        T = Tn; status = 1;                        // the end of the blocked interval
    }
    if (!all) return -4;
    for (int d = 0; d < 7; d++) {
        const double dp = xT[d] - x0[d];
        const double vc = need[d] ? sync_bisect(dp, x0[7 + d], xT[7 + d], al[d], jl[d], T, lo[d], hi[d]) : vco[d];
        const Eval e = eval_vc(dp, x0[7 + d], xT[7 + d], vc, al[d], jl[d]);
        fill_dof(tr, d, x0[d], x0[7 + d], xT[7 + d], vc, is_hold(dp, x0[7 + d], xT[7 + d]) ? T : e.tc, al[d], jl[d]);
    }
    tr->T = T; tr->status = status;
    return status;
}

inline void at_time(const mpcb_otg_traj* tr, double t, double* q, double* v, double* a) {
    for (int d = 0; d < 7; d++) dof_at_time(tr->t[d], tr->j[d], tr->q0[d], tr->v0[d], t, q + d, v + d, a + d);
}

}  // namespace otg
}  // namespace mpcb
