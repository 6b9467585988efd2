// CUDA kernels (sm_100a) of the batched OCP solver and their launchers.
//   sqp_stage_kernel: persistent CTAs, one OCP per CTA at a time: load | linearise + factorise + export | line search + step,
//                    out of shared memory -- phases in mpcb_qp.h / mpcb_sqp.h / mpcb_export.h
//   node_jac/val_kernel: the per-node rigid-body work (RNEA + derivatives), one thread per node of the whole batch
//   admm_kernel    : the hot loop, <= 700 box-ADMM iterations per QP with the interior inverses held in registers
//                    (mpcb_admm.h); QP data handed over through a per-OCP block in global memory
//   sample_kernel  : read-out + benchmark accounting (get_MPC_trajectory<N>, /root/reference/mpc_solver/motionPlanner.hpp:99-116;
//                    examples/benchmark.cpp:118-194)
//   fp64_peak_kernel: FP64 FMA throughput micro-benchmark (roofline denominator)
#include <cuda_runtime.h>
#include <stdint.h>
#include <mutex>

#include "mpcb_kernels.h"
#include "mpcb_gen.h"
#include "mpcb_otg.h"
#include "mpcb_sqp.h"

namespace mpcb {

// Per-OCP workspace in global memory: iterate z (kernel order) | multipliers | flags | node results | QP block (struct AL).
constexpr int MAX_TRIALS = 16;      // line_search_max_iter + 1 <= 16 (settings_supported)
template <class SH>
struct WS {
    static constexpr int Z = 0, LAM = SH::VP, FLAGS = LAM + SH::NDUAL + 8, JG = FLAGS + 4, GJ = JG + SH::NN * 176, GVAL = GJ + SH::NINEQ;
    static constexpr int QB = GVAL + MAX_TRIALS * SH::NINEQ;
    static constexpr size_t STRIDE = (size_t)QB + QBL<SH>::QPBLOCK;
};

template <class SH>
__device__ __forceinline__ void bind_problem(Ctx<SH>& c, Problem& pb, const KernelConst& kc, const BatchPtrs& bp, double* ws, int i) {
    c.z = ws + WS<SH>::Z;
    c.lam = ws + WS<SH>::LAM;
    c.gJg = ws + WS<SH>::JG; c.gGj = ws + WS<SH>::GJ; c.gval = ws + WS<SH>::GVAL;
    pb.limtab = kc.limtab;
    pb.x0 = bp.x0 + (size_t)14 * i; pb.xT = bp.xT + (size_t)14 * i;
    pb.gx = bp.gx + (size_t)14 * SH::NN * i; pb.gu = bp.gu + (size_t)7 * SH::NN * i;
    pb.glam = bp.glam ? bp.glam + (size_t)SH::NDUAL * i : nullptr;
    pb.gT = bp.gT[i];
    pb.sx = bp.sx + (size_t)14 * SH::NN * i; pb.su = bp.su + (size_t)7 * SH::NN * i; pb.sT = bp.sT + i;
    pb.slam = bp.slam ? bp.slam + (size_t)SH::NDUAL * i : nullptr;
    pb.stats = bp.stats ? bp.stats + i : nullptr;
}

// SQP stage kernel: kind 0 = stage_load, 1 = stage_linfac, 2 = stage_lsstep of SQP iteration `it`, 3 = stage_qp_generic (linearise,
// factorise and the ADMM iterations in one kernel: larger collocation shapes, adaptive rho).  gmat: one matrix block per CTA for the
// shapes whose KKT inverses do not fit shared memory (Layout::GMAT).
// GEN selects the kernel that holds kind 3 only, so that the code generation of the three light stages does not change with it.
#ifndef MPCB_STAGE_SMEM_OPAQUE
#define MPCB_STAGE_SMEM_OPAQUE 1
#endif
template <class SH, bool GEN>
__global__ void __launch_bounds__(SH::NT, 1)
sqp_stage_kernel(const __grid_constant__ KernelConst kc, const BatchPtrs bp, const int n, const int kind, const int it,
                 int* __restrict__ counter, double* __restrict__ workspace, double* __restrict__ gmat, long long* prof) {
    // (no static __shared__ variable next to the dynamic block: with two shared bases in one kernel ptxas 12.9 rematerialised the
    // dynamic base wrongly in the register-starved 1024-thread instances -- `ULEA URa, URb, URb, 0x18` -- and the CTA faulted;
    // tests/test_capi_cpu.py::test_no_miscompiled_shared_base greps the SASS for that pattern)
    extern __shared__ double smem[];
    const int tid = threadIdx.x;
    Ctx<SH> c;
#if MPCB_STAGE_SMEM_OPAQUE
    double* sbase = smem;                  // as in admm_kernel: the shared base held in a register instead of re-derived per phase
    asm volatile("" : "+l"(sbase));
    __builtin_assume(__isShared(sbase));
    c.bind(sbase, gmat + (size_t)blockIdx.x * Layout<SH>::CTA_BLOCK);
#else
    c.bind(smem, gmat + (size_t)blockIdx.x * Layout<SH>::CTA_BLOCK);
#endif
    volatile long long& s_next = *reinterpret_cast<volatile long long*>(c.isc + IS_NEXT);
    c.st = &kc.st;
    c.cheb = &kc.cheb;
    c.prof = prof;
    TS<SH> ts;
    setup_roles(c, ts, tid);
    for (;;) {
        if (tid == 0) s_next = atomicAdd(counter, 1);
        __syncthreads();
        const int i = (int)s_next;
        __syncthreads();
        if (i >= n) break;
        double* ws = workspace + WS<SH>::STRIDE * i;
        int* flags = reinterpret_cast<int*>(ws + WS<SH>::FLAGS);
        if (kind != 0 && flags[0]) continue;        // finished in an earlier SQP iteration
        Problem pb;
        bind_problem(c, pb, kc, bp, ws, i);
        if constexpr (GEN) {
            stage_qp_generic<SH>(c, ts, tid, pb, ws + WS<SH>::QB);
        } else {
            if (kind == 0) stage_load<SH>(c, ts, tid, pb, flags);
            else if (kind == 1) { if constexpr (SH::PO == 3 && SH::NS == 6) stage_linfac<SH>(c, ts, tid, pb, ws + WS<SH>::QB); }
            else stage_lsstep<SH>(c, ts, tid, pb, ws + WS<SH>::QB, flags, it);
        }
    }
}

// node kernels: one thread per (OCP, node) resp. (OCP, trial, node) of the whole batch (mpcb_node.h)
constexpr int NODE_NT = 64;
template <class SH>
__global__ void __launch_bounds__(NODE_NT)
node_jac_kernel(const int n, double* __restrict__ workspace) {
    const long long e = (long long)blockIdx.x * NODE_NT + threadIdx.x;
    if (e >= (long long)n * SH::NN) return;
    const int i = (int)(e / SH::NN), k = (int)(e - (long long)i * SH::NN);
    double* ws = workspace + WS<SH>::STRIDE * i;
    if (reinterpret_cast<const int*>(ws + WS<SH>::FLAGS)[0]) return;
    rbd::DerivScratch S;
    double Jg[176], g[8];
    node_jac<SH>(ws + WS<SH>::Z, k, S, Jg, g);
    for (int j = 0; j < 176; j++) ws[WS<SH>::JG + k * 176 + j] = Jg[j];
    for (int j = 0; j < 8; j++) ws[WS<SH>::GJ + k * 8 + j] = g[j];
}
template <class SH>
__global__ void __launch_bounds__(NODE_NT)
node_val_kernel(const int n, const int ntrial, const double tau, double* __restrict__ workspace) {
    const long long e = (long long)blockIdx.x * NODE_NT + threadIdx.x;
    const int per = ntrial * SH::NN;
    if (e >= (long long)n * per) return;
    const int i = (int)(e / per), r = (int)(e - (long long)i * per), t = r / SH::NN, k = r - t * SH::NN;
    double* ws = workspace + WS<SH>::STRIDE * i;
    if (reinterpret_cast<const int*>(ws + WS<SH>::FLAGS)[0]) return;
    double g[8];
    node_val<SH>(ws + WS<SH>::Z, ws + WS<SH>::QB + QBL<SH>::G_RES, t, tau, k, g);
    for (int j = 0; j < 8; j++) ws[WS<SH>::GVAL + (t * SH::NN + k) * 8 + j] = g[j];
}

#ifndef MPCB_SMEM_OPAQUE
#define MPCB_SMEM_OPAQUE 1
#endif
// the hot kernel: <= qp_max_iter box-ADMM iterations of one QP per CTA at a time
template <class SH>
__global__ void __launch_bounds__(ANT, 1)
admm_kernel(const __grid_constant__ KernelConst kc, const int n, int* __restrict__ counter, double* __restrict__ workspace, long long* prof) {
    extern __shared__ double smem[];
    __shared__ int s_next;
    __shared__ uint32_t s_tmem;
    const int ptid = threadIdx.x;       // physical thread: runs the logical threads ptid and ptid + ANT (mpcb_admm.h)
    ACtx<SH> c;
#if MPCB_SMEM_OPAQUE
    // the base of the dynamic shared memory as an opaque register value that is still known to be a shared-memory address: otherwise the
    // compiler re-derives it at every phase (S2UR SR_CgaCtaId + LEA + IMAD in front of the first load behind each barrier)
    double* sbase = smem;
    asm volatile("" : "+l"(sbase));
    __builtin_assume(__isShared(sbase));
    c.bind(sbase);
#else
    c.bind(smem);
#endif
    c.st = &kc.st;
    c.cheb = &kc.cheb;
    c.prof = prof;
    ATS<SH> ts0, ts1;
    // all 512 TMEM columns of the SM for this CTA (one CTA per SM): thread-private operand strips, mpcb_tmem.h
    if (ptid < 32) tm_alloc512(&s_tmem);
    tm_fence_before();
    __syncthreads();
    tm_fence_after();
    const uint32_t tmem_base = s_tmem;
    {
        const int w0 = ptid >> 5, w1 = (w0 + ANT / 32) & 15;      // logical warps; both sit in lane quadrant w0 % 4 (ts1 is unused when ANT == NT)
        const uint32_t lanes = (uint32_t)(32 * (w0 & 3)) << 16;
        ts0.tmaddr = tmem_base + lanes + 2 * tm_warp_base(w0);
        ts1.tmaddr = tmem_base + lanes + 2 * tm_warp_base(w1);
        // opaque to the compiler from here on: otherwise it re-derives the strip address from threadIdx in front of every tcgen05.ld
        // (5 integer instructions per load, 565 warp instructions per iteration, profiles/r2s_admm_kernel_instmix.txt)
        asm volatile("" : "+r"(ts0.tmaddr), "+r"(ts1.tmaddr));
    }
    for (;;) {
        if (ptid == 0) s_next = atomicAdd(counter, 1);
        __syncthreads();
        const int i = s_next;
        __syncthreads();
        if (i >= n) break;
        double* ws = workspace + WS<SH>::STRIDE * i;
        if (reinterpret_cast<const int*>(ws + WS<SH>::FLAGS)[0]) continue;
        c.g = ws + WS<SH>::QB;
        admm_run<SH>(c, ts0, ts1, ptid);
    }
    tm_fence_before();
    __syncthreads();
    if (ptid < 32) tm_free512(tmem_base);
}

// ---- read-out + accounting ------------------------------------------------------------------------------------------
constexpr int SAMPLE_NT = 256;

template <class SH>
__global__ void __launch_bounds__(SAMPLE_NT)
sample_kernel(const __grid_constant__ KernelConst kc, const double* __restrict__ sol_x, const double* __restrict__ sol_u,
              const double* __restrict__ sol_T, const double* __restrict__ target, const int npts, double* __restrict__ traj,
              mpcb_account* __restrict__ acc, const int n) {
    extern __shared__ double rows[];            // (npts+1) x 29
    __shared__ int flags[4];
    const int tid = threadIdx.x;
    for (int ocp = blockIdx.x; ocp < n; ocp += gridDim.x) {
        const double* sx = sol_x + (size_t)14 * SH::NN * ocp;
        const double* su = sol_u + (size_t)7 * SH::NN * ocp;
        const double T = sol_T[ocp];
        if (tid < 4) flags[tid] = 1;
        __syncthreads();
        for (int i = tid; i <= npts; i += SAMPLE_NT) {
            // polympc solution_x_at / solution_u_at: Lagrange interpolation on the segment containing t
            double t = 1.0 / npts * i;                      // motionPlanner.hpp:103
            double* row = rows + 29 * i;
            double tc = fmin(fmax(t, 0.0), 1.0);
            int s = (int)floor(tc * SH::NS);
            if (s > SH::NS - 1) s = SH::NS - 1;
            const double xi = 2.0 * (tc * SH::NS - s) - 1.0;
            double wgt[SH::PO + 1];
#pragma unroll
            for (int j = 0; j <= SH::PO; j++) {
                double num = 1, den = 1;
#pragma unroll
                for (int l = 0; l <= SH::PO; l++)
                    if (l != j) { num *= (xi - kc.cheb.xi[l]); den *= (kc.cheb.xi[j] - kc.cheb.xi[l]); }
                wgt[j] = num / den;
            }
            for (int cc = 0; cc < 21; cc++) row[1 + cc] = 0;
#pragma unroll
            for (int j = 0; j <= SH::PO; j++) {
                const int k = s * SH::PO + j;
                for (int cc = 0; cc < 7; cc++) {
                    row[1 + cc] += wgt[j] * sx[14 * k + cc];
                    row[8 + cc] += wgt[j] * sx[14 * k + 7 + cc];
                    row[15 + cc] += wgt[j] * su[7 * k + cc];
                }
            }
            double q[7], v[7], a[7], g[8];
            for (int cc = 0; cc < 7; cc++) { q[cc] = row[1 + cc]; v[cc] = row[8 + cc]; a[cc] = row[15 + cc]; }
            rbd::node_constraints_value(q, v, a, g);        // torque by RNEA (motionPlanner.hpp:110-113)
            for (int cc = 0; cc < 7; cc++) row[22 + cc] = g[cc];
            row[0] = t * T;                                 // motionPlanner.hpp:115
            if (acc) {
                // benchmark.cpp:140-156: Cartesian velocity limits and collision with the ground
                double oR[7][9], op[7][3], J[42], pos[3], tv[6];
                rbd::fk_world(q, oR, op);
                rbd::frame_jacobian(oR, op, PANDA(PANDA_TOOL_OFFSET), J, pos);
                for (int r = 0; r < 6; r++) { double sum = 0; for (int j = 0; j < 7; j++) sum += J[r * 7 + j] * v[j]; tv[r] = sum; }
                if (sqrt(tv[0] * tv[0] + tv[1] * tv[1] + tv[2] * tv[2]) > 1.7) flags[1] = 0;
                if (sqrt(tv[3] * tv[3] + tv[4] * tv[4] + tv[5] * tv[5]) > 2.5) flags[2] = 0;
                if (pos[2] < 0) flags[3] = 0;
            }
        }
        __syncthreads();
        if (acc) {
            const double dT = rows[29 * npts] / npts;       // benchmark.cpp:128
            for (int i = 1 + tid; i <= npts; i += SAMPLE_NT)
                for (int j = 0; j < 7; j++) {
                    const double jerk = (rows[29 * i + 15 + j] - rows[29 * (i - 1) + 15 + j]) / dT;
                    if (fabs(jerk) > 10 * kc.jmax[j]) flags[0] = 0;      // benchmark.cpp:135
                }
            if (tid < 28) {
                double mn = INFINITY, mx = -INFINITY;
                for (int i = 0; i <= npts; i++) { const double x = rows[29 * i + 1 + tid]; mn = fmin(mn, x); mx = fmax(mx, x); }
                acc[ocp].mins[tid] = mn; acc[ocp].maxs[tid] = mx;
            }
            if (tid < 14) acc[ocp].final_err[tid] = rows[29 * npts + 1 + tid] - target[(size_t)14 * ocp + tid];
        }
        if (traj)
            for (int e = tid; e < 29 * (npts + 1); e += SAMPLE_NT) traj[(size_t)ocp * 29 * (npts + 1) + e] = rows[e];
        __syncthreads();
        if (acc && tid < 4) acc[ocp].flags[tid] = flags[tid];
        __syncthreads();
    }
}

// ---- warm start (MotionPlanner::warm_start_RK, /root/reference/mpc_solver/motionPlanner.cpp:146-175) ---------------------------
// 8 lanes per problem, lane d < 7 owns joint d (lane 7 idles): step 1 per lane, max over the group, step 2 per lane, retry with
// a longer T while any joint sits in a blocked interval, then every lane samples its joint at the node times.  The per-joint
// routines are the ones the host path uses (csrc/mpcb_otg.h), so both paths take the same decisions.
constexpr int OTG_NT = 128;
// The profile of one problem by its group of 8 lanes (lane d < 7 owns joint d): on return every lane knows the common duration T
// and whether the group fell back to the cubic-Hermite profile; lane d holds the 7 phases (t, j) of its joint otherwise.
struct OtgLane { double q0, v0, q1, v1, vl, al, T, t[7], j[7]; bool act, fallback; };
__device__ __forceinline__ void otg_group_profile(const OtgConst& oc, const double* __restrict__ x0, const double* __restrict__ xT, int d, unsigned gmask,
                                                  OtgLane& L) {
    const bool act = d < 7;
    const int dd = act ? d : 0;
    const double q0 = x0[dd], v0 = x0[7 + dd], q1 = xT[dd], v1 = xT[7 + dd];
    const double vl = oc.vl[dd], al = oc.al[dd], jl = oc.jl[dd];
    double Tmin = 0, vco = 0;
    bool bad = false;
    if (act) {
        bad = !(vl > 0 && al > 0 && jl > 0) || fabs(v0) > vl * (1 + 1e-12) || fabs(v1) > vl * (1 + 1e-12);
        if (!bad) bad = !otg::min_time(q1 - q0, v0, v1, vl, al, jl, &Tmin, &vco);
    }
    bool fallback = (__ballot_sync(gmask, bad) & gmask) != 0;
    double T = act && !bad ? Tmin : 0.0, vc = vco;
    for (int o = 1; o < 8; o <<= 1) T = fmax(T, __shfl_xor_sync(gmask, T, o));
    if (!fallback) {
        bool all = false, need = false;
        double lo = 0, hi = 0;
        for (int round = 0; round < MPCB_OTG_ROUNDS && !all; round++) {
            bool good = true;
            double tn = T;
            need = act && Tmin < T * (1 - 1e-15) && !otg::is_hold(q1 - q0, v0, v1);
            if (need) good = otg::sync_scan(q1 - q0, v0, v1, vl, al, jl, T, vco, &lo, &hi, &tn);
            all = (__ballot_sync(gmask, !good) & gmask) == 0;
            if (all) break;
            for (int o = 1; o < 8; o <<= 1) tn = fmax(tn, __shfl_xor_sync(gmask, tn, o));
            if (!(tn < INFINITY)) break;              // some joint has no profile that lasts T or longer
            T = tn;                                   // the end of the blocked interval
        }
        fallback = !all || !(T > 0 && T <= oc.p_ub);
        if (!fallback && need) vc = otg::sync_bisect(q1 - q0, v0, v1, al, jl, T, lo, hi);
    }
    if (!fallback) {
        if (act) {
            const otg::Eval e = otg::eval_vc(q1 - q0, v0, v1, vc, al, jl);
            otg::dof_phases(v0, v1, vc, e.tc, al, jl, L.t, L.j);
        }
    } else {
        T = 0.05;
        for (;;) {
            const bool ok = !act || otg::hermite_ok(q1 - q0, v0, v1, T, vl, al);
            const bool allok = (__ballot_sync(gmask, !ok) & gmask) == 0;
            if (allok || T >= oc.p_ub) break;
            T *= 1.05;
            if (T > oc.p_ub) T = oc.p_ub;
        }
    }
    L.q0 = q0; L.v0 = v0; L.q1 = q1; L.v1 = v1; L.vl = vl; L.al = al; L.T = T; L.act = act; L.fallback = fallback;
}
// state of the lane's joint at time `time` in [0, T] (profile or fallback)
__device__ __forceinline__ void otg_lane_at(const OtgLane& L, double time, double* q, double* v, double* a) {
    if (!L.fallback) otg::dof_at_time(L.t, L.j, L.q0, L.v0, time, q, v, a);
    else otg::hermite_eval(L.q0, L.v0, L.q1, L.v1, L.T, L.T > 0 ? time / L.T : 0.0, q, v, a);
}

__global__ void __launch_bounds__(OTG_NT)
otg_guess_kernel(const __grid_constant__ OtgConst oc, const int n, const double* __restrict__ x0, const double* __restrict__ xT,
                 double* __restrict__ gx, double* __restrict__ gu, double* __restrict__ gT) {
    const int gt = blockIdx.x * OTG_NT + threadIdx.x, i = gt >> 3, d = gt & 7, lane = threadIdx.x & 31;
    if (i >= n) return;                                   // whole groups leave together
    const unsigned gmask = 0xFFu << (lane & 24);
    OtgLane L;
    otg_group_profile(oc, x0 + (size_t)14 * i, xT + (size_t)14 * i, d, gmask, L);
    const int N = oc.N;
    if (L.act)
        for (int k = 0; k < N; k++) {
            double q, v, a;
            if (!L.fallback) otg::dof_at_time(L.t, L.j, L.q0, L.v0, oc.tn[k] * L.T, &q, &v, &a);
            else otg::hermite_eval(L.q0, L.v0, L.q1, L.v1, L.T, oc.tn[k], &q, &v, &a);
            gx[((size_t)i * N + k) * 14 + d] = q; gx[((size_t)i * N + k) * 14 + 7 + d] = v; gu[((size_t)i * N + k) * 7 + d] = a;
        }
    if (d == 0) gT[i] = L.T;
}

// ---- the warm-start ("Ruckig") half of the benchmark: get_ruckig_trajectory<N> + the checks of examples/benchmark.cpp:72-114 ------
// One CTA per problem: the first 8 lanes build the profile (same routines as the warm start), every joint's samples go to shared
// memory, then the threads take one sample each: RNEA torque, Cartesian tool velocity, tool height, and the extrema / jerk pass.
constexpr int PROFILE_NT = 256;
__global__ void __launch_bounds__(PROFILE_NT)
sample_profile_kernel(const __grid_constant__ OtgConst oc, const __grid_constant__ KernelConst kc, const int n, const double* __restrict__ x0,
                      const double* __restrict__ xT, const int npts, double* __restrict__ traj, mpcb_account* __restrict__ acc, double* __restrict__ Tout) {
    extern __shared__ double rows[];            // (npts+1) x 29
    __shared__ int flags[4];
    __shared__ double sT;
    const int tid = threadIdx.x;
    for (int ocp = blockIdx.x; ocp < n; ocp += gridDim.x) {
        if (tid < 4) flags[tid] = 1;
        if (tid < 8) {
            OtgLane L;
            otg_group_profile(oc, x0 + (size_t)14 * ocp, xT + (size_t)14 * ocp, tid, 0xFFu, L);
            if (tid == 0) sT = L.T;
            if (L.act)
                for (int i = 0; i <= npts; i++) {
                    double q, v, a;
                    otg_lane_at(L, L.T / npts * i, &q, &v, &a);          // motionPlanner.hpp:77-85
                    rows[29 * i + 1 + tid] = q; rows[29 * i + 8 + tid] = v; rows[29 * i + 15 + tid] = a;
                }
        }
        __syncthreads();
        const double T = sT;
        for (int i = tid; i <= npts; i += PROFILE_NT) {
            double* row = rows + 29 * i;
            double q[7], v[7], a[7], g[8];
            for (int cc = 0; cc < 7; cc++) { q[cc] = row[1 + cc]; v[cc] = row[8 + cc]; a[cc] = row[15 + cc]; }
            rbd::node_constraints_value(q, v, a, g);        // torque by RNEA (motionPlanner.hpp:91-94)
            for (int cc = 0; cc < 7; cc++) row[22 + cc] = g[cc];
            row[0] = T / npts * i;
            if (acc) {
                double oR[7][9], op[7][3], J[42], pos[3], tv[6];
                rbd::fk_world(q, oR, op);
                rbd::frame_jacobian(oR, op, PANDA(PANDA_TOOL_OFFSET), J, pos);
                for (int r = 0; r < 6; r++) { double sum = 0; for (int j = 0; j < 7; j++) sum += J[r * 7 + j] * v[j]; tv[r] = sum; }
                if (sqrt(tv[0] * tv[0] + tv[1] * tv[1] + tv[2] * tv[2]) > 1.7) flags[1] = 0;      // benchmark.cpp:96-103
                if (sqrt(tv[3] * tv[3] + tv[4] * tv[4] + tv[5] * tv[5]) > 2.5) flags[2] = 0;
                if (pos[2] < 0) flags[3] = 0;                                                   // benchmark.cpp:108-113
            }
        }
        __syncthreads();
        if (acc) {
            const double dT = rows[29 * npts] / npts;       // benchmark.cpp:82
            for (int i = 1 + tid; i <= npts; i += PROFILE_NT)
                for (int j = 0; j < 7; j++) {
                    const double jerk = (rows[29 * i + 15 + j] - rows[29 * (i - 1) + 15 + j]) / dT;
                    if (fabs(jerk) > 10 * kc.jmax[j]) flags[0] = 0;      // benchmark.cpp:89
                }
            if (tid < 28) {
                double mn = INFINITY, mx = -INFINITY;
                for (int i = 0; i <= npts; i++) { const double x = rows[29 * i + 1 + tid]; mn = fmin(mn, x); mx = fmax(mx, x); }
                acc[ocp].mins[tid] = mn; acc[ocp].maxs[tid] = mx;
            }
            if (tid < 14) acc[ocp].final_err[tid] = rows[29 * npts + 1 + tid] - xT[(size_t)14 * ocp + tid];
        }
        if (traj)
            for (int e = tid; e < 29 * (npts + 1); e += PROFILE_NT) traj[(size_t)ocp * 29 * (npts + 1) + e] = rows[e];
        if (Tout && tid == 0) Tout[ocp] = T;
        __syncthreads();
        if (acc && tid < 4) acc[ocp].flags[tid] = flags[tid];
        __syncthreads();
    }
}

// ---- problem generator: one problem per thread (the routines of mpcb_gen.h, shared with mpcb_generate_problems) -----------------
constexpr int GEN_NT = 32;       // small CTAs: the rejection loop and the 6x6 solve are long serial chains, spread them over the SMs
__global__ void __launch_bounds__(GEN_NT)
gen_problems_kernel(const __grid_constant__ mpcb_limits lim, const int kind, const unsigned long long seed, const unsigned long long first,
                    const int n, double* __restrict__ x0, double* __restrict__ xT) {
    const int i = blockIdx.x * GEN_NT + threadIdx.x;
    if (i >= n) return;
    double a[14], b[14];
    gen::generate_one(&lim, kind, seed, first + (unsigned long long)i, a, b);
    for (int k = 0; k < 14; k++) { x0[(size_t)14 * i + k] = a[k]; xT[(size_t)14 * i + k] = b[k]; }
}

// ---- debug: the per-node device functions on arbitrary states (per-function parity tests) -------------------
__global__ void __launch_bounds__(NODE_NT)
node_debug_kernel(const int n, const double* __restrict__ qva, double* __restrict__ g_val, double* __restrict__ g_jac, double* __restrict__ Jg) {
    const int i = blockIdx.x * NODE_NT + threadIdx.x;
    if (i >= n) return;
    double q[7], v[7], a[7], g[8], J[176];
    for (int j = 0; j < 7; j++) { q[j] = qva[21 * i + j]; v[j] = qva[21 * i + 7 + j]; a[j] = qva[21 * i + 14 + j]; }
    rbd::node_constraints_value(q, v, a, g);
    for (int j = 0; j < 8; j++) g_val[8 * i + j] = g[j];
    rbd::DerivScratch S;
    rbd::node_constraints_jac(q, v, a, S, g, J);
    for (int j = 0; j < 8; j++) g_jac[8 * i + j] = g[j];
    for (int j = 0; j < 176; j++) Jg[176 * (size_t)i + j] = J[j];
}

// ---- FP64 FMA peak ---------------------------------------------------------------------------------------------------
__global__ void fp64_peak_kernel(double* out, int iters) {
    double a0 = threadIdx.x * 1e-3, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double b = 0.999999, cc = 1e-7;
    for (int i = 0; i < iters; i++) {
        a0 = fma(a0, b, cc); a1 = fma(a1, b, cc); a2 = fma(a2, b, cc); a3 = fma(a3, b, cc);
        a4 = fma(a4, b, cc); a5 = fma(a5, b, cc); a6 = fma(a6, b, cc); a7 = fma(a7, b, cc);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

// ---- launchers -------------------------------------------------------------------------------------------------------
// Instantiated collocation shapes (POLY_ORDER, NUM_SEG): the reference's (3,6) and BASELINE config 4's (3,12), (5,6), (7,4).
typedef Shape<3, 6> SH36;
#define MPCB_FOR_SHAPE(po, ns, ...)                                                           \
    do {                                                                                      \
        if ((po) == 3 && (ns) == 6) { typedef Shape<3, 6> SH; __VA_ARGS__; }                  \
        else if ((po) == 3 && (ns) == 12) { typedef Shape<3, 12> SH; __VA_ARGS__; }           \
        else if ((po) == 5 && (ns) == 6) { typedef Shape<5, 6> SH; __VA_ARGS__; }             \
        else if ((po) == 7 && (ns) == 4) { typedef Shape<7, 4> SH; __VA_ARGS__; }             \
        else return cudaErrorInvalidValue;                                                    \
    } while (0)

template <class SH>
static size_t admm_smem_of() {
    if constexpr (SH::PO == 3 && SH::NS == 6) return AL<SH>::BYTES;
    else return 0;
}
template <class SH>
static cudaError_t prepare_shape() {
    cudaError_t e = cudaFuncSetAttribute(sqp_stage_kernel<SH, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Layout<SH>::BYTES);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(sqp_stage_kernel<SH, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Layout<SH>::BYTES);
    if (e != cudaSuccess) return e;
    if constexpr (SH::PO == 3 && SH::NS == 6) e = cudaFuncSetAttribute(admm_kernel<SH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)AL<SH>::BYTES);
    return e;
}

int shape_supported(int po, int ns) { return (po == 3 && (ns == 6 || ns == 12)) || (po == 5 && ns == 6) || (po == 7 && ns == 4); }

cudaError_t shape_info(int po, int ns, ShapeInfo* si) {
    MPCB_FOR_SHAPE(po, ns, {
        si->stage_smem = Layout<SH>::BYTES; si->stage_threads = SH::NT; si->ws_doubles = WS<SH>::STRIDE;
        si->gmat_doubles = (size_t)Layout<SH>::CTA_BLOCK;
        si->admm_smem = admm_smem_of<SH>();
    });
    return cudaSuccess;
}

cudaError_t solve_prepare(int po, int ns) {
    MPCB_FOR_SHAPE(po, ns, return prepare_shape<SH>());
    return cudaSuccess;
}
cudaError_t launch_sqp_stage(const KernelConst& kc, const BatchPtrs& bp, int n, int kind, int it, int grid, int* counter, double* workspace,
                             double* gmat, long long* prof, cudaStream_t stream) {
    MPCB_FOR_SHAPE(kc.st.poly_order, kc.st.num_seg, {
        if (kind == 3) sqp_stage_kernel<SH, true><<<grid, SH::NT, Layout<SH>::BYTES, stream>>>(kc, bp, n, kind, it, counter, workspace, gmat, prof);
        else sqp_stage_kernel<SH, false><<<grid, SH::NT, Layout<SH>::BYTES, stream>>>(kc, bp, n, kind, it, counter, workspace, gmat, prof);
    });
    return cudaGetLastError();
}
cudaError_t launch_node_jac(const KernelConst& kc, int n, double* workspace, cudaStream_t stream) {
    MPCB_FOR_SHAPE(kc.st.poly_order, kc.st.num_seg, {
        const long long threads = (long long)n * SH::NN;
        node_jac_kernel<SH><<<(unsigned)((threads + NODE_NT - 1) / NODE_NT), NODE_NT, 0, stream>>>(n, workspace);
    });
    return cudaGetLastError();
}
cudaError_t launch_node_val(const KernelConst& kc, int n, double* workspace, cudaStream_t stream) {
    const int ntrial = kc.st.line_search_max_iter + 1;
    MPCB_FOR_SHAPE(kc.st.poly_order, kc.st.num_seg, {
        const long long threads = (long long)n * ntrial * SH::NN;
        node_val_kernel<SH><<<(unsigned)((threads + NODE_NT - 1) / NODE_NT), NODE_NT, 0, stream>>>(n, ntrial, kc.st.sqp_tau, workspace);
    });
    return cudaGetLastError();
}
cudaError_t launch_admm(const KernelConst& kc, int n, int grid, int* counter, double* workspace, long long* prof, cudaStream_t stream) {
    admm_kernel<SH36><<<grid, ANT, AL<SH36>::BYTES, stream>>>(kc, n, counter, workspace, prof);
    return cudaGetLastError();
}
cudaError_t launch_sample(const KernelConst& kc, const double* sx, const double* su, const double* sT, const double* target, int npts,
                          double* traj, mpcb_account* acc, int n, int grid, cudaStream_t stream) {
    const size_t smem = (size_t)29 * (npts + 1) * sizeof(double);
    MPCB_FOR_SHAPE(kc.st.poly_order, kc.st.num_seg, {
        cudaError_t e = cudaFuncSetAttribute(sample_kernel<SH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        sample_kernel<SH><<<grid, SAMPLE_NT, smem, stream>>>(kc, sx, su, sT, target, npts, traj, acc, n);
    });
    return cudaGetLastError();
}
cudaError_t launch_otg_guess(const OtgConst& oc, int n, const double* x0, const double* xT, double* gx, double* gu, double* gT, cudaStream_t stream) {
    const long long threads = 8ll * n;
    otg_guess_kernel<<<(unsigned)((threads + OTG_NT - 1) / OTG_NT), OTG_NT, 0, stream>>>(oc, n, x0, xT, gx, gu, gT);
    return cudaGetLastError();
}
cudaError_t launch_sample_profile(const OtgConst& oc, const KernelConst& kc, int n, const double* x0, const double* xT, int npts, double* traj,
                                  mpcb_account* acc, double* Tout, int grid, cudaStream_t stream) {
    const size_t smem = (size_t)29 * (npts + 1) * sizeof(double);
    cudaError_t e = cudaFuncSetAttribute(sample_profile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    sample_profile_kernel<<<grid, PROFILE_NT, smem, stream>>>(oc, kc, n, x0, xT, npts, traj, acc, Tout);
    return cudaGetLastError();
}
cudaError_t launch_gen_problems(const mpcb_limits& lim, int kind, unsigned long long seed, unsigned long long first, int n, double* x0, double* xT,
                                cudaStream_t stream) {
    gen_problems_kernel<<<(unsigned)((n + GEN_NT - 1) / GEN_NT), GEN_NT, 0, stream>>>(lim, kind, seed, first, n, x0, xT);
    return cudaGetLastError();
}
cudaError_t launch_node_debug(int n, const double* qva, double* g_val, double* g_jac, double* Jg, cudaStream_t stream) {
    node_debug_kernel<<<(unsigned)((n + NODE_NT - 1) / NODE_NT), NODE_NT, 0, stream>>>(n, qva, g_val, g_jac, Jg);
    return cudaGetLastError();
}
cudaError_t launch_fp64_peak(double* out, int blocks, int threads, int iters, cudaStream_t stream) {
    fp64_peak_kernel<<<blocks, threads, 0, stream>>>(out, iters);
    return cudaGetLastError();
}

// ---- scratch of the no-allocation solve (mpcb_solve_batch_noalloc; no counterpart in the reference, added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops) -----------------------------------------
// The per-OCP workspace (iterate, multipliers, node Jacobians, factorised QP block) and the work counters of the persistent kernels
// live in file-scope device arrays instead of the handle's cudaMalloc'ed buffers: MPCB_NOALLOC_MAX_OCPS problems of the base shape
// (4.3 MB; every GPU that loads libmpcb holds it until exit).  In host code the array names are host-side stand-ins, not device
// addresses: the device addresses are looked up once per device with cudaGetSymbolAddress (that device current) and handed to the
// same kernels the allocating path launches.  All handles and all calls share the arrays.
static constexpr size_t NOALLOC_WS_DOUBLES = (size_t)MPCB_NOALLOC_MAX_OCPS * WS<SH36>::STRIDE;
static constexpr int NOALLOC_COUNTERS = 64;
static __device__ __align__(256) double g_noalloc_workspace[NOALLOC_WS_DOUBLES];
static __device__ __align__(256) int g_noalloc_counters[NOALLOC_COUNTERS];

cudaError_t noalloc_scratch(int device, double** workspace, size_t* workspace_doubles, int** counters, int* n_counters) {
    static std::mutex mu;
    static double* ws_of[64];
    static int* ctr_of[64];
    static bool have[64];
    if (device < 0 || device >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!have[device]) {
        int cur = -1;
        cudaError_t e = cudaGetDevice(&cur);
        if (e != cudaSuccess) return e;
        if (cur != device) return cudaErrorInvalidDevice;       // the caller makes the handle's device current first
        void *pw = nullptr, *pc = nullptr;
        if ((e = cudaGetSymbolAddress(&pw, g_noalloc_workspace)) != cudaSuccess) return e;
        if ((e = cudaGetSymbolAddress(&pc, g_noalloc_counters)) != cudaSuccess) return e;
        ws_of[device] = static_cast<double*>(pw); ctr_of[device] = static_cast<int*>(pc);
        have[device] = true;
    }
    *workspace = ws_of[device]; *workspace_doubles = NOALLOC_WS_DOUBLES;
    *counters = ctr_of[device]; *n_counters = NOALLOC_COUNTERS;
    return cudaSuccess;
}

}  // namespace mpcb
