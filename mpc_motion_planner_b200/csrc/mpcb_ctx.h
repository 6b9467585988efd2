// Per-CTA context, per-thread persistent state and the lockstep-phase machinery of the OCP kernel.
//
// One CTA (NT threads) solves one OCP at a time.  The solver is written as a sequence of PHASES: inside a phase every
// thread works on its own items and never reads what another thread writes in the same phase; phases are separated by
// a CTA barrier.  The same source compiles in two modes:
//   * nvcc (default): __device__ code, a phase is `stmt; __syncthreads();`                      <- the product
//   * -DMPCB_EMUL (g++ only, tests/emul): a phase is a loop over all thread ids on the host     <- test-only lockstep
//     emulation used by the CPU test-suite to check the kernel's index logic without a GPU.  It is never linked into
//     libmpcb.so; the product has no CPU path.
#pragma once
#include <string.h>
#include "mpcb.h"
#include "mpcb_common.h"
#include "mpcb_rbd.h"
#include "mpcb_shape.h"

namespace mpcb {

constexpr int NT = 512;   // threads per CTA of the (3,6) kernels (mpcb_admm.h, mpcb_export.h); the stage phases use SH::NT

#ifdef MPCB_EMUL
#define MPCB_DEV inline
#define MPCB_DRIVER_PARAMS(SH) Ctx<SH>& c, TS<SH>* tsv
#define MPCB_DRIVER_ARGS c, tsv
#define MPCB_PHASE(...)                       \
    for (int tid = 0; tid < SH::NT; ++tid) {  \
        auto& ts = tsv[tid];                  \
        (void)ts;                             \
        __VA_ARGS__;                          \
    }
// a phase that needs no CTA barrier after it on the device (the next phase only exchanges registers inside a warp)
#define MPCB_PHASE_NS(...) MPCB_PHASE(__VA_ARGS__)
// value of `field` in the per-thread state of lane (tid ^ k)
#define MPCB_PEER(field, k) (tsv[tid ^ (k)].field)
#else
#define MPCB_DEV __device__ __forceinline__
#define MPCB_DRIVER_PARAMS(SH) Ctx<SH>& c, TS<SH>& ts, const int tid
#define MPCB_DRIVER_ARGS c, ts, tid
#define MPCB_PHASE(...)  \
    {                    \
        __VA_ARGS__;     \
        __syncthreads(); \
    }
#define MPCB_PHASE_NS(...) \
    {                      \
        __VA_ARGS__;       \
    }
#define MPCB_PEER(field, k) __shfl_xor_sync(0xffffffffu, ts.field, k)
#endif

// Per-phase cycle accounting (debug builds with -DMPCB_PROFILE only): thread 0 of CTA 0 accumulates clock64 deltas.
#if defined(MPCB_PROFILE) && !defined(MPCB_EMUL)
#define MPCB_PROF_DECL long long prof_last = clock64();
#define MPCB_PROF(slot)                                                  \
    if (tid == 0 && blockIdx.x == 0 && c.prof) {                         \
        const long long now_ = clock64();                                \
        c.prof[slot] += now_ - prof_last;                                \
        prof_last = now_;                                                \
    }
#else
#define MPCB_PROF_DECL
#define MPCB_PROF(slot)
#endif

// uniform double scalars in shared memory
enum {
    SC_PBAR = 0,   // p at the linearisation point
    SC_MTP,        // -ts * pbar
    SC_HDP,        // regularised H[p,p]
    SC_RHSP,       // rhs[p] gather part
    SC_MU, SC_V_L1, SC_V_LINF, SC_PHI0, SC_DPHI, SC_ALPHA, SC_COST_TRIAL, SC_PN, SC_DN, SC_GRADP,
    SC_RP, SC_RD, SC_EP, SC_ED,
    SC_COUNT = 24
};
// uniform int scalars
enum { IS_QP_DONE = 0, IS_QP_STATUS, IS_QP_ITERS, IS_LS_ACCEPT, IS_SQP_DONE, IS_NEXT = 7 /* the persistent CTA's next problem index */, IS_COUNT = 8 };
// atomic-max slots
enum { UM_RP = 0, UM_RD, UM_NAX, UM_NX, UM_NZ, UM_NHX, UM_NATY, UM_NH, UM_A, UM_B, UM_V, UM_COUNT = 12 };

// Tiling of the symmetric sweeps (mpcb_qp.h): 5 x 5 register tiles of the lower triangle, one thread per tile.
template <class SH>
struct Sweep {
    static constexpr int TILE = 5;
    static constexpr int tiles(int M) { return ((M + TILE - 1) / TILE) * ((M + TILE - 1) / TILE + 1) / 2; }
    static constexpr int MAXM = SH::AUGMAX > SH::NB ? SH::AUGMAX : SH::NB;
    static constexpr int COLS = ((MAXM + TILE - 1) / TILE * TILE + 1 + 7) / 8 * 8;     // one published column + 1/pivot in the last slot
    static constexpr int SLOT_I = tiles(SH::AUGMAX), SLOT_C = tiles(SH::NB);           // threads per interior matrix / for the separator
    static constexpr int PAR0 = SH::NT / SLOT_I;
    static constexpr int PAR = PAR0 > SH::NS ? SH::NS : (PAR0 > 6 ? 6 : PAR0);          // interior matrices swept per round
    // scratch of the blocked sweep per matrix: raw panel M x 5 | scaled panel M x 5 | inverse of the diagonal tile (25)
    static constexpr int PANEL = 2 * TILE * ((MAXM + TILE - 1) / TILE * TILE) + 32;
    static_assert(SLOT_I <= SH::NT && SLOT_C <= SH::NT && PAR >= 1, "sweep tiling: one thread per 5 x 5 tile of the largest matrix");
};

template <class SH>
struct Layout {   // offsets in doubles into the dynamic shared-memory block
    // The matrix area (W | F | C | BB) sits in shared memory when it fits (the base shape); for the larger collocation shapes
    // it lives in a per-CTA block of global memory (Ctx::bind gets the pointer) and only vectors + Jacobians stay on chip.
    static constexpr bool GMAT = (SH::MATSIZE + SH::BBSIZE) > 22000;
    static constexpr int MATDOUBLES = SH::MATSIZE + SH::BBSIZE + SH::FULLSIZE;     // packed factors + full copies (GMAT shapes)
    // per-CTA block in global memory: the matrices of the GMAT shapes (a token block otherwise)
    static constexpr int CTA_BLOCK = GMAT ? MATDOUBLES : 8;
    static constexpr int MAT = 0;                                  // W | F | C   (also: QP result)
    static constexpr int W = MAT, F = W + SH::WSIZE, C = F + SH::FSIZE;
    static constexpr int BB = MAT + SH::MATSIZE;
    static constexpr int JG = GMAT ? 0 : BB + SH::BBSIZE;
    static constexpr int VW = JG + SH::NN * 8 * 22;                // w      (rows)
    static constexpr int VRHS = VW + SH::VP;                       // rhs    (vars; rows in the residual check)
    static constexpr int VXT = VRHS + SH::VP;                      // x~     (vars)
    static constexpr int HA = VXT + SH::VP;                        // Hessian arrow (vars, kernel order)
    static constexpr int PCOL = HA + SH::VP;                       // A[:, p] of the equality rows
    static constexpr int PART = PCOL + SH::NEQ;                    // F^T rhs partials
    static constexpr int PSUM = PART + SH::PARTSIZE + 1;           // 2 x 32 partial sums
    static constexpr int COLBUF = (PSUM + 64 + 1) / 2 * 2;         // 2 x PAR published pivot columns
    // (shapes whose matrices live in global memory keep the panels of the blocked sweep here too; the base shape parks them in
    // the matrix area itself, whose content sits in the register tiles while a sweep runs)
    static constexpr int COLBUF_SIZE = 2 * Sweep<SH>::PAR * Sweep<SH>::COLS + (GMAT ? Sweep<SH>::PAR * Sweep<SH>::PANEL : 0);
    static constexpr int LIMS = COLBUF + COLBUF_SIZE;   // per-joint bounds 5x7 + 5, x0, xT
    static constexpr int SCAL = LIMS + 72;
    static constexpr int UMAX = SCAL + SC_COUNT;
    static constexpr int ISC = UMAX + UM_COUNT;
    static constexpr int TOTAL = ISC + IS_COUNT;                   // ints stored in 8-byte slots for simplicity
    static constexpr size_t BYTES = (size_t)TOTAL * 8;
    static_assert(SH::NDUAL + SH::VP <= SH::MATSIZE, "QP result must fit the matrix area");
    static_assert(BYTES <= 227 * 1024, "shared-memory block of the stage kernel exceeds 227 KB");
};
// Result block of one QP in global memory (per OCP): step p (kernel order, VP) | multipliers (NDUAL + 4) | info (4).
// The base shape embeds it in the ADMM kernel's QP block (specialisation in mpcb_export.h).
template <class SH>
struct QBL {
    static constexpr int G_RES = 0, R_LAM = SH::VP, R_INFO = SH::VP + SH::NDUAL + 4;
    static constexpr int QPBLOCK = G_RES + R_INFO + 4;
};
// LIMS sub-offsets
enum { LM_QLO = 0, LM_QHI = 7, LM_V = 14, LM_A = 21, LM_TAU = 28, LM_MINH = 35, LM_EPS, LM_PLB, LM_PUB, LM_X0 = 40, LM_XT = 54 };

template <class SH>
struct Ctx {
    double* sm;                    // shared-memory block
    double *W, *F, *C, *BB, *Jg, *w, *rhs, *xt, *Ha, *pcol, *part, *psum, *colbuf, *lims, *scal;
    double *Wf, *Fc, *Cf;          // full copies of the operators (global block; generic ADMM path only, null otherwise)
    unsigned long long* umax;
    long long* isc;
    double* z;                     // global workspace: iterate, kernel order (NVAR)
    double* lam;                   // global workspace: multipliers [eq | ineq | box kernel order]
    // global workspace filled by the node kernels (mpcb_node.h): inequality Jacobians and values at the iterate, and the
    // inequality values at the line-search trial points  z + alpha_t p,  t = 0 (alpha = 0), 1 (alpha = 1), 2 (tau), ..
    const double* gJg;             // NN x 176
    const double* gGj;             // NN x 8   (values returned by the derivative pass)
    const double* gval;            // (line_search_max_iter + 1) x NN x 8
    const mpcb_settings* st;
    const ChebConst* cheb;
    long long* prof;               // MPCB_PROFILE builds: cycle counters of CTA 0 (slots 16..31 belong to the SQP stages)
    // results of the last QP, aliased onto the matrix area after the ADMM loop
    double* qp_p;                  // step p, kernel order
    double* qp_lam;                // QP multipliers
    // block: this CTA's block in global memory (Layout::CTA_BLOCK doubles)
    MPCB_HD void bind(double* smem, double* block) {
        typedef Layout<SH> L;
        sm = smem;
        double* gmat = block;
        double* mat = L::GMAT ? gmat : sm + L::MAT;
        W = mat; F = mat + SH::WSIZE; C = F + SH::FSIZE; BB = mat + SH::MATSIZE; Jg = sm + L::JG; w = sm + L::VW; rhs = sm + L::VRHS;
        Wf = L::GMAT ? gmat + SH::MATSIZE + SH::BBSIZE : nullptr; Fc = L::GMAT ? Wf + SH::WFSIZE : nullptr; Cf = L::GMAT ? Fc + SH::FSIZE : nullptr;
        xt = sm + L::VXT; Ha = sm + L::HA; pcol = sm + L::PCOL; part = sm + L::PART; psum = sm + L::PSUM;
        colbuf = sm + L::COLBUF; lims = sm + L::LIMS; scal = sm + L::SCAL;
        umax = reinterpret_cast<unsigned long long*>(sm + L::UMAX);
        isc = reinterpret_cast<long long*>(sm + L::ISC);
        qp_p = mat; qp_lam = mat + SH::VP;
    }
};

// Per-thread persistent state ("registers" on the device).
template <class SH>
struct TS {
    // ---- row role: thread r < NA owns constraint row r
    double l, u, rho, zA, yA, pc;
    double dco[SH::PO + 1];
    int ridx[SH::PO + 2];
    // ---- variable role: thread i < NVAR owns variable i (kernel order)
    double x, zB, yB, lb, ub, rhoB, ha, hd;
    double gco[2 * SH::PO];
    int grow[2 * SH::PO];
    int gn, gfrow, gk, gcol;
    int vk, vcomp, vseg, vloc;
    // ---- factorisation: the 5 x 5 tile of an augmented segment matrix this thread sweeps in registers (mpcb_qp.h)
    double tile[25];
    int tI, tJ, tseg;
#ifdef MPCB_EMUL
    double acc;                  // host emulation of a cross-lane exchange (mpcb_gadmm.h)
#endif
};

// ---- small helpers ------------------------------------------------------------------------------------------------
MPCB_DEV void umax_accum(unsigned long long* slot, double v) {
#ifdef MPCB_EMUL
    unsigned long long b;
    memcpy(&b, &v, 8);
    if (b > *slot) *slot = b;
#else
    unsigned long long b = (unsigned long long)__double_as_longlong(v);
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        unsigned long long t = __shfl_xor_sync(0xffffffffu, b, o);
        b = (t > b) ? t : b;
    }
    if ((threadIdx.x & 31) == 0) atomicMax(slot, b);
#endif
}
// four maxima at once: the four shuffle butterflies are independent, interleaved they cost the latency of one
MPCB_DEV void umax_accum4(unsigned long long* s0, double v0, unsigned long long* s1, double v1, unsigned long long* s2, double v2,
                          unsigned long long* s3, double v3) {
#ifdef MPCB_EMUL
    umax_accum(s0, v0); umax_accum(s1, v1); umax_accum(s2, v2); umax_accum(s3, v3);
#else
    unsigned long long b0 = (unsigned long long)__double_as_longlong(v0), b1 = (unsigned long long)__double_as_longlong(v1);
    unsigned long long b2 = (unsigned long long)__double_as_longlong(v2), b3 = (unsigned long long)__double_as_longlong(v3);
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        const unsigned long long t0 = __shfl_xor_sync(0xffffffffu, b0, o), t1 = __shfl_xor_sync(0xffffffffu, b1, o);
        const unsigned long long t2 = __shfl_xor_sync(0xffffffffu, b2, o), t3 = __shfl_xor_sync(0xffffffffu, b3, o);
        b0 = (t0 > b0) ? t0 : b0; b1 = (t1 > b1) ? t1 : b1; b2 = (t2 > b2) ? t2 : b2; b3 = (t3 > b3) ? t3 : b3;
    }
    if ((threadIdx.x & 31) == 0) { atomicMax(s0, b0); atomicMax(s1, b1); atomicMax(s2, b2); atomicMax(s3, b3); }
#endif
}
MPCB_DEV double umax_get(const unsigned long long* slot) {
    double v;
    unsigned long long b = *slot;
    memcpy(&v, &b, 8);
    return v;
}
MPCB_DEV double dmax(double a, double b) { return a > b ? a : b; }
MPCB_DEV double dmin(double a, double b) { return a < b ? a : b; }

// Box bounds of variable (k, comp) -- MotionPlanner::set_constraint_margins / set_current_state / set_target_state
// (/root/reference/mpc_solver/motionPlanner.cpp:27-54,64-80).  Pins are always applied last.
template <class SH>
MPCB_DEV void var_bounds(const Ctx<SH>& c, int k, int comp, double& lo, double& hi) {
    const double* L = c.lims;
    if (k < 0) { lo = L[LM_PLB]; hi = L[LM_PUB]; return; }
    if (comp < 14 && k == 0) { lo = hi = L[LM_X0 + comp]; return; }
    if (comp < 14 && k == SH::NN - 1) { lo = L[LM_XT + comp] - L[LM_EPS]; hi = L[LM_XT + comp] + L[LM_EPS]; return; }
    if (comp < 7) { lo = L[LM_QLO + comp]; hi = L[LM_QHI + comp]; return; }
    if (comp < 14) { hi = L[LM_V + comp - 7]; lo = -hi; return; }
    hi = L[LM_A + comp - 14]; lo = -hi;
}
// Bounds of inequality output rr (0-6 torque, 7 tool height) -- motionPlanner.cpp:82-97
template <class SH>
MPCB_DEV void ineq_bounds(const Ctx<SH>& c, int rr, double& lo, double& hi) {
    if (rr < 7) { hi = c.lims[LM_TAU + rr]; lo = -hi; }
    else { lo = c.lims[LM_MINH]; hi = INFINITY; }
}
// boxADMM's per-row penalty: loose rows 1e-6, equality rows 1e3*rho, else rho
MPCB_DEV double row_rho(double lo, double hi, double rho) {
    if (lo < -1e16 && hi > 1e16) return 1e-6;
    return (fabs(hi - lo) < 1e-4) ? 1e3 * rho : rho;
}

// Static (problem-independent) part of the per-thread roles.  All TS arrays are indexed with compile-time constants
// after unrolling so that the struct stays in registers.
template <class SH>
MPCB_DEV void setup_roles(const Ctx<SH>& c, TS<SH>& ts, int tid) {
    const ChebConst& cb = *c.cheb;
    MPCB_UNROLL
    for (int l = 0; l < SH::PO + 2; l++) ts.ridx[l] = 0;
    MPCB_UNROLL
    for (int l = 0; l <= SH::PO; l++) ts.dco[l] = 0;
    if (tid < SH::NEQ) {
        const int k = tid / 14, cc = tid % 14, s = k / SH::PO, j = k - s * SH::PO;
        MPCB_UNROLL
        for (int l = 0; l <= SH::PO; l++) { ts.ridx[l] = SH::kvar(s * SH::PO + l, cc); ts.dco[l] = cb.D[j][l]; }
        ts.ridx[SH::PO + 1] = (cc < 7) ? SH::kvar(k, 7 + cc) : SH::kvar(k, 14 + (cc - 7));
    } else if (tid < SH::NA) {
        const int k = (tid - SH::NEQ) / 8;
        ts.ridx[0] = SH::kvar(k, 0);
        ts.ridx[1] = SH::kvar(k, 14);
    }
    ts.gn = 0; ts.gfrow = -1; ts.gk = -1; ts.gcol = 0; ts.vk = -1; ts.vcomp = 0; ts.vseg = -1; ts.vloc = 0;
    MPCB_UNROLL
    for (int n = 0; n < 2 * SH::PO; n++) { ts.grow[n] = 0; ts.gco[n] = 0; }
    ts.l = ts.u = ts.rho = ts.zA = ts.yA = ts.pc = 0;
    ts.x = ts.zB = ts.yB = ts.lb = ts.ub = ts.rhoB = ts.ha = ts.hd = 0;
    if (tid < SH::NVAR) {
        int k, comp, seg, loc;
        SH::decode(tid, k, comp, seg, loc);
        ts.vk = k; ts.vcomp = comp; ts.vseg = seg; ts.vloc = loc;
        if (k >= 0) {
            ts.gk = k; ts.gcol = comp;
            if (comp < 14) {
                // collocation rows containing x_{k,comp}: rows of segment s1 (k is its local node l1) in slots [0,PO),
                // and, for a junction, rows of segment s1-1 (k is its end node) in slots [PO,2PO)
                const int s1 = k / SH::PO, l1 = k - s1 * SH::PO;
                const bool own = k < SH::NN - 1, prev = (l1 == 0 && k > 0);
                MPCB_UNROLL
                for (int j = 0; j < SH::PO; j++) {
                    if (own) { ts.grow[j] = 14 * (s1 * SH::PO + j) + comp; ts.gco[j] = cb.D[j][l1]; }
                    if (prev) { ts.grow[SH::PO + j] = 14 * ((s1 - 1) * SH::PO + j) + comp; ts.gco[SH::PO + j] = cb.D[j][SH::PO]; }
                }
                if (comp >= 7 && k < SH::NN - 1) ts.gfrow = 14 * k + (comp - 7);
            } else if (k < SH::NN - 1) {
                ts.gfrow = 14 * k + 7 + (comp - 14);
            }
        }
    }
}

}  // namespace mpcb
