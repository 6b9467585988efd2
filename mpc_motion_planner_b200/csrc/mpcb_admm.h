// The box-ADMM iteration kernel body (polympc boxADMM::solve restated in SURVEY.md A.5) on top of the factorisation
// of mpcb_qp.h.  This is the hot loop of the whole path: <= 700 iterations per QP, 2 QPs per OCP
// (/root/reference/mpc_solver/motionPlanner.cpp:15-20,184).
//
//   rhs   = sigma x - h + rho_b z_b - y_b + A^T (rho_A z_A - y_A)                                   phase A
//   x~    = S^-1 rhs  via  t = W rhs_I | part = S_BI t (= F^T rhs_I, but sparse) | x~_B = C^-1 (rhs_B - part) | x~_I = t - F x~_B
//                                                                                                     phases B, B', C, D
//   z~_A  = A x~ ; x = a x~ + (1-a) x ; z = clip(a z~ + (1-a) z + y/rho) ; y += rho (a z~ + (1-a) z_old - z)   phase E
// Every `check_termination` iterations the OSQP residual test runs (chk_*).
//
// Data placement (one CTA of NT = 512 threads per OCP, one CTA per SM):
//   * every thread owns (part of) one row of an interior inverse W_s and keeps 32 of its coefficients in REGISTERS for the
//     whole QP (16 K doubles of the 18.5 K of all W_s); the rest of W, F, C^-1 and the inequality Jacobian sit in shared
//     memory in conflict-free layouts; shared memory delivers 16 doubles/clk against 64 FMA/clk of the FP64 pipe, so
//     everything that can live in registers does;
//   * row r's (z_A, y_A, l, u, rho) and variable i's (x, z_b, y_b, lb, ub, rho_b) live in registers of thread r / i;
//   * vectors use the "padded node order": index(k, comp) = 66 s + 21 j + comp with s = min(k/3, 5), j = k - 3 s, p at 414;
//     every segment interior and every junction state is contiguous and starts on an even index (16-byte loads); the
//     segment stride 66 (= 4 banks mod 32) keeps the broadcast loads of two segments in one warp on different banks.
// The QP data arrives from the factorisation kernel through a per-OCP block in global memory (layout: struct AL).
#pragma once
#include "mpcb_ctx.h"
#include "mpcb_tmem.h"

namespace mpcb {

template <class SH>
struct AL {
    static_assert(SH::PO == 3 && SH::NS == 6, "the ADMM kernel's thread mapping is specialised for POLY_ORDER=3, NUM_SEG=6");
    static constexpr int SEG = 66, NV = 416, IP = 414;  // padded node order: segment stride, length, index of p
#ifndef MPCB_ADMM_WIDE
#define MPCB_ADMM_WIDE 1                               // 1: one logical thread per physical thread (512 x 128 registers); 0: two (256 x 255)
#endif
    // W coefficients per thread held in registers; the rest of a thread's block sits in its TMEM strip in chunks of 16
    static constexpr int WREG = MPCB_ADMM_WIDE ? 20 : 32;
    // number of 16-slot overflow chunks of the three thread kinds (segment 0: 32 coefficients, middle: 49, segment 5: 36)
    MPCB_HD static constexpr int wchunks(int kind) { return ((kind == 0 ? 32 : (kind == 1 ? 49 : 36)) - WREG + 15) / 16; }
    // ---- shared-memory image (copied verbatim from the global QP block) --------------------------------------------
    // inequality Jacobian, twice: row-major for A x (row stride 26, node stride 208; a row starts one slot late when its
    // node's first variable sits on an odd index, so that 16-byte loads of the row line up with 16-byte loads of x) and
    // transposed for A^T y (8 row entries per variable, stride 10, node stride 210).  The strides keep the eight 16-byte
    // pieces of a quarter warp on distinct bank groups.
    static constexpr int I_JG = 0, JGR = 26, JGS = 8 * JGR;
    static constexpr int I_GT = I_JG + 19 * JGS, GTC = 10, GTS = 21 * GTC;
    static constexpr int I_HD = I_GT + 19 * GTS;       // regularised Hessian diagonal (padded node order)
    static constexpr int I_HA = I_HD + NV;             // Hessian arrow H[p, .]
    static constexpr int I_PCOL = I_HA + NV;           // A[:, p] of the equality rows (256)
    static constexpr int I_PCV = I_PCOL + 256;         // A[r, p] of every row r (416, zero padded)
    static constexpr int I_FPC = I_PCV + 416;          // F_s[row, p] at the padded-node-order index of the interior variable (416)
    static constexpr int I_DT = I_FPC + 416;           // D (16) | -ts*pbar | pad
    static constexpr int I_SBI = I_DT + 24;            // S_BI units: [SBI_E][SBI_T] coefficients (phase B')
    static constexpr int IMG = I_SBI + 7 * 288;
    // ---- thread-private strips in tensor memory (mpcb_tmem.h), slot = one double -----------------------------------
    //   warps 0..3   (segment-0 blocks) : consts 0..7 | W overflow chunks | F half row (8) | C^-1 quarter row (18)
    //   warps 4..10  (middle blocks)    : consts 0..7 | W overflow chunks | F row (29)
    //   warps 11..15 (segment-5 blocks) : consts 0..7 | W overflow chunks | F half row (8) | C^-1 quarter row (18)
    // consts = l, u, rho, 1/rho of the thread's row | lb, ub, rho_b, 1/rho_b of the thread's variable
    static constexpr int TM_CST = 0;
    MPCB_HD static int tm_size(int w) { return (w >= 4 && w <= 10) ? 72 : 56; }
    MPCB_HD static int tm_wov(int tid) { return 8; }
    MPCB_HD static int tm_frow(int tid) { return 8 + 16 * wchunks(tid < 128 ? 0 : (tid < 352 ? 1 : 2)); }
    MPCB_HD static int tm_c(int tid) { return tm_frow(tid) + 8; }
    static_assert(8 + 16 * wchunks(0) + 8 + 18 <= 56 && 8 + 16 * wchunks(1) + 29 <= 72 && 8 + 16 * wchunks(2) + 8 + 18 <= 56, "TMEM strips");
    // C^-1 unit of a thread (4 units per separator row): threads 0..127 and 352..511, -1 elsewhere / beyond 4 NB
    MPCB_HD static int cunit(int tid) { const int u = tid < 128 ? tid : (tid >= 352 ? tid - 224 : -1); return u < 4 * SH::NB ? u : -1; }
    // ---- working vectors behind the image --------------------------------------------------------------------------
    // S_BI units (phase B'): 4 lanes per junction variable, SBI_E candidate entries per lane; V_SBI holds their vector indices (ints)
    static constexpr int SBI_T = 288, SBI_E = 7;
    static constexpr int V_RHS = IMG, V_XT = V_RHS + NV, V_W = V_XT + NV, V_SBI = V_W + NV, V_RB = V_SBI + (SBI_T * SBI_E + 1) / 2;
    static constexpr int V_PSUM = V_RB + 72, V_SC = V_PSUM + 64, V_UMAX = V_SC + 8, V_ISC = V_UMAX + UM_COUNT;
    static constexpr int PTID = 416;                   // thread that owns the variable p (lane 0 of warp 13)
    static constexpr int TOTAL = V_ISC + 8;
    static constexpr size_t BYTES = (size_t)TOTAL * 8;
    // ---- global QP block ---------------------------------------------------------------------------------------------
    static constexpr int G_WREG = 0;                   // [32][NT]
    static constexpr int G_TM = G_WREG + WREG * NT;    // [TM_SLOTS][NT]  the strips, slot-major
    static constexpr int G_IMG = G_TM + TM_SLOTS * NT;
    static constexpr int G_RES = G_IMG + IMG;       // step p (kernel order, VP) | multipliers (NDUAL+4) | info (4)
    static constexpr int R_LAM = SH::VP, R_INFO = SH::VP + SH::NDUAL + 4;
    static constexpr int QPBLOCK = G_RES + R_INFO + 4;

    // row role: thread tid owns constraint row rowof(tid).  The inequality rows start on a warp boundary (thread 256): a warp that mixed
    // equality and inequality rows would run both branches of the row product one after the other and be the last one of phase E.
#ifndef MPCB_ROW_ALIGN
#define MPCB_ROW_ALIGN 1
#endif
    static constexpr int ROW_GAP = MPCB_ROW_ALIGN ? (32 - SH::NEQ % 32) % 32 : 0;
    MPCB_HD static int rowof(int tid) {
        if (tid < SH::NEQ) return tid;
        const int r = tid - ROW_GAP;
        return (r >= SH::NEQ && r < SH::NA) ? r : -1;
    }
    MPCB_HD static int rtid(int r) { return r < SH::NEQ ? r : r + ROW_GAP; }
    MPCB_HD static int vidx(int k, int comp) { const int s = k / 3 > 5 ? 5 : k / 3; return SEG * s + 21 * (k - 3 * s) + comp; }
    MPCB_HD static int sepidx(int b) { return b < 70 ? SEG * (b / 14 + 1) + b % 14 : IP; }
    MPCB_HD static int ibase(int s) { return s == 0 ? 0 : SEG * s + 14; }                        // first interior index of segment s
    // decode a padded-node-order index: false for pads
    MPCB_HD static bool decode(int i, int& k, int& comp) {
        if (i == IP) { k = -1; comp = 0; return true; }
        if (i > IP) return false;
        const int s = (i / SEG) > 5 ? 5 : (i / SEG), rem = i - SEG * s;
        if (s < 5 && rem >= 63) return false;
        const int j = rem / 21;
        k = 3 * s + j; comp = rem - 21 * j;
        return true;
    }
};

// Role of a thread in the interior mat-vecs (phases B and D).
//   Phase B (t = W rhs_I) works on BLOCKS: 8 lanes share a group of R rows of one W_s; lane q multiplies the columns q, q + 8,
//   q + 16, .. of all R rows (R independent accumulators, one shared-memory load of the vector entry per R multiply-adds; the 8
//   lanes of a group read 8 consecutive vector entries per step), then a transpose-reduce over the 8 lanes (xor 4, 2, 1) leaves
//   every lane with the full sum of ONE row:
//       segment 0      : 16 groups of R = 4 rows, 8 column steps  -> 32 coefficients per thread (registers)
//       middle segments: 7 groups of R = 7 rows, 7 column steps   -> 49 coefficients per thread (32 registers + 17 TMEM)
//       segment 5      : 18 groups of R = 4 rows, 9 column steps  -> 36 coefficients per thread (32 registers + 4 TMEM)
//   The row a lane ends up with is the row it owns in phase D (x~_I = t - F x~_B): with R = 4 two lanes end with the same row
//   and split its F row in halves (paired), with R = 7 one lane per row (the eighth lane of the group idles in phase D).
struct WRole {
    bool valid, paired;     // phase D: the lane owns (half of) row `row`
    int seg, row, half;
    int out;                // padded-node-order index of the row
    int kind;               // 0: segment 0, 1: middle segment, 2: segment 5
    int grow0, q;           // phase B: first row of the lane's group, column lane
};
template <class SH>
MPCB_HD WRole wrole(int tid) {
    typedef AL<SH> L;
    WRole r;
    r.q = tid & 7;
    const int b2 = (r.q >> 2) & 1, b1 = (r.q >> 1) & 1, b0 = r.q & 1;
    if (tid < 128) {
        r.kind = 0; r.seg = 0; r.grow0 = 4 * (tid >> 3);
        r.row = r.grow0 + 2 * b2 + b1; r.half = b0; r.paired = true; r.valid = r.row < 63; r.out = r.row;
    } else if (tid < 352) {
        const int idx = tid - 128;
        r.kind = 1; r.seg = 1 + idx / 56; r.grow0 = 7 * ((idx % 56) >> 3);
        const int rl = 4 * b2 + 2 * b1 + b0;
        r.row = r.grow0 + rl; r.half = 0; r.paired = false; r.valid = rl < 7;
        r.out = L::SEG * r.seg + 14 + r.row;
    } else {
        const int idx = tid - 352;
        r.kind = 2; r.seg = 5; r.grow0 = 4 * (idx >> 3);
        r.row = r.grow0 + 2 * b2 + b1; r.half = b0; r.paired = true; r.valid = idx < 144 && r.row < 70;
        r.out = 5 * L::SEG + 14 + r.row;
    }
    if (!r.valid) { r.row = 0; r.out = 0; }
    return r;
}
// rows per group / column steps of the three kinds
MPCB_HD int wrows(int kind) { return kind == 1 ? 7 : 4; }
MPCB_HD int wsteps(int kind) { return kind == 0 ? 8 : (kind == 1 ? 7 : 9); }
// (row, col) of W_seg that coefficient slot e of the thread multiplies: e = step * R + r (e < 32: registers, e >= 32: TMEM
// overflow); false for padding slots.  The exporter and the kernel share this map.
template <class SH>
MPCB_HD bool wslot(const WRole& r, int tid, int e, int& row, int& col) {
    const int R = wrows(r.kind);
    if (e >= R * wsteps(r.kind) || (r.kind == 2 && tid >= 352 + 144)) return false;
    row = r.grow0 + e % R; col = r.q + 8 * (e / R);
    return row < SH::m(r.seg) && col < SH::m(r.seg);
}

// part = S_BI t (phase B'): the rows of the KKT matrix S that belong to the junction states x_{3 sj} (sj = 1..5), restricted to
// the interior variables.  F^T rhs_I = S_BI W rhs_I = S_BI t, and S_BI is sparse: junction variable (3 sj, c) of joint j = c % 7
// only meets (q_j, v_j, u_j) at the other nodes of its two segments (collocation rows of joint j) and the controls of its own
// node (inequality rows of node 3 sj).  Candidate e = 0..24 of junction variable b = 14 (sj - 1) + c:
//      0..1   x_{0, j}, x_{0, 7+j}                        (left segment is segment 0: its first state is interior)
//      2      u_{3 sj - 3, j}
//      3..8   (q_j, v_j, u_j) at nodes 3 sj - 2, 3 sj - 1
//      9..15  u_{3 sj, 0..6}
//      16..21 (q_j, v_j, u_j) at nodes 3 sj + 1, 3 sj + 2
//      22..24 (q_j, v_j, u_j) at node 18                  (right segment is segment 5: the final node is interior)
// A superset of the structural non-zeros (entries that are structurally zero carry a zero coefficient).  Four lanes share
// one junction variable: lane q takes the candidates q, q + 4, .., q + 24.  The exporter and the kernel share this map.
template <class SH>
MPCB_HD bool sbi_cand(int b, int e, int& seg, int& k, int& comp) {
    const int sj = b / 14 + 1, c = b % 14, j = c % 7, kj = 3 * sj;
    if (e < 9) {
        seg = sj - 1;
        if (e < 2) { k = 0; comp = e == 0 ? j : 7 + j; return sj == 1; }
        if (e == 2) { k = kj - 3; comp = 14 + j; return true; }
        const int q = e - 3;
        k = kj - 2 + q / 3; comp = 7 * (q % 3) + j;
        return true;
    }
    seg = sj;
    if (e < 16) { k = kj; comp = 14 + (e - 9); return true; }
    if (e < 22) { const int q = e - 16; k = kj + 1 + q / 3; comp = 7 * (q % 3) + j; return true; }
    if (e < 25) { k = 18; comp = 7 * (e - 22) + j; return sj == 5; }
    return false;
}
// boundary column of junction variable b = 14 (sj - 1) + c inside segment `seg` (sj - 1: its right junction, sj: its left junction)
template <class SH>
MPCB_HD int sbi_col(int b, int seg) {
    const int sj = b / 14 + 1, c = b % 14;
    return seg == sj ? c : (seg > 0 ? 14 : 0) + c;
}

template <class SH>
struct ACtx {
    double* sm;
    double *Jg, *Gt, *hd, *ha, *pcol, *pcv, *fpc, *dt, *sbc;
    double *rhs, *xt, *w, *rB, *psum, *sc;
    int* sbi;                      // [SBI_E][SBI_T] vector indices of the S_BI units
    unsigned long long* umax;
    long long* isc;
    double* g;                     // this OCP's QP block in global memory
    const mpcb_settings* st;
    long long* prof;               // MPCB_PROFILE builds: per-phase cycle counters of CTA 0 (else unused)
    int dbg_iter;                  // MPCB_PROFILE builds: the iteration counter, for time stamps inside the phase functions
    MPCB_HD void bind(double* smem) {
        typedef AL<SH> L;
        sm = smem;
        Jg = sm + L::I_JG; Gt = sm + L::I_GT; hd = sm + L::I_HD;
        ha = sm + L::I_HA; pcol = sm + L::I_PCOL; pcv = sm + L::I_PCV; fpc = sm + L::I_FPC; dt = sm + L::I_DT; sbc = sm + L::I_SBI;
        rhs = sm + L::V_RHS; xt = sm + L::V_XT; w = sm + L::V_W; sbi = reinterpret_cast<int*>(sm + L::V_SBI); rB = sm + L::V_RB;
        psum = sm + L::V_PSUM; sc = sm + L::V_SC;
        umax = reinterpret_cast<unsigned long long*>(sm + L::V_UMAX);
        isc = reinterpret_cast<long long*>(sm + L::V_ISC);
    }
    // collocation differentiation matrix: from the constant bank (a separate path from the shared-memory pipe that bounds phases A and E)
    // or from the shared-memory image (MPCB_DCONST 0)
#ifndef MPCB_DCONST
#define MPCB_DCONST 1
#endif
    const ChebConst* cheb;
    MPCB_HD double D(int j, int l) const { return MPCB_DCONST ? cheb->D[j][l] : dt[4 * j + l]; }
    MPCB_HD double mtp() const { return dt[16]; }
};

// Per-thread persistent state of the ADMM kernel.
template <class SH>
struct ATS {
    double wreg[AL<SH>::WREG];
    double zA, yA;                         // row role (thread r < NA); l, u, rho, 1/rho live in the thread's TMEM strip
    double x, zB, yB, rhoB;                // variable role (thread i < 404 in padded node order; p on thread PTID); lb, ub, 1/rho_b in TMEM
    double t, acc, acc2;                   // interior row: W rhs_I ; cross-lane exchange slots
#ifdef MPCB_EMUL
    double pv[8];                          // host emulation of the transpose-reduce of phase B: the lane's R partial sums
    double tm[TM_SLOTS];                   // host emulation of the thread's TMEM strip
    const double* tma() const { return tm; }
#else
    uint32_t tmaddr;                       // TMEM address of the thread's strip (lane quadrant | first column)
    __device__ __forceinline__ uint32_t tma() const { return tmaddr; }
#endif
};

// The ADMM kernel runs its NT = 512 LOGICAL threads (the unit of every role map in this file) on ANT = 256 physical
// threads: physical thread p executes logical threads p and p + 256 one after the other inside every phase.  A CTA of 256
// threads may use 255 registers per thread, so each logical thread's phase body has ~90 registers for operands in flight
// next to the 2 x 64 registers of W coefficients (with 512 physical threads it had ~45, which serialised every phase into
// load -> use round trips, profiles/r1j_admm_kernel_hotlines.txt).  Logical warps w and w + 8 share physical warp w: same
// lanes (shuffles are unchanged) and the same TMEM lane quadrant (w % 4).
constexpr int ANT = MPCB_ADMM_WIDE ? NT : NT / 2;
#ifdef MPCB_EMUL
#define MPCB_ADRV_PARAMS(SH) ACtx<SH>& c, ATS<SH>* tsv
#define MPCB_ADRV_ARGS c, tsv
#define MPCB_APHASE(...) MPCB_PHASE(__VA_ARGS__)
#define MPCB_APHASE_NS(...) MPCB_PHASE_NS(__VA_ARGS__)
#else
#define MPCB_ADRV_PARAMS(SH) ACtx<SH>& c, ATS<SH>& ts0, ATS<SH>& ts1, const int ptid
#define MPCB_ADRV_ARGS c, ts0, ts1, ptid
#if MPCB_ADMM_WIDE
#define MPCB_APHASE_NS(...)                                                   \
    {                                                                         \
        { const int tid = ptid; ATS<SH>& ts = ts0; (void)ts; (void)tid; __VA_ARGS__; }   \
    }
#else
#define MPCB_APHASE_NS(...)                                                   \
    {                                                                         \
        { const int tid = ptid; ATS<SH>& ts = ts0; (void)ts; (void)tid; __VA_ARGS__; }   \
        { const int tid = ptid + ANT; ATS<SH>& ts = ts1; (void)ts; (void)tid; __VA_ARGS__; } \
    }
#endif
#define MPCB_APHASE(...)             \
    {                                \
        MPCB_APHASE_NS(__VA_ARGS__)  \
        __syncthreads();             \
    }
#endif

// Prefetch of thread-private constant operands (tensor memory) across the barrier that ends the previous phase: the device issues
// the load before the barrier and hands the buffer to the consumer; the lock-step host emulation (one thread at a time, a phase
// at a time) declares the buffer inside the consuming phase and loads it there (same values: the strips are constant during a QP).
// Which phases prefetch (device): bit 0 = first W overflow chunk (phase B), bit 4 = the inequality-Jacobian column of phase A (a_fetch).
// Measured and NOT done: C^-1 (phase C), the F row (D), the row / box constants (E) -- no gain, the
// tensor-memory latency is not what bounds those intervals; the collocation coefficients of phase A and part of the inequality row
// of phase E -- their 14..24 extra live registers cost more in spills than the shorter request burst gains (33.7 k -> 32.7 k / 33.0 k).
#ifndef MPCB_PRE_MASK
#define MPCB_PRE_MASK 17
#endif
constexpr bool PRE_B = MPCB_PRE_MASK & 1, PRE_A = MPCB_PRE_MASK & 16;
#ifdef MPCB_EMUL
#define MPCB_PRE_ON false
#define MPCB_PRE_DECL(type, name) type name;
#define MPCB_PRE(...)
#else
#define MPCB_PRE_ON true
#define MPCB_PRE_DECL(type, name) type name;
#define MPCB_PRE(...) __VA_ARGS__
#endif

struct d2 { double a, b; };
MPCB_DEV d2 ld2(const double* p) {          // 16-byte load (p is 16-byte aligned by construction of the layouts)
#ifdef MPCB_EMUL
    d2 r; r.a = p[0]; r.b = p[1]; return r;
#else
    const double2 v = *reinterpret_cast<const double2*>(p);
    d2 r; r.a = v.x; r.b = v.y; return r;
#endif
}
// variable role of a thread: padded-node-order index i (and node / component), false if the thread owns no variable
template <class SH>
MPCB_DEV bool vrole(int tid, int& i, int& k, int& comp) {
    typedef AL<SH> L;
    if (tid == L::PTID) { i = L::IP; k = -1; comp = 0; return true; }
    if (tid >= L::IP) return false;
    i = tid;
    return L::decode(tid, k, comp);
}

// ---- sparse products with the constraint Jacobian A ------------------------------------------------------------------
// (A x)_r ; x in padded node order.  Equality row r = 14 k + c (collocation at node k of segment s = k / 3):
//     sum_l D[j,l] x_{3s+l,c} - ts pbar f_{k,c} + A[r,p] x_p        (robot_ocp.hpp:64-70 for f)
// inequality row r = 252 + 8 k + rr: the hand-filled 8 x 22 Jacobian of node k (robot_ocp.hpp:129-160).
template <class SH>
MPCB_DEV double row_times(const ACtx<SH>& c, int r, const double* x) {
    typedef AL<SH> L;
    if (r < SH::NEQ) {
        const int k = r / 14, cc = r - 14 * k, s = k / 3, j = k - 3 * s, base = L::SEG * s + cc;
        const double x0 = x[base], x1 = x[base + 21], x2 = x[base + 42], x3 = x[base + (s < 5 ? L::SEG : 63)];
        const double xf = x[base + 21 * j + 7], xp = x[L::IP];
        double a0 = c.D(j, 0) * x0, a1 = c.D(j, 1) * x1, a2 = c.mtp() * xf;
        a0 += c.D(j, 2) * x2; a1 += c.D(j, 3) * x3; a2 += c.pcol[r] * xp;
        return (a0 + a1) + a2;
    }
    const int q = r - SH::NEQ, k = q >> 3, rr = q & 7;
    const double* G = c.Jg + k * L::JGS + rr * L::JGR;
    const double* xs = x + (L::vidx(k, 0) & ~1);
    double a0 = 0, a1 = 0, a2 = 0, a3 = c.pcv[r] * x[L::IP];
    MPCB_UNROLL
    for (int cc = 0; cc < 20; cc += 4) {
        const d2 g = ld2(G + cc), g2 = ld2(G + cc + 2), v = ld2(xs + cc), v2 = ld2(xs + cc + 2);
        a0 += g.a * v.a; a1 += g.b * v.b; a2 += g2.a * v2.a; a3 += g2.b * v2.b;
    }
    {   // slots 20, 21 (the 21 coefficients sit in slots shift .. shift + 20, shift <= 1); 22 .. 25 are pads
        const d2 g = ld2(G + 20), v = ld2(xs + 20);
        a0 += g.a * v.a; a1 += g.b * v.b;
    }
    return (a0 + a1) + (a2 + a3);
}
// (A^T y)_i for variable (k, comp), i != p ; y in row order
template <class SH>
MPCB_DEV double col_times(const ACtx<SH>& c, int k, int comp, const double* y) {
    typedef AL<SH> L;
    double g0 = 0, g1 = 0, g2_ = 0, g3 = 0;
    if (comp < 14) {
        const int s1 = k / 3, l1 = k - 3 * s1;
        if (k < SH::NN - 1) {
            const double* yy = y + 14 * (3 * s1) + comp;
            g0 += c.D(0, l1) * yy[0]; g1 += c.D(1, l1) * yy[14]; g2_ += c.D(2, l1) * yy[28];
        }
        if (l1 == 0 && k > 0) {
            const double* yy = y + 14 * (3 * (s1 - 1)) + comp;
            g0 += c.D(0, 3) * yy[0]; g1 += c.D(1, 3) * yy[14]; g2_ += c.D(2, 3) * yy[28];
        }
    }
    if (comp >= 7 && k < SH::NN - 1) g3 += c.mtp() * y[14 * k + comp - 7];
    const double* G = c.Gt + k * L::GTS + comp * L::GTC;
    const double* yi = y + SH::NEQ + 8 * k;
    MPCB_UNROLL
    for (int rr = 0; rr < 8; rr += 4) {
        const d2 g = ld2(G + rr), g2 = ld2(G + rr + 2), v = ld2(yi + rr), v2 = ld2(yi + rr + 2);
        g0 += g.a * v.a; g1 += g.b * v.b; g2_ += g2.a * v2.a; g3 += g2.b * v2.b;
    }
    return (g0 + g1) + (g2_ + g3);
}
// A[r, p]: equality rows from pcol, inequality rows from the last Jacobian column
template <class SH>
MPCB_DEV double pcoef(const ACtx<SH>& c, int r) {
    if (r < SH::NEQ) return c.pcol[r];
    return c.pcv[r];
}
// Dot product of two vectors (length 416, zero padded) by the 32 lanes of warp `w`; returned on lane 0 (other lanes: junk).
// Device: 13 strided products per lane + a 5-step shuffle butterfly.  The host emulation (lanes run one after the other)
// parks the lane sums in scratch32 and lets the last lane add them; it returns the total on lane 31 instead (`last`).
MPCB_DEV double tree32(const double* p);
template <class SH>
MPCB_DEV bool warp_vdot(int tid, int w, const double* a, const double* b, double* scratch32, double& out) {
    if ((tid >> 5) != w) return false;
    const int lane = tid & 31;
    const double* p = a + lane;
    const double* q = b + lane;
    double a0 = p[0] * q[0], a1 = p[32] * q[32], a2 = p[64] * q[64], a3 = p[96] * q[96];
    a0 += p[128] * q[128]; a1 += p[160] * q[160]; a2 += p[192] * q[192]; a3 += p[224] * q[224];
    a0 += p[256] * q[256]; a1 += p[288] * q[288]; a2 += p[320] * q[320]; a3 += p[352] * q[352];
    a0 += p[384] * q[384];
    double s = (a0 + a1) + (a2 + a3);
#ifdef MPCB_EMUL
    scratch32[lane] = s;
    if (lane != 31) return false;
    out = tree32(scratch32);
    return true;
#else
    (void)scratch32;
    MPCB_UNROLL
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    out = s;
    return lane == 0;
#endif
}
// sum of 32 values, pairwise tree
MPCB_DEV double tree32(const double* p) {
    double s[16];
    MPCB_UNROLL
    for (int i = 0; i < 16; i++) s[i] = p[2 * i] + p[2 * i + 1];
    MPCB_UNROLL
    for (int n = 8; n >= 1; n >>= 1) {
        MPCB_UNROLL
        for (int i = 0; i < n; i++) s[i] = s[2 * i] + s[2 * i + 1];
    }
    return s[0];
}

// ---- load ------------------------------------------------------------------------------------------------------------
template <class SH>
MPCB_DEV void a_load(ACtx<SH>& c, ATS<SH>& ts, int tid) {
    typedef AL<SH> L;
    const double* g = c.g;
    MPCB_UNROLL
    for (int e = 0; e < L::WREG; e++) ts.wreg[e] = g[L::G_WREG + e * NT + tid];
    // the thread's strip: global (slot-major, coalesced over tid) -> tensor memory, 8 slots per store
    const int nslot = L::tm_size(tid >> 5);
    for (int q = 0; q < nslot; q += 8) {
        double v[8];
        MPCB_UNROLL
        for (int j = 0; j < 8; j++) v[j] = g[L::G_TM + (q + j) * NT + tid];
#ifdef MPCB_EMUL
        for (int j = 0; j < 8; j++) ts.tm[q + j] = v[j];
#else
        tm_st8(ts.tmaddr, q, v);
#endif
    }
    ts.rhoB = g[L::G_TM + (L::TM_CST + 6) * NT + tid];
    for (int e = 2 * tid; e < L::IMG; e += 2 * NT) { c.sm[e] = g[L::G_IMG + e]; c.sm[e + 1] = g[L::G_IMG + e + 1]; }
    if (tid < L::NV) { c.rhs[tid] = 0; c.xt[tid] = 0; c.w[tid] = 0; }
    ts.zA = ts.yA = ts.x = ts.zB = ts.yB = ts.t = ts.acc = ts.acc2 = 0;     // boxADMM starts cold (warm_start = false)
    if (tid < 72) c.rB[tid] = 0;
    if (tid < L::SBI_T) {      // vector indices of the thread's S_BI candidates (static; absent candidates point at slot 0 with a zero coefficient)
        for (int e = 0; e < L::SBI_E; e++) {
            int seg, k, comp;
            const bool ok = (tid >> 2) < SH::NB - 1 && sbi_cand<SH>(tid >> 2, (tid & 3) + 4 * e, seg, k, comp);
            c.sbi[L::SBI_T * e + tid] = ok ? L::vidx(k, comp) : 0;
        }
    }
    if (tid < 64) c.psum[tid] = 0;
    if (tid < UM_COUNT) c.umax[tid] = 0;
    if (tid == 0) { c.isc[IS_QP_DONE] = 0; c.isc[IS_QP_STATUS] = 2; c.isc[IS_QP_ITERS] = 0; }
#ifndef MPCB_EMUL
    tm_wait_st();
#endif
}

// 32 strided partial sums of a dot product of two vectors of length 416 by warp `w` (no shuffle butterfly: ~230 cycles of pure latency
// at the end of an interval); one lane adds them up in a later interval (b_sbi)
template <class SH>
MPCB_DEV void p_partial(int tid, int w, const double* a, const double* b, double* out32) {
    if ((tid >> 5) != w) return;
    const double* p = a + (tid & 31);
    const double* q = b + (tid & 31);
    double a0 = p[0] * q[0], a1 = p[32] * q[32], a2 = p[64] * q[64], a3 = p[96] * q[96];
    a0 += p[128] * q[128]; a1 += p[160] * q[160]; a2 += p[192] * q[192]; a3 += p[224] * q[224];
    a0 += p[256] * q[256]; a1 += p[288] * q[288]; a2 += p[320] * q[320]; a3 += p[352] * q[352];
    a0 += p[384] * q[384];
    out32[tid & 31] = (a0 + a1) + (a2 + a3);
}
// ---- phase A: rhs = sigma x - h + rho_b z_b - y_b + A^T w  (the p entry of A^T w is added in phase C, see b_sbi) ----
// Straight-line (branch-free) so that the compiler can interleave the loads of the two logical threads a physical thread
// runs: conditions become zero coefficients (dt[20] holds 0.0) and clamped addresses, the store is predicated.
// the 8 inequality-Jacobian coefficients of the thread's variable (its column of the node's 8 x 22 block, transposed copy): constant
// during a QP.  On the device they are read BEFORE the barrier that precedes phase A (end of phase E / of the residual test), so that
// the burst of shared-memory requests behind that barrier is shorter (a_fetch is unconditional: every thread has a valid address).
struct AGt { d2 ga, gb, gc, gd; };
template <class SH>
MPCB_DEV void a_fetch(const ACtx<SH>& c, int tid, AGt& q) {
    typedef AL<SH> L;
    const int ii = tid < L::IP ? tid : 0;
    const int s = (ii / L::SEG) > 5 ? 5 : (ii / L::SEG), rem = ii - L::SEG * s;
    const int j = rem / 21, comp = rem - 21 * j, k = 3 * s + j;
    const double* G = c.Gt + k * L::GTS + comp * L::GTC;
    q.ga = ld2(G); q.gb = ld2(G + 2); q.gc = ld2(G + 4); q.gd = ld2(G + 6);
}
template <class SH, bool PRE>
MPCB_DEV void a_rhs(ACtx<SH>& c, ATS<SH>& ts, int tid, AGt& q) {
    typedef AL<SH> L;
    if (tid >= L::PTID + 32) return;                  // warps 14, 15 own no variable (warp-uniform)
    const bool isp = tid == L::PTID;
    const int ii = tid < L::IP ? tid : 0;
    const int s = (ii / L::SEG) > 5 ? 5 : (ii / L::SEG), rem = ii - L::SEG * s;
    const int j = rem / 21, comp = rem - 21 * j, k = 3 * s + j;        // pads of segments 0..4 decode to (k = 3 s + 3, comp < 3): valid addresses
    const bool valid = (tid < L::IP) && !(s < 5 && rem >= 63);
    const double* y = c.w;
    const int s1 = k / 3, l1 = k - 3 * s1;
    const bool st = comp < 14;
    const bool c1 = st && k < SH::NN - 1, c2 = st && l1 == 0 && k > 0, c3 = comp >= 7 && k < SH::NN - 1;
    const int cc = st ? comp : 0, Z = 20;
    const double* yy1 = y + 42 * (c1 ? s1 : 0) + cc;
    const double* yy2 = y + 42 * (c2 ? s1 - 1 : 0) + cc;
    const double* y3 = y + (c3 ? 14 * k + comp - 7 : 0);
    const double* yi = y + SH::NEQ + 8 * k;
    if constexpr (!PRE) a_fetch<SH>(c, tid, q);
#if MPCB_DCONST
    const double d10 = c1 ? c.cheb->D[0][l1] : 0.0, d11 = c1 ? c.cheb->D[1][l1] : 0.0, d12 = c1 ? c.cheb->D[2][l1] : 0.0;
    const double d20 = c2 ? c.cheb->D[0][3] : 0.0, d21 = c2 ? c.cheb->D[1][3] : 0.0, d22 = c2 ? c.cheb->D[2][3] : 0.0;
    const double mt = c.dt[c3 ? 16 : Z];
#else
    const double d10 = c.dt[c1 ? l1 : Z], d11 = c.dt[c1 ? 4 + l1 : Z], d12 = c.dt[c1 ? 8 + l1 : Z];
    const double d20 = c.dt[c2 ? 3 : Z], d21 = c.dt[c2 ? 7 : Z], d22 = c.dt[c2 ? 11 : Z];
    const double mt = c.dt[c3 ? 16 : Z];
#endif
    const d2 ga = q.ga, gb = q.gb, gc = q.gc, gd = q.gd;
    const d2 va = ld2(yi), vb = ld2(yi + 2), vc = ld2(yi + 4), vd = ld2(yi + 6);
    double g0 = 0, g1 = 0, g2 = 0, g3 = 0;
    g0 += d10 * yy1[0]; g1 += d11 * yy1[14]; g2 += d12 * yy1[28];
    g0 += d20 * yy2[0]; g1 += d21 * yy2[14]; g2 += d22 * yy2[28];
    g3 += mt * y3[0];
    g0 += ga.a * va.a; g1 += ga.b * va.b; g2 += gb.a * vb.a; g3 += gb.b * vb.b;
    g0 += gc.a * vc.a; g1 += gc.b * vc.b; g2 += gd.a * vd.a; g3 += gd.b * vd.b;
    const double g = isp ? 0.0 : (g0 + g1) + (g2 + g3);              // p: the gather arrives through sc[0] in phase B2
    const double h = isp ? 1.0 : 0.0;                                  // gradient of the Mayer cost p (robot_ocp.hpp:206)
    const double r = c.st->qp_sigma * ts.x - h + ts.rhoB * ts.zB - ts.yB + g;
    if (valid || isp) c.rhs[isp ? L::IP : tid] = r;
}
// ---- phase B: t = W rhs_I from the register-resident rows ------------------------------------------------------------
// transpose-reduce of R partial sums over the 8 lanes of a group: lane q ends with the full sum of local row
// (R = 8: 4 b2 + 2 b1 + b0 ; R = 4: 2 b2 + b1, on both lanes of a pair) where b2 b1 b0 are the bits of q
#ifndef MPCB_EMUL
template <int R>
__device__ __forceinline__ double wreduce(const double* v, int q) {
    const bool h4 = q & 4, h2 = q & 2, h1 = q & 1;
    if (R == 8) {
        double u[4], w[2];
        MPCB_UNROLL
        for (int i = 0; i < 4; i++) { const double keep = h4 ? v[4 + i] : v[i], send = h4 ? v[i] : v[4 + i]; u[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4); }
        MPCB_UNROLL
        for (int i = 0; i < 2; i++) { const double keep = h2 ? u[2 + i] : u[i], send = h2 ? u[i] : u[2 + i]; w[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2); }
        const double keep = h1 ? w[1] : w[0], send = h1 ? w[0] : w[1];
        return keep + __shfl_xor_sync(0xffffffffu, send, 1);
    } else {
        double u[2];
        MPCB_UNROLL
        for (int i = 0; i < 2; i++) { const double keep = h4 ? v[2 + i] : v[i], send = h4 ? v[i] : v[2 + i]; u[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4); }
        const double keep = h2 ? u[1] : u[0], send = h2 ? u[0] : u[1];
        const double w = keep + __shfl_xor_sync(0xffffffffu, send, 2);
        return w + __shfl_xor_sync(0xffffffffu, w, 1);
    }
}
#endif
// R x STEPS block of W times the vector: coefficient slot e = step * R + r comes from the registers (e < 32) or from the TMEM
// overflow buffers (ov(e - 32))
// multiply-adds of the coefficient slots [E0, E1) of an R x STEPS block: slot e = step * R + r multiplies x[step] into acc[r];
// coef(e) delivers the coefficient (registers or a TMEM chunk)
template <int R, int STEPS, int E0, int E1, class COEF>
MPCB_DEV void w_slots(const double* x, const COEF& coef, double* acc) {
    MPCB_UNROLL
    for (int e = E0; e < (E1 < R * STEPS ? E1 : R * STEPS); e++) acc[e % R] += coef(e) * x[e / R];
}
// MPCB_PROFILE builds: time stamp `slot` (7..14) of the thread's warp inside a phase function, iteration 100 of CTA 0
#if defined(MPCB_PROFILE) && !defined(MPCB_EMUL)
#define MPCB_ISTAMP(slot) if (c.dbg_iter == 100 && blockIdx.x == 0 && c.prof && (tid & 31) == 0) c.prof[64 + 16 * (slot) + (tid >> 5)] = clock64();
#else
#define MPCB_ISTAMP(slot)
#endif
// one thread's R x STEPS block of W times the vector.  The first TMEM overflow chunk (16 slots) arrives in `pre`: on the device it
// is loaded BEFORE the barrier that ends phase A (w_prefetch: thread-private constant data, so the barrier wait hides the tensor-memory
// latency); the host emulation loads it here.  Order: chunk 0 | second chunk requested | register coefficients (cover its latency) |
// second chunk.
template <class SH, int KIND, int R, int STEPS, bool PRE>
MPCB_DEV void w_block(ACtx<SH>& c, ATS<SH>& ts, int tid, const double* b, double* acc, TmBuf<16>& ch0) {
    typedef AL<SH> L;
    constexpr int W = L::WREG, NCH = L::wchunks(KIND);
    static_assert(NCH >= 1 && NCH <= 2, "one or two overflow chunks");
    if constexpr (!PRE) tm_ld(ts.tma(), L::tm_wov(tid), ch0);
    double x[STEPS];
    MPCB_UNROLL
    for (int t = 0; t < STEPS; t++) x[t] = b[8 * t];
    tm_wait(ch0);
    MPCB_ISTAMP(7)
    w_slots<R, STEPS, W, W + 16>(x, [&](int e) { return ch0.get(e - W); }, acc);
    MPCB_ISTAMP(8)
    if constexpr (NCH > 1) {
        TmBuf<16> ch1;
        tm_ld_after(ts.tma(), L::tm_wov(tid) + 16, ch1, x);
        w_slots<R, STEPS, 0, W>(x, [&](int e) { return ts.wreg[e]; }, acc);
        MPCB_ISTAMP(9)
        tm_wait(ch1);
        w_slots<R, STEPS, W + 16, W + 32>(x, [&](int e) { return ch1.get(e - W - 16); }, acc);
    } else {
        w_slots<R, STEPS, 0, W>(x, [&](int e) { return ts.wreg[e]; }, acc);
        MPCB_ISTAMP(9)
    }
    MPCB_ISTAMP(10)
}
// device: request the first overflow chunk of the thread's W block (every kind has one, at the same strip offset)
template <class SH>
MPCB_DEV void w_prefetch(ATS<SH>& ts, int tid, TmBuf<16>& ch0) {
    tm_ld(ts.tma(), AL<SH>::tm_wov(tid), ch0);
}
template <class SH, bool PRE>
MPCB_DEV void b_w(ACtx<SH>& c, ATS<SH>& ts, int tid, TmBuf<16>& ch0) {
    typedef AL<SH> L;
    const WRole r = wrole<SH>(tid);
    const double* b = c.rhs + L::ibase(r.seg) + r.q;
    double v[8];
    MPCB_UNROLL
    for (int i = 0; i < 8; i++) v[i] = 0;
    if (r.kind == 0) w_block<SH, 0, 4, 8, PRE>(c, ts, tid, b, v, ch0);
    else if (r.kind == 1) w_block<SH, 1, 7, 7, PRE>(c, ts, tid, b, v, ch0);
    else w_block<SH, 2, 4, 9, PRE>(c, ts, tid, b, v, ch0);
#ifdef MPCB_EMUL
    for (int i = 0; i < 8; i++) ts.pv[i] = v[i];
#else
    ts.acc = r.kind == 1 ? wreduce<8>(v, r.q) : wreduce<4>(v, r.q);
    MPCB_ISTAMP(11)
#endif
}
// ---- phase B': separator right-hand side r_B = rhs_B - S_BI t (4 lanes per junction variable), r_B[p] -------------------------
template <class SH>
MPCB_DEV void b_sbi(ACtx<SH>& c, ATS<SH>& ts, int tid) {
    typedef AL<SH> L;
    double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    if (tid < L::SBI_T) {
        // the unit's 7 coefficients S[b, cand] (zero for absent candidates / spare lanes) and the vector entries they multiply
        const int* ix = c.sbi + tid;
        const double* k = c.sbc + tid;
        double tv[L::SBI_E], kv[L::SBI_E];
        MPCB_UNROLL
        for (int e = 0; e < L::SBI_E; e++) { tv[e] = c.xt[ix[L::SBI_T * e]]; kv[e] = k[L::SBI_T * e]; }
        a0 += kv[0] * tv[0]; a1 += kv[1] * tv[1]; a2 += kv[2] * tv[2]; a3 += kv[3] * tv[3];
        a0 += kv[4] * tv[4]; a1 += kv[5] * tv[5]; a2 += kv[6] * tv[6];
    }
    ts.acc = (a0 + a1) + (a2 + a3);
    // the two dot products of r_B[p] = rhs_p + (A^T w)_p - (F^T rhs_I)_p arrive as 32 partial sums each (phase A: p_partial by warp 14,
    // which owns no variable; phase B: by warp 3, whose W blocks are the shortest); two threads of warps that idle in this interval add
    // them up, phase C combines them while it loads r_B.  (Both dot products as shuffle butterflies of idle warps in this interval
    // were measured: 32.5 k -> 31.9 k OCP/s, the butterfly warps become the last ones of the interval.)
    if (tid == 10 * 32) c.sc[0] = tree32(c.psum);
    if (tid == 11 * 32) c.sc[1] = tree32(c.psum + 32);
}
// ---- phase C: x~_B = C^-1 r_B, 4 lanes per row ------------------------------------------------------------------------
template <class SH>
MPCB_DEV void c_separator(ACtx<SH>& c, ATS<SH>& ts, int tid) {
    typedef AL<SH> L;
    double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    if (tid < 128 || tid >= 352) {                      // warp-uniform: tcgen05.ld is warp-collective
        // unit cu: row cu / 4 of C^-1, columns 4 jj + cu % 4 (jj = 0..17, zero beyond NB) from the thread's TMEM strip
        // (the four spare lanes of warp 15 hold zeros)
        TmBuf<8> c0, c1;
        TmBuf<2> c2;
        const int so = L::tm_c(tid);
        tm_ld(ts.tma(), so, c0); tm_ld(ts.tma(), so + 8, c1); tm_ld(ts.tma(), so + 16, c2);
        const double* rb = c.rB + (tid & 3);
        double rv[18];                                   // read while the tensor-memory loads are in flight
        MPCB_UNROLL
        for (int jj = 0; jj < 17; jj++) rv[jj] = rb[4 * jj];
        {   // columns 68..71: r_B[p] (column 70) is formed here from its three pieces; 71 is a pad
            const double rp = (c.rhs[L::IP] + c.sc[0]) - c.sc[1];
            const int l = tid & 3;
            rv[17] = l == 2 ? rp : (l < 2 ? rb[68] : 0.0);
        }
        tm_wait(c0, c1, c2);
        MPCB_UNROLL
        for (int jj = 0; jj < 8; jj += 4) {
            a0 += c0.get(jj) * rv[jj]; a1 += c0.get(jj + 1) * rv[jj + 1];
            a2 += c0.get(jj + 2) * rv[jj + 2]; a3 += c0.get(jj + 3) * rv[jj + 3];
        }
        MPCB_UNROLL
        for (int jj = 0; jj < 8; jj += 4) {
            a0 += c1.get(jj) * rv[8 + jj]; a1 += c1.get(jj + 1) * rv[8 + jj + 1];
            a2 += c1.get(jj + 2) * rv[8 + jj + 2]; a3 += c1.get(jj + 3) * rv[8 + jj + 3];
        }
        a0 += c2.get(0) * rv[16]; a1 += c2.get(1) * rv[17];
    }
    ts.acc = (a0 + a1) + (a2 + a3);
}
// ---- phase D: x~_I = t - F x~_B ----------------------------------------------------------------------------------------
template <class SH>
MPCB_DEV void d_interior(ACtx<SH>& c, ATS<SH>& ts, int tid) {
    typedef AL<SH> L;
    const WRole r = wrole<SH>(tid);
    const int s = r.seg, so = L::tm_frow(tid);
    double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    if (r.paired) {
        // 15 boundary columns: 14 junction states + p, split over the two lanes of the row (8 | 7); strip: F[row, 8 half + j]
        TmBuf<8> fh;
        tm_ld(ts.tma(), so, fh);
        const double* xj = c.xt + (s == 0 ? L::SEG : 5 * L::SEG) + 8 * r.half;
        const d2 v0 = ld2(xj), v1 = ld2(xj + 2), v2 = ld2(xj + 4);
        const double xp = c.xt[L::IP];
        tm_wait(fh);
        a0 += fh.get(0) * v0.a; a1 += fh.get(1) * v0.b; a2 += fh.get(2) * v1.a; a3 += fh.get(3) * v1.b;
        a0 += fh.get(4) * v2.a; a1 += fh.get(5) * v2.b;
        if (r.half == 0) { const d2 v3 = ld2(xj + 6); a2 += fh.get(6) * v3.a; a3 += fh.get(7) * v3.b; }
        else a2 += fh.get(6) * xp;
    } else {
        // strip: the 29 boundary columns of the row (left junction 0..13, right junction 14..27, p)
        TmBuf<8> f0, f1, f2;
        TmBuf<4> f3;
        TmBuf<1> f4;
        tm_ld(ts.tma(), so, f0); tm_ld(ts.tma(), so + 8, f1);
        const double* x1 = c.xt + L::SEG * s;
        const double* x2 = c.xt + L::SEG * (s + 1);
        tm_wait(f0, f1);
        tm_ld(ts.tma(), so + 16, f2); tm_ld(ts.tma(), so + 24, f3); tm_ld(ts.tma(), so + 28, f4);
        MPCB_UNROLL
        for (int cc = 0; cc < 8; cc += 4) {
            const d2 v = ld2(x1 + cc), v2 = ld2(x1 + cc + 2);
            a0 += f0.get(cc) * v.a; a1 += f0.get(cc + 1) * v.b; a2 += f0.get(cc + 2) * v2.a; a3 += f0.get(cc + 3) * v2.b;
        }
        { const d2 v = ld2(x1 + 8), v2 = ld2(x1 + 10); a0 += f1.get(0) * v.a; a1 += f1.get(1) * v.b; a2 += f1.get(2) * v2.a; a3 += f1.get(3) * v2.b; }
        { const d2 v = ld2(x1 + 12); a0 += f1.get(4) * v.a; a1 += f1.get(5) * v.b; }
        { const d2 v = ld2(x2); a0 += f1.get(6) * v.a; a1 += f1.get(7) * v.b; }
        tm_wait(f2, f3, f4);
        { const d2 v = ld2(x2 + 2); a2 += f2.get(0) * v.a; a3 += f2.get(1) * v.b; }
        { const d2 v = ld2(x2 + 4), v2 = ld2(x2 + 6); a0 += f2.get(2) * v.a; a1 += f2.get(3) * v.b; a2 += f2.get(4) * v2.a; a3 += f2.get(5) * v2.b; }
        { const d2 v = ld2(x2 + 8); a0 += f2.get(6) * v.a; a1 += f2.get(7) * v.b; }
        { const d2 v = ld2(x2 + 10), v2 = ld2(x2 + 12); a2 += f3.get(0) * v.a; a3 += f3.get(1) * v.b; a2 += f3.get(2) * v2.a; a3 += f3.get(3) * v2.b; }
        a0 += f4.get(0) * c.xt[L::IP];
    }
    ts.acc = r.valid ? (a0 + a1) + (a2 + a3) : 0.0;
}
// ---- phase E: rows and variables update --------------------------------------------------------------------
// y/rho is evaluated as y * (1/rho) with the reciprocal taken once per QP (an FP64 division costs ~130 cycles of latency).
#ifndef MPCB_PW
#define MPCB_PW 3          // the warp that forms the partial sums of (F^T rhs_I)_p in phase B
#endif
#ifndef MPCB_KLATE
#define MPCB_KLATE 1          // the row / box constants of phase E are requested behind the row product (+0.6 %: their 16 registers stay free for loads in flight)
#endif
#ifndef MPCB_EFLAT
#define MPCB_EFLAT 1
#endif
template <class SH>
MPCB_DEV void e_update(ACtx<SH>& c, ATS<SH>& ts, int tid) {
    typedef AL<SH> L;
    const double al = c.st->qp_alpha;
    TmBuf<8> k;                                  // l, u, rho, 1/rho | lb, ub, rho_b, 1/rho_b
#if !MPCB_KLATE
    tm_ld(ts.tma(), L::TM_CST, k);
#endif
    double ax = 0, xt = 0;
    const int row = L::rowof(tid);
    if (row >= 0) ax = row_times(c, row, c.xt);
#if MPCB_KLATE
    tm_ld(ts.tma(), L::TM_CST, k);               // behind the row product: its 16 registers are free for loads in flight until here
#endif
    int i, kk, comp;
    const bool var = vrole<SH>(tid, i, kk, comp);
#if MPCB_EFLAT
    // straight-line: the row and the variable update are two independent dependent chains of ~10 FP64 operations each; as two guarded
    // blocks they run one after the other, as one block they interleave.  Threads without a row / a variable update state nobody reads
    // (every reader is guarded by the same role test), with finite operands (their state and strip constants are zero).
    xt = c.xt[var ? i : 0];
    tm_wait(k);
    const double om = 1 - al;
    const double zhA = al * ax + om * ts.zA, zhB = al * xt + om * ts.zB;
    const double znA = dmin(dmax(zhA + ts.yA * k.get(3), k.get(0)), k.get(1));
    const double znB = dmin(dmax(zhB + ts.yB * k.get(7), k.get(4)), k.get(5));
    ts.x = al * xt + om * ts.x;
    ts.yA += k.get(2) * (zhA - znA);
    ts.yB += k.get(6) * (zhB - znB);
    ts.zA = znA; ts.zB = znB;
    if (row >= 0) c.w[row] = k.get(2) * znA - ts.yA;
#else
    if (var) xt = c.xt[i];
    tm_wait(k);
    if (row >= 0) {
        const double zh = al * ax + (1 - al) * ts.zA;
        const double zn = dmin(dmax(zh + ts.yA * k.get(3), k.get(0)), k.get(1));
        ts.yA += k.get(2) * (zh - zn);
        ts.zA = zn;
        c.w[row] = k.get(2) * ts.zA - ts.yA;
    }
    if (var) {
        ts.x = al * xt + (1 - al) * ts.x;
        const double zh = al * xt + (1 - al) * ts.zB;
        const double zn = dmin(dmax(zh + ts.yB * k.get(7), k.get(4)), k.get(5));
        ts.yB += k.get(6) * (zh - zn);
        ts.zB = zn;
    }
#endif
}
// ---- residual test (every check_termination iterations) ----------------------------------------------------------------
template <class SH>
MPCB_DEV void chk_a(ACtx<SH>& c, ATS<SH>& ts, int tid) {
    typedef AL<SH> L;
    const int row = L::rowof(tid);
    if (row >= 0) c.rhs[row] = ts.yA;           // the rhs slot is free until the next phase A
    if (tid >= SH::NA && tid < 416) c.rhs[tid] = 0;
    int i, k, comp;
    if (vrole<SH>(tid, i, k, comp)) c.xt[i] = ts.x;
}
template <class SH>
MPCB_DEV void chk_b(ACtx<SH>& c, ATS<SH>& ts, int tid) {
    typedef AL<SH> L;
    double rp = 0, nax = 0, nz = 0, nx = 0, rd = 0, nhx = 0, naty = 0;
    const int row = L::rowof(tid);
    if (row >= 0) {
        const double ax = row_times(c, row, c.xt);
        rp = fabs(ax - ts.zA); nax = fabs(ax); nz = fabs(ts.zA);
    }
    int i, k, comp;
    if (vrole<SH>(tid, i, k, comp)) {
        nx = fabs(ts.x); nz = dmax(nz, fabs(ts.zB)); rp = dmax(rp, fabs(ts.x - ts.zB));
        if (k >= 0) {
            const double aty = col_times(c, k, comp, c.rhs) + ts.yB;
            const double hx = c.hd[i] * ts.x + c.ha[i] * c.xt[L::IP];
            naty = fabs(aty); nhx = fabs(hx); rd = fabs(hx + aty);
        }
    }
    { double s; if (warp_vdot<SH>(tid, 13, c.pcv, c.rhs, c.psum, s)) c.sc[2] = s; }   // (A^T y)_p
    if (tid >= NT - 32) {
        double s = 0;
        for (int j = tid - (NT - 32); j < L::IP; j += 32) s += c.ha[j] * c.xt[j];     // H[p,.] x ; pads hold zeros
        c.psum[32 + tid - (NT - 32)] = s;
    }
    // the test only uses max(|Ax|, |x|, |z|) and max(|Hx|, |A^T y|): four warp reductions instead of seven (the slots UM_NZ, UM_NX
    // and UM_NATY stay zero)
    umax_accum4(c.umax + UM_RP, rp, c.umax + UM_NAX, dmax(dmax(nax, nz), nx), c.umax + UM_RD, rd, c.umax + UM_NHX, dmax(nhx, naty));
}
template <class SH>
MPCB_DEV void chk_c(ACtx<SH>& c, ATS<SH>& ts, int tid, int iter) {
    typedef AL<SH> L;
    if (tid != L::PTID) return;
    const double s1 = c.sc[2], s2 = tree32(c.psum + 32);
    const double aty = s1 + ts.yB, hx = c.hd[L::IP] * ts.x + s2;
    const double rd = dmax(umax_get(c.umax + UM_RD), fabs(hx + 1.0 + aty));
    const double nhx = dmax(umax_get(c.umax + UM_NHX), fabs(hx)), naty = dmax(umax_get(c.umax + UM_NATY), fabs(aty));
    const double rp = umax_get(c.umax + UM_RP);
    const double np_ = dmax(dmax(umax_get(c.umax + UM_NAX), umax_get(c.umax + UM_NX)), umax_get(c.umax + UM_NZ));
    const double nd_ = dmax(dmax(nhx, naty), 1.0);      // |h|_inf = 1
    const double ep = c.st->qp_eps_abs + c.st->qp_eps_rel * np_, ed = c.st->qp_eps_abs + c.st->qp_eps_rel * nd_;
    if (rp <= ep && rd <= ed) { c.isc[IS_QP_DONE] = 1; c.isc[IS_QP_STATUS] = 1; c.isc[IS_QP_ITERS] = iter; }
    for (int j = 0; j < UM_COUNT; j++) c.umax[j] = 0;
}
// ---- result: step and multipliers back to global memory in kernel order (what the SQP stage expects) -----------------
template <class SH>
MPCB_DEV void a_finish(ACtx<SH>& c, ATS<SH>& ts, int tid, int max_iter) {
    typedef AL<SH> L;
    double* res = c.g + L::G_RES;
    if (L::rowof(tid) >= 0) res[L::R_LAM + L::rowof(tid)] = ts.yA;
    int i, k, comp;
    if (vrole<SH>(tid, i, k, comp)) {
        const int io = (k < 0) ? SH::IP : SH::kvar(k, comp);
        res[io] = ts.x;
        res[L::R_LAM + SH::NA + io] = ts.yB;
    }
    if (tid == 0) {
        res[L::R_INFO] = (double)c.isc[IS_QP_STATUS];
        res[L::R_INFO + 1] = (double)(c.isc[IS_QP_DONE] ? c.isc[IS_QP_ITERS] : max_iter);
    }
}

// Per-warp time stamps of one iteration (MPCB_PROFILE builds): slot s, warp w -> prof[64 + 16 s + w] = clock at the END of
// the warp's work in phase s (before the barrier); slot 15 holds the common origin (after the barrier that starts phase A).
#if defined(MPCB_PROFILE) && !defined(MPCB_EMUL)
#define MPCB_WSTAMP(slot)                                                                                        \
    if (iter == 100 && blockIdx.x == 0 && c.prof && (tid & 31) == 0) c.prof[64 + 16 * (slot) + (tid >> 5)] = clock64();
#else
#define MPCB_WSTAMP(slot)
#endif

// host emulation of wreduce (same association of the sums as the shuffle network of the device)
#ifdef MPCB_EMUL
#define MPCB_WREDUCE_EMUL()                                                                                                     \
    {                                                                                                                           \
        const WRole r_ = wrole<SH>(tid);                                                                                        \
        const int q_ = r_.q, b2_ = (q_ >> 2) & 1, b1_ = (q_ >> 1) & 1, b0_ = q_ & 1;                                            \
        const int rl_ = r_.kind == 1 ? 4 * b2_ + 2 * b1_ + b0_ : 2 * b2_ + b1_;                                                 \
        auto P_ = [&](int x) { return tsv[(tid & ~7) | x].pv[rl_]; };                                                           \
        auto S1_ = [&](int x) { return P_(x) + P_(x ^ 4); };                                                                    \
        auto S2_ = [&](int x) { return S1_(x) + S1_(x ^ 2); };                                                                  \
        ts.acc = S2_(q_) + S2_(q_ ^ 1);                                                                                         \
    }
#else
#define MPCB_WREDUCE_EMUL()
#endif

// One QP: load the block, iterate, store the result.
template <class SH>
MPCB_DEV void admm_run(MPCB_ADRV_PARAMS(SH)) {
    typedef AL<SH> L;
#ifndef MPCB_EMUL
    const int tid = ptid;            // for the profiling hooks outside the phases
    (void)tid;
#endif
    MPCB_PROF_DECL
    MPCB_APHASE(a_load(c, ts, tid));
    MPCB_PROF(0)
    MPCB_PRE_DECL(AGt, agt)
    MPCB_APHASE_NS({ MPCB_PRE(if constexpr (PRE_A) a_fetch<SH>(c, tid, agt)); });
    const int max_iter = c.st->qp_max_iter, chk = c.st->qp_check_termination;
    int until_chk = chk > 0 ? chk : -1;          // iterations left until the next residual test (a countdown: no integer division per iteration)
    for (int iter = 1; iter <= max_iter; iter++) {
#if defined(MPCB_PROFILE) && !defined(MPCB_EMUL)
        c.dbg_iter = iter;
#endif
        MPCB_WSTAMP(15)
        MPCB_APHASE_NS((a_rhs<SH, MPCB_PRE_ON && PRE_A>(c, ts, tid, agt)));
        MPCB_PRE_DECL(TmBuf<16>, wch0)
        MPCB_APHASE({ p_partial<SH>(tid, 14, c.pcv, c.w, c.psum); MPCB_PRE(if constexpr (PRE_B) w_prefetch(ts, tid, wch0)); MPCB_WSTAMP(0) });
        MPCB_PROF(1)
        // t = W rhs_I, published in the x~ slots of the interior variables (phase D overwrites them with x~_I = t - F x~_B)
        MPCB_APHASE_NS((b_w<SH, MPCB_PRE_ON && PRE_B>(c, ts, tid, wch0)));
        MPCB_APHASE({
            MPCB_WREDUCE_EMUL()
            const WRole r = wrole<SH>(tid);
            ts.t = ts.acc;
            if (r.valid && r.half == 0) c.xt[r.out] = ts.t;
            p_partial<SH>(tid, MPCB_PW, c.fpc, c.rhs, c.psum + 32);
            MPCB_WSTAMP(1)
        });
        MPCB_PROF(2)
        MPCB_APHASE_NS({ b_sbi(c, ts, tid); MPCB_ISTAMP(12) });
        MPCB_APHASE_NS({ ts.acc2 = ts.acc + MPCB_PEER(acc, 1); MPCB_ISTAMP(13) });
        MPCB_APHASE({
            const double s4 = ts.acc2 + MPCB_PEER(acc2, 2);
            if (tid < 4 * (SH::NB - 1) && (tid & 3) == 0) c.rB[tid >> 2] = c.rhs[L::sepidx(tid >> 2)] - s4;
            MPCB_WSTAMP(2)
        });
        MPCB_PROF(3)
        MPCB_APHASE_NS(c_separator(c, ts, tid));
        MPCB_APHASE_NS({ ts.acc2 = ts.acc + MPCB_PEER(acc, 1); });
        MPCB_APHASE({
            const double s4 = ts.acc2 + MPCB_PEER(acc2, 2);
            const int cu = L::cunit(tid);
            if (cu >= 0 && (cu & 3) == 0) c.xt[L::sepidx(cu >> 2)] = s4;
            MPCB_WSTAMP(3)
        });
        MPCB_PROF(4)
        MPCB_APHASE_NS(d_interior(c, ts, tid));
        MPCB_APHASE({
            const double o = MPCB_PEER(acc, 1);
            const WRole r = wrole<SH>(tid);
            const double fx = r.paired ? ts.acc + o : ts.acc;
            if (r.valid && r.half == 0) c.xt[r.out] = ts.t - fx;
            MPCB_WSTAMP(4)
        });
        MPCB_PROF(5)
        MPCB_APHASE({ e_update(c, ts, tid); MPCB_PRE(if constexpr (PRE_A) a_fetch<SH>(c, tid, agt)); MPCB_WSTAMP(5) });
        MPCB_WSTAMP(6)
        MPCB_PROF(6)
        if (--until_chk == 0) {
            until_chk = chk;
            MPCB_APHASE(chk_a(c, ts, tid));
            MPCB_APHASE(chk_b(c, ts, tid));
            MPCB_APHASE(chk_c(c, ts, tid, iter));
            MPCB_PROF(7)
            if (c.isc[IS_QP_DONE]) break;
        }
    }
    MPCB_APHASE(a_finish(c, ts, tid, max_iter));
    MPCB_PROF(8)
}

}  // namespace mpcb
