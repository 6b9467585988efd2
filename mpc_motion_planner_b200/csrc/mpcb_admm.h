// The box-ADMM iteration kernel body (polympc boxADMM::solve restated in SURVEY.md A.5) on top of the factorisation
// of mpcb_qp.h.  This is the hot loop of the whole path: <= 700 iterations per QP, 2 QPs per OCP
// (/root/reference/mpc_solver/motionPlanner.cpp:15-20,184).
//
//   rhs   = sigma x - h + rho_b z_b - y_b + A^T (rho_A z_A - y_A)                                   phase A
//   x~    = S^-1 rhs  via  t = W rhs_I, part = F^T rhs_I | x~_B = C^-1 (rhs_B - part) | x~_I = t - F x~_B    phases B, B', C, D
//   z~_A  = A x~ ; x = a x~ + (1-a) x ; z = clip(a z~ + (1-a) z + y/rho) ; y += rho (a z~ + (1-a) z_old - z)   phase E
// Every `check_termination` iterations the OSQP residual test runs (chk_*).
//
// Data placement (one CTA of NT = 512 threads per OCP, one CTA per SM):
//   * every thread owns (part of) one row of an interior inverse W_s and keeps 32 of its coefficients in REGISTERS for the
//     whole QP (16 K doubles of the 18.5 K of all W_s); the rest of W, F, C^-1 and the inequality Jacobian sit in shared
//     memory in conflict-free layouts; shared memory delivers 16 doubles/clk against 64 FMA/clk of the FP64 pipe, so
//     everything that can live in registers does;
//   * row r's (z_A, y_A, l, u, rho) and variable i's (x, z_b, y_b, lb, ub, rho_b) live in registers of thread r / i;
//   * vectors use the "padded node order": index(k, comp) = 64 s + 21 j + comp with s = min(k/3, 5), j = k - 3 s, p at 404;
//     every segment interior and every junction state is contiguous and starts on an even index (16-byte loads).
// The QP data arrives from the factorisation kernel through a per-OCP block in global memory (layout: struct AL).
#pragma once
#include "mpcb_ctx.h"

namespace mpcb {

template <class SH>
struct AL {
    static_assert(SH::PO == 3 && SH::NS == 6, "the ADMM kernel's thread mapping is specialised for POLY_ORDER=3, NUM_SEG=6");
    static constexpr int NV = 408, IP = 404;           // padded node order
    static constexpr int WREG = 32;                    // W coefficients per thread held in registers
    // ---- shared-memory image (copied verbatim from the global QP block) --------------------------------------------
    static constexpr int I_WOVM = 0;                   // [18][224]  columns 32..49 of the middle-segment rows
    static constexpr int I_WOV5 = I_WOVM + 18 * 224;   // [4][160]   tail columns of the last segment's half rows
    static constexpr int I_F = I_WOV5 + 4 * 160;       // F_s column-major with odd leading dimension
    static constexpr int FSZ = 15 * 63 + 4 * 29 * 49 + 15 * 71 + 2;
    static constexpr int I_C = I_F + FSZ;              // [18][284]  C^-1, 4 lanes per row
    static constexpr int I_JG = I_C + 18 * 284;        // 19 nodes x (8 x 22), node stride 177
    static constexpr int JGS = 177;
    static constexpr int I_HD = I_JG + 19 * JGS + 1;   // regularised Hessian diagonal (padded node order)
    static constexpr int I_HA = I_HD + NV;             // Hessian arrow H[p, .]
    static constexpr int I_PCOL = I_HA + NV;           // A[:, p] of the equality rows (256)
    static constexpr int I_DT = I_PCOL + 256;          // D (16) | -ts*pbar | pad
    static constexpr int IMG = I_DT + 24;
    // ---- working vectors behind the image --------------------------------------------------------------------------
    static constexpr int V_RHS = IMG, V_XT = V_RHS + NV, V_W = V_XT + NV, V_PARTF = V_W + NV, V_RB = V_PARTF + 296;
    static constexpr int V_PSUM = V_RB + 72, V_SC = V_PSUM + 64, V_UMAX = V_SC + 8, V_ISC = V_UMAX + UM_COUNT;
    static constexpr int TOTAL = V_ISC + 8;
    static constexpr size_t BYTES = (size_t)TOTAL * 8;
    // ---- global QP block ---------------------------------------------------------------------------------------------
    static constexpr int G_WREG = 0;                   // [32][NT]
    static constexpr int G_IMG = G_WREG + WREG * NT;
    static constexpr int G_ROW = G_IMG + IMG;          // l | u | rho | pc   (NV each, row order)
    static constexpr int G_VAR = G_ROW + 4 * NV;       // lb | ub | rho_b    (NV each, padded node order)
    static constexpr int G_RES = G_VAR + 3 * NV;       // step p (kernel order, VP) | multipliers (NDUAL+4) | info (4)
    static constexpr int R_LAM = SH::VP, R_INFO = SH::VP + SH::NDUAL + 4;
    static constexpr int QPBLOCK = G_RES + R_INFO + 4;

    MPCB_HD static int fld(int s) { return s == 0 ? 63 : (s == 5 ? 71 : 49); }                  // leading dimension of F_s
    MPCB_HD static int foff(int s) { return s == 0 ? 0 : 15 * 63 + (s - 1) * 29 * 49; }
    MPCB_HD static int vidx(int k, int comp) { const int s = k / 3 > 5 ? 5 : k / 3; return 64 * s + 21 * (k - 3 * s) + comp; }
    MPCB_HD static int sepidx(int b) { return b < 70 ? 64 * (b / 14 + 1) + b % 14 : IP; }
    MPCB_HD static int ibase(int s) { return s == 0 ? 0 : 64 * s + 14; }                        // first interior index of segment s
    // decode a padded-node-order index: false for pads
    MPCB_HD static bool decode(int i, int& k, int& comp) {
        if (i == IP) { k = -1; comp = 0; return true; }
        if (i > IP) return false;
        const int s = (i >> 6) > 5 ? 5 : (i >> 6), rem = i - 64 * s;
        if (s < 5 && rem == 63) return false;
        const int j = rem / 21;
        k = 3 * s + j; comp = rem - 21 * j;
        return true;
    }
};

// Role of a thread in the interior mat-vecs (phases B and D): which row of which W_s, and which columns it covers.
struct WRole {
    bool valid, paired;
    int seg, row, half;
    int c0;          // first column of the register part (32 columns)
    int nov;         // number of (padded) overflow columns read from shared memory, starting at column c0 + 32
    int out;         // padded-node-order index of the row
};
template <class SH>
MPCB_HD WRole wrole(int tid) {
    WRole r;
    if (tid < 128) {
        r.seg = 0; r.row = tid >> 1; r.half = tid & 1; r.valid = r.row < 63; r.paired = true; r.c0 = 32 * r.half; r.nov = 0; r.out = r.row;
    } else if (tid < 352) {
        const int idx = tid - 128;
        r.valid = idx < 196; r.paired = false; r.half = 0;
        r.seg = 1 + idx / 49; r.row = idx % 49; r.c0 = 0; r.nov = 18;
        if (!r.valid) { r.seg = 1; r.row = 0; }
        r.out = 64 * r.seg + 14 + r.row;
    } else {
        const int idx = tid - 352;
        r.seg = 5; r.row = idx >> 1; r.half = idx & 1; r.valid = idx < 140; r.paired = true; r.c0 = 36 * r.half; r.nov = 4;
        if (!r.valid) r.row = 0;
        r.out = 334 + r.row;
    }
    return r;
}
// number of real columns of row-half (seg, half): used by the exporter to zero the padding
template <class SH>
MPCB_HD int wcols_end(const WRole& r) {
    if (r.seg == 0) return r.half ? 63 : 32;
    if (r.seg == 5) return r.half ? 70 : 36;
    return 49;
}

// Role of a thread in part = F^T rhs_I (phase B): one column of one F_s over a chunk of rows.
struct FRole {
    int unit;        // -1: none
    int seg, col, r0, len;
};
template <class SH>
MPCB_HD FRole frole(int tid) {
    FRole f;
    f.unit = -1; f.seg = 0; f.col = 0; f.r0 = 0; f.len = 0;
    int u = -1;
    if (tid < 128) u = tid;
    else if (tid >= 352) u = tid - 352 + 128;
    else if (tid >= 324 && tid < 328) u = 288 + tid - 324;
    if (u < 0) return f;
    f.unit = u;
    if (u < 232) {
        f.seg = 1 + u / 58; const int rem = u % 58, ch = rem / 29;
        f.col = rem % 29; f.r0 = 26 * ch; f.len = ch ? 23 : 26;
    } else if (u < 262) {
        const int rem = u - 232, ch = rem / 15;
        f.seg = 0; f.col = rem % 15; f.r0 = 32 * ch; f.len = ch ? 31 : 32;
    } else {
        const int rem = u - 262, ch = rem / 15;
        f.seg = 5; f.col = rem % 15; f.r0 = 36 * ch; f.len = ch ? 34 : 36;
    }
    return f;
}
// unit index of (seg, col, chunk)
template <class SH>
MPCB_HD int funit(int s, int col, int ch) {
    if (s == 0) return 232 + 15 * ch + col;
    if (s == 5) return 262 + 15 * ch + col;
    return 58 * (s - 1) + 29 * ch + col;
}

template <class SH>
struct ACtx {
    double* sm;
    double *wovm, *wov5, *F, *C, *Jg, *hd, *ha, *pcol, *dt;
    double *rhs, *xt, *w, *partF, *rB, *psum, *sc;
    unsigned long long* umax;
    long long* isc;
    double* g;                     // this OCP's QP block in global memory
    const mpcb_settings* st;
    long long* prof;               // MPCB_PROFILE builds: per-phase cycle counters of CTA 0 (else unused)
    MPCB_HD void bind(double* smem) {
        typedef AL<SH> L;
        sm = smem;
        wovm = sm + L::I_WOVM; wov5 = sm + L::I_WOV5; F = sm + L::I_F; C = sm + L::I_C; Jg = sm + L::I_JG; hd = sm + L::I_HD;
        ha = sm + L::I_HA; pcol = sm + L::I_PCOL; dt = sm + L::I_DT;
        rhs = sm + L::V_RHS; xt = sm + L::V_XT; w = sm + L::V_W; partF = sm + L::V_PARTF; rB = sm + L::V_RB; psum = sm + L::V_PSUM;
        sc = sm + L::V_SC;
        umax = reinterpret_cast<unsigned long long*>(sm + L::V_UMAX);
        isc = reinterpret_cast<long long*>(sm + L::V_ISC);
    }
    MPCB_HD double D(int j, int l) const { return dt[4 * j + l]; }
    MPCB_HD double mtp() const { return dt[16]; }
};

// Per-thread persistent state of the ADMM kernel.
template <class SH>
struct ATS {
    double wreg[AL<SH>::WREG];
    double l, u, rho, zA, yA, pc;          // row role (thread r < NA)
    double x, zB, yB, lb, ub, rhoB;        // variable role (thread i, padded node order)
    double t, acc, acc2;                   // interior row: W rhs_I ; cross-lane exchange slots
};

#ifdef MPCB_EMUL
#define MPCB_ADRV_PARAMS(SH) ACtx<SH>& c, ATS<SH>* tsv
#define MPCB_ADRV_ARGS c, tsv
#else
#define MPCB_ADRV_PARAMS(SH) ACtx<SH>& c, ATS<SH>& ts, const int tid
#define MPCB_ADRV_ARGS c, ts, tid
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

// ---- sparse products with the constraint Jacobian A ------------------------------------------------------------------
// (A x)_r ; x in padded node order.  Equality row r = 14 k + c (collocation at node k of segment s = k / 3):
//     sum_l D[j,l] x_{3s+l,c} - ts pbar f_{k,c} + A[r,p] x_p        (robot_ocp.hpp:64-70 for f)
// inequality row r = 252 + 8 k + rr: the hand-filled 8 x 22 Jacobian of node k (robot_ocp.hpp:129-160).
template <class SH>
MPCB_DEV double row_times(const ACtx<SH>& c, int r, double pc, const double* x) {
    typedef AL<SH> L;
    if (r < SH::NEQ) {
        const int k = r / 14, cc = r - 14 * k, s = k / 3, j = k - 3 * s, base = 64 * s + cc;
        double acc = c.D(j, 0) * x[base];
        acc += c.D(j, 1) * x[base + 21];
        acc += c.D(j, 2) * x[base + 42];
        acc += c.D(j, 3) * x[base + 63 + (s < 5 ? 1 : 0)];
        acc += c.mtp() * x[base + 21 * j + 7];
        acc += pc * x[L::IP];
        return acc;
    }
    const int q = r - SH::NEQ, k = q >> 3, rr = q & 7;
    const double* G = c.Jg + k * L::JGS + rr * 22;
    const double* xs = x + L::vidx(k, 0);
    double acc = 0;
    MPCB_UNROLL
    for (int cc = 0; cc < 21; cc++) acc += G[cc] * xs[cc];
    acc += G[21] * x[L::IP];
    return acc;
}
// (A^T y)_i for variable (k, comp), i != p ; y in row order
template <class SH>
MPCB_DEV double col_times(const ACtx<SH>& c, int k, int comp, const double* y) {
    typedef AL<SH> L;
    double g = 0;
    if (comp < 14) {
        const int s1 = k / 3, l1 = k - 3 * s1;
        if (k < SH::NN - 1) {
            const double* yy = y + 14 * (3 * s1) + comp;
            g += c.D(0, l1) * yy[0]; g += c.D(1, l1) * yy[14]; g += c.D(2, l1) * yy[28];
        }
        if (l1 == 0 && k > 0) {
            const double* yy = y + 14 * (3 * (s1 - 1)) + comp;
            g += c.D(0, 3) * yy[0]; g += c.D(1, 3) * yy[14]; g += c.D(2, 3) * yy[28];
        }
    }
    if (comp >= 7 && k < SH::NN - 1) g += c.mtp() * y[14 * k + comp - 7];
    const double* G = c.Jg + k * L::JGS + comp;
    const double* yi = y + SH::NEQ + 8 * k;
    MPCB_UNROLL
    for (int rr = 0; rr < 8; rr++) g += G[rr * 22] * yi[rr];
    return g;
}
// partial sums of (A^T y)_p = sum_r A[r,p] y_r  by the last 32 threads
template <class SH>
MPCB_DEV void pcol_partial(const ACtx<SH>& c, int tid, const double* y, double* out) {
    if (tid < NT - 32) return;
    const int j = tid - (NT - 32);
    double s = 0;
    for (int r = j; r < SH::NEQ; r += 32) s += c.pcol[r] * y[r];
    for (int r = j; r < SH::NINEQ; r += 32) s += c.Jg[(r >> 3) * AL<SH>::JGS + (r & 7) * 22 + 21] * y[SH::NEQ + r];
    out[j] = s;
}

// ---- load ------------------------------------------------------------------------------------------------------------
template <class SH>
MPCB_DEV void a_load(ACtx<SH>& c, ATS<SH>& ts, int tid) {
    typedef AL<SH> L;
    const double* g = c.g;
    MPCB_UNROLL
    for (int e = 0; e < L::WREG; e++) ts.wreg[e] = g[L::G_WREG + e * NT + tid];
    for (int e = 2 * tid; e < L::IMG; e += 2 * NT) { c.sm[e] = g[L::G_IMG + e]; c.sm[e + 1] = g[L::G_IMG + e + 1]; }
    ts.l = ts.u = ts.rho = ts.pc = 0; ts.lb = ts.ub = ts.rhoB = 0;
    if (tid < L::NV) {
        ts.l = g[L::G_ROW + tid]; ts.u = g[L::G_ROW + L::NV + tid]; ts.rho = g[L::G_ROW + 2 * L::NV + tid]; ts.pc = g[L::G_ROW + 3 * L::NV + tid];
        ts.lb = g[L::G_VAR + tid]; ts.ub = g[L::G_VAR + L::NV + tid]; ts.rhoB = g[L::G_VAR + 2 * L::NV + tid];
        c.rhs[tid] = 0; c.xt[tid] = 0; c.w[tid] = 0;
    }
    ts.zA = ts.yA = ts.x = ts.zB = ts.yB = ts.t = ts.acc = ts.acc2 = 0;     // boxADMM starts cold (warm_start = false)
    if (tid < 72) c.rB[tid] = 0;
    if (tid < 296) c.partF[tid] = 0;
    if (tid < 64) c.psum[tid] = 0;
    if (tid < UM_COUNT) c.umax[tid] = 0;
    if (tid == 0) { c.isc[IS_QP_DONE] = 0; c.isc[IS_QP_STATUS] = 2; c.isc[IS_QP_ITERS] = 0; }
}

// ---- phase A: rhs ------------------------------------------------------------------------------------------------------
template <class SH>
MPCB_DEV void a_rhs(ACtx<SH>& c, ATS<SH>& ts, int tid) {
    typedef AL<SH> L;
    int k, comp;
    if (tid <= L::IP && L::decode(tid, k, comp)) {
        const double g = (k >= 0) ? col_times(c, k, comp, c.w) : 0.0;
        const double h = (k < 0) ? 1.0 : 0.0;              // gradient of the Mayer cost p (robot_ocp.hpp:206)
        c.rhs[tid] = c.st->qp_sigma * ts.x - h + ts.rhoB * ts.zB - ts.yB + g;
    }
    pcol_partial(c, tid, c.w, c.psum);
}

// ---- phase B: t = W rhs_I (register-resident rows) and part = F^T rhs_I ----------------------------------------------
template <class SH>
MPCB_DEV void b_interior(ACtx<SH>& c, ATS<SH>& ts, int tid) {
    typedef AL<SH> L;
    const WRole r = wrole<SH>(tid);
    {
        const double* b = c.rhs + L::ibase(r.seg) + r.c0;
        double a0 = 0, a1 = 0;
        MPCB_UNROLL
        for (int e = 0; e < L::WREG; e += 2) {
            const d2 v = ld2(b + e);
            a0 += ts.wreg[e] * v.a; a1 += ts.wreg[e + 1] * v.b;
        }
        if (r.nov == 18) {
            const double* wo = c.wovm + (tid - 128);
            MPCB_UNROLL
            for (int e = 0; e < 18; e += 2) {
                const d2 v = ld2(b + 32 + e);
                a0 += wo[e * 224] * v.a; a1 += wo[(e + 1) * 224] * v.b;
            }
        } else if (r.nov == 4) {
            const double* wo = c.wov5 + (tid - 352);
            MPCB_UNROLL
            for (int e = 0; e < 4; e += 2) {
                const d2 v = ld2(b + 32 + e);
                a0 += wo[e * 160] * v.a; a1 += wo[(e + 1) * 160] * v.b;
            }
        }
        ts.acc = a0 + a1;
    }
    const FRole f = frole<SH>(tid);
    if (f.unit >= 0) {
        const double* Fc = c.F + L::foff(f.seg) + f.col * L::fld(f.seg) + f.r0;
        const double* b = c.rhs + L::ibase(f.seg) + f.r0;
        double a0 = 0, a1 = 0;
        int i = 0;
        for (; i + 1 < f.len; i += 2) {
            const d2 v = ld2(b + i);
            a0 += Fc[i] * v.a; a1 += Fc[i + 1] * v.b;
        }
        if (i < f.len) a0 += Fc[i] * b[i];
        c.partF[f.unit] = a0 + a1;
    }
    if (tid == NT - 1) {
        double s = 0;
        for (int j = 0; j < 32; j++) s += c.psum[j];
        c.rhs[L::IP] += s;
    }
}
// ---- phase B': separator right-hand side r_B = rhs_B - sum_s (F_s^T rhs_I) -------------------------------------------
template <class SH>
MPCB_DEV void b_seprhs(ACtx<SH>& c, int tid) {
    typedef AL<SH> L;
    if (tid >= SH::NB) return;
    double r = c.rhs[L::sepidx(tid)];
    if (tid == SH::NB - 1) {
        for (int s = 0; s < SH::NS; s++) { const int col = SH::nb(s) - 1; r -= c.partF[funit<SH>(s, col, 0)]; r -= c.partF[funit<SH>(s, col, 1)]; }
    } else {
        const int sj = tid / 14 + 1, cc = tid % 14;     // junction x_{3 sj}: right boundary of segment sj-1, left boundary of sj
        const int cprev = (sj - 1 > 0 ? 14 : 0) + cc;
        r -= c.partF[funit<SH>(sj - 1, cprev, 0)]; r -= c.partF[funit<SH>(sj - 1, cprev, 1)];
        r -= c.partF[funit<SH>(sj, cc, 0)]; r -= c.partF[funit<SH>(sj, cc, 1)];
    }
    c.rB[tid] = r;
}
// ---- phase C: x~_B = C^-1 r_B, 4 lanes per row ------------------------------------------------------------------------
template <class SH>
MPCB_DEV void c_separator(ACtx<SH>& c, ATS<SH>& ts, int tid) {
    double a0 = 0, a1 = 0;
    if (tid < 4 * SH::NB) {
        const double* Cr = c.C + tid;
        const double* rb = c.rB + (tid & 3);
        MPCB_UNROLL
        for (int jj = 0; jj < 18; jj += 2) { a0 += Cr[jj * 284] * rb[4 * jj]; a1 += Cr[(jj + 1) * 284] * rb[4 * jj + 4]; }
    }
    ts.acc = a0 + a1;
}
// ---- phase D: x~_I = t - F x~_B ----------------------------------------------------------------------------------------
template <class SH>
MPCB_DEV void d_interior(ACtx<SH>& c, ATS<SH>& ts, int tid) {
    typedef AL<SH> L;
    const WRole r = wrole<SH>(tid);
    const int s = r.seg, ld = L::fld(s);
    const double* Fr = c.F + L::foff(s) + r.row;
    double a0 = 0, a1 = 0;
    if (s == 0 || s == 5) {
        // 15 boundary columns: 14 junction states + p, split over the two lanes of the row (8 | 7)
        const double* xj = c.xt + (s == 0 ? 64 : 320);
        if (r.half == 0) {
            MPCB_UNROLL
            for (int cc = 0; cc < 8; cc += 2) { const d2 v = ld2(xj + cc); a0 += Fr[cc * ld] * v.a; a1 += Fr[(cc + 1) * ld] * v.b; }
        } else {
            MPCB_UNROLL
            for (int cc = 8; cc < 14; cc += 2) { const d2 v = ld2(xj + cc); a0 += Fr[cc * ld] * v.a; a1 += Fr[(cc + 1) * ld] * v.b; }
            a0 += Fr[14 * ld] * c.xt[L::IP];
        }
    } else {
        const double* x1 = c.xt + 64 * s;
        const double* x2 = c.xt + 64 * (s + 1);
        MPCB_UNROLL
        for (int cc = 0; cc < 14; cc += 2) { const d2 v = ld2(x1 + cc); a0 += Fr[cc * ld] * v.a; a1 += Fr[(cc + 1) * ld] * v.b; }
        MPCB_UNROLL
        for (int cc = 0; cc < 14; cc += 2) { const d2 v = ld2(x2 + cc); a0 += Fr[(14 + cc) * ld] * v.a; a1 += Fr[(15 + cc) * ld] * v.b; }
        a0 += Fr[28 * ld] * c.xt[L::IP];
    }
    ts.acc = r.valid ? a0 + a1 : 0.0;
}
// ---- phase E: rows and variables update --------------------------------------------------------------------
template <class SH>
MPCB_DEV void e_update(ACtx<SH>& c, ATS<SH>& ts, int tid) {
    typedef AL<SH> L;
    const double al = c.st->qp_alpha;
    if (tid < SH::NA) {
        const double ax = row_times(c, tid, ts.pc, c.xt);
        const double zh = al * ax + (1 - al) * ts.zA;
        const double zn = dmin(dmax(zh + ts.yA / ts.rho, ts.l), ts.u);
        ts.yA += ts.rho * (zh - zn);
        ts.zA = zn;
        c.w[tid] = ts.rho * ts.zA - ts.yA;
    }
    int k, comp;
    if (tid <= L::IP && L::decode(tid, k, comp)) {
        const double xt = c.xt[tid];
        ts.x = al * xt + (1 - al) * ts.x;
        const double zh = al * xt + (1 - al) * ts.zB;
        const double zn = dmin(dmax(zh + ts.yB / ts.rhoB, ts.lb), ts.ub);
        ts.yB += ts.rhoB * (zh - zn);
        ts.zB = zn;
    }
}
// ---- residual test (every check_termination iterations) ----------------------------------------------------------------
template <class SH>
MPCB_DEV void chk_a(ACtx<SH>& c, ATS<SH>& ts, int tid) {
    typedef AL<SH> L;
    if (tid < SH::NA) c.rhs[tid] = ts.yA;       // the rhs slot is free until the next phase A
    int k, comp;
    if (tid <= L::IP && L::decode(tid, k, comp)) c.xt[tid] = ts.x;
}
template <class SH>
MPCB_DEV void chk_b(ACtx<SH>& c, ATS<SH>& ts, int tid) {
    typedef AL<SH> L;
    double rp = 0, nax = 0, nz = 0, nx = 0, rd = 0, nhx = 0, naty = 0;
    if (tid < SH::NA) {
        const double ax = row_times(c, tid, ts.pc, c.xt);
        rp = fabs(ax - ts.zA); nax = fabs(ax); nz = fabs(ts.zA);
    }
    int k, comp;
    if (tid <= L::IP && L::decode(tid, k, comp)) {
        nx = fabs(ts.x); nz = dmax(nz, fabs(ts.zB)); rp = dmax(rp, fabs(ts.x - ts.zB));
        if (k >= 0) {
            const double aty = col_times(c, k, comp, c.rhs) + ts.yB;
            const double hx = c.hd[tid] * ts.x + c.ha[tid] * c.xt[L::IP];
            naty = fabs(aty); nhx = fabs(hx); rd = fabs(hx + aty);
        }
    }
    pcol_partial(c, tid, c.rhs, c.psum);
    if (tid >= NT - 32) {
        double s = 0;
        for (int i = tid - (NT - 32); i < L::IP; i += 32) s += c.ha[i] * c.xt[i];     // pads hold zeros
        c.psum[32 + tid - (NT - 32)] = s;
    }
    umax_accum(c.umax + UM_RP, rp); umax_accum(c.umax + UM_NAX, nax); umax_accum(c.umax + UM_NZ, nz); umax_accum(c.umax + UM_NX, nx);
    umax_accum(c.umax + UM_RD, rd); umax_accum(c.umax + UM_NHX, nhx); umax_accum(c.umax + UM_NATY, naty);
}
template <class SH>
MPCB_DEV void chk_c(ACtx<SH>& c, ATS<SH>& ts, int tid, int iter) {
    typedef AL<SH> L;
    if (tid != L::IP) return;
    double s1 = 0, s2 = 0;
    for (int j = 0; j < 32; j++) { s1 += c.psum[j]; s2 += c.psum[32 + j]; }
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
    if (tid < SH::NA) res[L::R_LAM + tid] = ts.yA;
    int k, comp;
    if (tid <= L::IP && L::decode(tid, k, comp)) {
        const int i = (k < 0) ? SH::IP : SH::kvar(k, comp);
        res[i] = ts.x;
        res[L::R_LAM + SH::NA + i] = ts.yB;
    }
    if (tid == 0) {
        res[L::R_INFO] = (double)c.isc[IS_QP_STATUS];
        res[L::R_INFO + 1] = (double)(c.isc[IS_QP_DONE] ? c.isc[IS_QP_ITERS] : max_iter);
    }
}

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

// One QP: load the block, iterate, store the result.
template <class SH>
MPCB_DEV void admm_run(MPCB_ADRV_PARAMS(SH)) {
    typedef AL<SH> L;
    MPCB_PROF_DECL
    MPCB_PHASE(a_load(c, ts, tid));
    MPCB_PROF(0)
    const int max_iter = c.st->qp_max_iter, chk = c.st->qp_check_termination;
    for (int iter = 1; iter <= max_iter; iter++) {
        MPCB_PHASE(a_rhs(c, ts, tid));
        MPCB_PROF(1)
        MPCB_PHASE_NS(b_interior(c, ts, tid));
        MPCB_PHASE({
            const double o = MPCB_PEER(acc, 1);
            ts.t = wrole<SH>(tid).paired ? ts.acc + o : ts.acc;
        });
        MPCB_PROF(2)
        MPCB_PHASE(b_seprhs(c, tid));
        MPCB_PROF(3)
        MPCB_PHASE_NS(c_separator(c, ts, tid));
        MPCB_PHASE_NS({ ts.acc2 = ts.acc + MPCB_PEER(acc, 1); });
        MPCB_PHASE({
            const double s4 = ts.acc2 + MPCB_PEER(acc2, 2);
            if (tid < 4 * SH::NB && (tid & 3) == 0) c.xt[L::sepidx(tid >> 2)] = s4;
        });
        MPCB_PROF(4)
        MPCB_PHASE_NS(d_interior(c, ts, tid));
        MPCB_PHASE({
            const double o = MPCB_PEER(acc, 1);
            const WRole r = wrole<SH>(tid);
            const double fx = r.paired ? ts.acc + o : ts.acc;
            if (r.valid && r.half == 0) c.xt[r.out] = ts.t - fx;
        });
        MPCB_PROF(5)
        MPCB_PHASE(e_update(c, ts, tid));
        MPCB_PROF(6)
        if (chk > 0 && iter % chk == 0) {
            MPCB_PHASE(chk_a(c, ts, tid));
            MPCB_PHASE(chk_b(c, ts, tid));
            MPCB_PHASE(chk_c(c, ts, tid, iter));
            MPCB_PROF(7)
            if (c.isc[IS_QP_DONE]) break;
        }
    }
    MPCB_PHASE(a_finish(c, ts, tid, max_iter));
    MPCB_PROF(8)
}

}  // namespace mpcb
