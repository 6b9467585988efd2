// Hand-over between the SQP-stage kernels (linearise + factorise, line search + step; shared-memory context Ctx of
// mpcb_ctx.h) and the ADMM kernel (mpcb_admm.h): the factorised QP is written to a per-OCP block in global memory in
// exactly the layouts the ADMM kernel wants (struct AL), and the QP result comes back the same way.
// The blocks stay in L2/HBM between the kernels: ~340 KB per OCP and SQP iteration, i.e. a few hundred MB per thousand
// problems -- negligible against the ~1400 ADMM iterations each block is then used for.
#pragma once
#include "mpcb_admm.h"
#include "mpcb_qp.h"

namespace mpcb {

template <>
struct QBL<Shape<3, 6>> {
    typedef AL<Shape<3, 6>> L;
    static constexpr int G_RES = L::G_RES, R_LAM = L::R_LAM, R_INFO = L::R_INFO, QPBLOCK = L::QPBLOCK;
};

template <class SH>
MPCB_DEV double sym_packed(const double* P, int i, int j) {
    const int hi = i > j ? i : j, lo = i > j ? j : i;
    return P[hi * (hi + 1) / 2 + lo];
}

// W rows: 32 coefficients per ADMM thread (register part); the overflow columns of the middle / last segments go to the
// thread's TMEM strip.  The factorisation holds -S_II^-1 (packed lower, per segment); the ADMM kernel wants +S_II^-1.
template <class SH>
MPCB_DEV void exp_w(const Ctx<SH>& c, int tid, double* qb) {
    typedef AL<SH> L;
    const WRole r = wrole<SH>(tid);
    const double* Wn = c.W + SH::woff(r.seg);
    int row, col;
    for (int e = 0; e < L::WREG; e++)
        qb[L::G_WREG + e * NT + tid] = wslot<SH>(r, tid, e, row, col) ? -sym_packed<SH>(Wn, row, col) : 0.0;
    double* tm = qb + L::G_TM + tid;
    const int nov = 16 * L::wchunks(r.kind);           // overflow slots behind the register part, whole 16-slot chunks
    for (int e = 0; e < nov; e++)
        tm[(L::tm_wov(tid) + e) * NT] = wslot<SH>(r, tid, L::WREG + e, row, col) ? -sym_packed<SH>(Wn, row, col) : 0.0;
}
// F_s (row-major i*nb + col in the factorisation) -> per-thread pieces in the TMEM strips: the row (or half row) the
// thread multiplies in phase D; C^-1 -> 4 units per row; the inequality Jacobian stays in the shared-memory image
template <class SH>
MPCB_DEV void exp_fc(const Ctx<SH>& c, int tid, double* qb) {
    typedef AL<SH> L;
    double* img = qb + L::G_IMG;
    double* tm = qb + L::G_TM + tid;
    {
        const WRole r = wrole<SH>(tid);
        const double* Fs = c.F + SH::foff(r.seg) + r.row * SH::nb(r.seg);
        const int so = L::tm_frow(tid);
        if (r.paired) {
            for (int j = 0; j < 8; j++) {
                const int col = 8 * r.half + j;
                tm[(so + j) * NT] = (r.valid && col < SH::nb(r.seg)) ? Fs[col] : 0.0;
            }
        } else {
            for (int j = 0; j < 29; j++) tm[(so + j) * NT] = r.valid ? Fs[j] : 0.0;
        }
    }
    if (tid < 128 || tid >= 352) {
        const int cu = L::cunit(tid), so = L::tm_c(tid);
        for (int jj = 0; jj < 18; jj++) {
            const int col = 4 * jj + (tid & 3);
            tm[(so + jj) * NT] = (cu >= 0 && col < SH::NB) ? -sym_packed<SH>(c.C, cu >> 2, col) : 0.0;
        }
    }
    for (int e = tid; e < 19 * L::JGS; e += NT) {        // rows for A x, shifted by the parity of the node's first index
        const int k = e / L::JGS, o = e - k * L::JGS, rr = o / L::JGR, cc = o - rr * L::JGR - (L::vidx(k, 0) & 1);
        img[L::I_JG + e] = (cc >= 0 && cc < 21) ? c.Jg[k * 176 + rr * 22 + cc] : 0.0;
    }
    for (int e = tid; e < 19 * L::GTS; e += NT) {        // transposed for A^T y
        const int k = e / L::GTS, o = e - k * L::GTS, comp = o / L::GTC, rr = o - comp * L::GTC;
        img[L::I_GT + e] = rr < 8 ? c.Jg[k * 176 + rr * 22 + comp] : 0.0;
    }
}
// vectors: Hessian diagonal / arrow, A[:,p], D, row data, box data
template <class SH>
MPCB_DEV void exp_vec(const Ctx<SH>& c, const TS<SH>& ts, int tid, double* qb) {
    typedef AL<SH> L;
    double* img = qb + L::G_IMG;
    double* tm = qb + L::G_TM;
    if (tid < L::NV) {
        int k, comp;
        if (!L::decode(tid, k, comp)) { img[L::I_HD + tid] = 0; img[L::I_HA + tid] = 0; }       // pads
    }
    {   // constants of ADMM threads that own no row / no variable
        int i, k, comp;
        if (L::rowof(tid) < 0) { tm[(L::TM_CST + 0) * NT + tid] = 0; tm[(L::TM_CST + 1) * NT + tid] = 0; tm[(L::TM_CST + 2) * NT + tid] = 1; tm[(L::TM_CST + 3) * NT + tid] = 1; }
        if (!vrole<SH>(tid, i, k, comp)) { tm[(L::TM_CST + 4) * NT + tid] = 0; tm[(L::TM_CST + 5) * NT + tid] = 0; tm[(L::TM_CST + 6) * NT + tid] = 1; tm[(L::TM_CST + 7) * NT + tid] = 1; }
    }
    if (tid < SH::NVAR) {                     // thread tid owns variable tid (kernel order); its ADMM thread is the padded-node-order index
        const int i = (ts.vk < 0) ? L::IP : L::vidx(ts.vk, ts.vcomp);
        const int at = (ts.vk < 0) ? L::PTID : i;
        img[L::I_HD + i] = ts.hd; img[L::I_HA + i] = ts.ha;
        tm[(L::TM_CST + 4) * NT + at] = ts.lb; tm[(L::TM_CST + 5) * NT + at] = ts.ub;
        tm[(L::TM_CST + 6) * NT + at] = ts.rhoB; tm[(L::TM_CST + 7) * NT + at] = 1.0 / ts.rhoB;
    }
    if (tid < SH::NA) {                       // thread tid owns row tid; its ADMM thread is rtid(tid)
        const int at = L::rtid(tid);
        tm[(L::TM_CST + 0) * NT + at] = ts.l; tm[(L::TM_CST + 1) * NT + at] = ts.u;
        tm[(L::TM_CST + 2) * NT + at] = ts.rho; tm[(L::TM_CST + 3) * NT + at] = 1.0 / ts.rho;
    }
    if (tid < 256) img[L::I_PCOL + tid] = (tid < SH::NEQ) ? c.pcol[tid] : 0.0;
    if (tid < 416) {
        // A[r, p] of every row; F_s[row, p] at the padded-node-order index of every interior variable
        img[L::I_PCV + tid] = tid < SH::NEQ ? c.pcol[tid] : (tid < SH::NA ? c.Jg[(tid - SH::NEQ) * 22 + 21] : 0.0);
        int k, comp;
        double fp = 0.0;
        if (tid < L::IP && L::decode(tid, k, comp)) {
            const int s = k / 3 > 5 ? 5 : k / 3, row = tid - L::ibase(s);
            if (s == 0 || row >= 0) fp = c.F[SH::foff(s) + row * SH::nb(s) + (SH::nb(s) - 1)];
        }
        img[L::I_FPC + tid] = fp;
    }
    if (tid < 24) img[L::I_DT + tid] = (tid < 16) ? c.cheb->D[tid >> 2][tid & 3] : (tid == 16 ? c.scal[SC_MTP] : 0.0);
}
// BEFORE the sweeps the F area still holds S_IB (the assembled KKT entries between interior and separator variables): the
// sparse rows S_BI the ADMM kernel uses for part = S_BI t (mpcb_admm.h, sbi_cand) go out here.
template <class SH>
MPCB_DEV void exp_sbi(const Ctx<SH>& c, int tid, double* qb) {
    typedef AL<SH> L;
    double* img = qb + L::G_IMG;
    if (tid < L::SBI_T) {
        const int b = tid >> 2, q = tid & 3;
        for (int t = 0; t < L::SBI_E; t++) {
            double v = 0.0;
            int seg, k, comp;
            if (b < SH::NB - 1 && sbi_cand<SH>(b, q + 4 * t, seg, k, comp)) {
                const int l = SH::aug_index(seg, k, comp), nb = SH::nb(seg);       // interior row of the candidate in its segment
                v = c.F[SH::foff(seg) + l * nb + sbi_col<SH>(b, seg)];
            }
            img[L::I_SBI + L::SBI_T * t + tid] = v;
        }
    }
}
template <class SH>
MPCB_DEV void export_pre(MPCB_DRIVER_PARAMS(SH), double* qb) {
    MPCB_PHASE(exp_sbi(c, tid, qb));
}
template <class SH>
MPCB_DEV void export_qp(MPCB_DRIVER_PARAMS(SH), double* qb) {
    MPCB_PHASE({ exp_w(c, tid, qb); exp_fc(c, tid, qb); exp_vec(c, ts, tid, qb); });
}
// QP result (kernel order) -> the shared-memory slots the SQP stage reads (aliased onto the dead factor)
template <class SH>
MPCB_DEV void import_qp(Ctx<SH>& c, int tid, const double* qb) {
    typedef QBL<SH> L;
    const double* res = qb + L::G_RES;
    if (tid < SH::NVAR) c.qp_p[tid] = res[tid];
    for (int i = tid; i < SH::NDUAL; i += SH::NT) c.qp_lam[i] = res[L::R_LAM + i];
    if (tid == 0) { c.isc[IS_QP_STATUS] = (long long)res[L::R_INFO]; c.isc[IS_QP_ITERS] = (long long)res[L::R_INFO + 1]; }
}

}  // namespace mpcb
