// Hand-over between the SQP-stage kernels (linearise + factorise, line search + step; shared-memory context Ctx of
// mpcb_ctx.h) and the ADMM kernel (mpcb_admm.h): the factorised QP is written to a per-OCP block in global memory in
// exactly the layouts the ADMM kernel wants (struct AL), and the QP result comes back the same way.
// The blocks stay in L2/HBM between the kernels: ~340 KB per OCP and SQP iteration, i.e. a few hundred MB per thousand
// problems -- negligible against the ~1400 ADMM iterations each block is then used for.
#pragma once
#include "mpcb_admm.h"
#include "mpcb_qp.h"

namespace mpcb {

template <class SH>
MPCB_DEV double sym_packed(const double* P, int i, int j) {
    const int hi = i > j ? i : j, lo = i > j ? j : i;
    return P[hi * (hi + 1) / 2 + lo];
}

// W rows: 32 coefficients per ADMM thread (register part) + the overflow columns of the middle / last segments.
// The factorisation holds -S_II^-1 (packed lower, per segment); the ADMM kernel wants +S_II^-1.
template <class SH>
MPCB_DEV void exp_w(const Ctx<SH>& c, int tid, double* qb) {
    typedef AL<SH> L;
    const WRole r = wrole<SH>(tid);
    const double* Wn = c.W + SH::woff(r.seg);
    for (int e = 0; e < L::WREG; e++) {
        const int col = wcol<SH>(r, e);
        qb[L::G_WREG + e * NT + tid] = col >= 0 ? -sym_packed<SH>(Wn, r.row, col) : 0.0;
    }
    double* img = qb + L::G_IMG;
    if (tid >= 128 && tid < 352) {
        for (int e = 0; e < 18; e++) {
            const int col = wcol<SH>(r, 32 + e);
            img[L::I_WOVM + e * 224 + (tid - 128)] = col >= 0 ? -sym_packed<SH>(Wn, r.row, col) : 0.0;
        }
    } else if (tid >= 352) {
        for (int e = 0; e < 4; e++) {
            const int col = wcol<SH>(r, 32 + e);
            img[L::I_WOV5 + e * 160 + (tid - 352)] = col >= 0 ? -sym_packed<SH>(Wn, r.row, col) : 0.0;
        }
    }
}
// F_s (row-major i*nb + col in the factorisation) -> column-major with odd leading dimension; C^-1 -> 4 lanes per row
template <class SH>
MPCB_DEV void exp_fc(const Ctx<SH>& c, int tid, double* qb) {
    typedef AL<SH> L;
    double* img = qb + L::G_IMG;
    for (int e = tid; e < L::FSZ; e += NT) {
        double v = 0.0;
        {
            const int s = e < 945 ? 0 : (e < 945 + 4 * L::FMID ? 1 + (e - 945) / L::FMID : 5);
            const int off = e - L::foff(s), ld = L::fld(s), col = off / ld, i = off - col * ld;
            if (i < SH::m(s) && col < SH::nb(s)) v = c.F[SH::foff(s) + i * SH::nb(s) + col];
        }
        img[L::I_F + e] = v;
    }
    for (int e = tid; e < 18 * 284; e += NT) {
        const int jj = e / 284, t = e - jj * 284, b = t >> 2, col = 4 * jj + (t & 3);
        img[L::I_C + e] = (col < SH::NB) ? -sym_packed<SH>(c.C, b, col) : 0.0;
    }
    for (int e = tid; e < 19 * L::JGS + 1; e += NT) {
        const int k = e / L::JGS, o = e - k * L::JGS;
        img[L::I_JG + e] = (k < SH::NN && o < 176) ? c.Jg[k * 176 + o] : 0.0;
    }
}
// vectors: Hessian diagonal / arrow, A[:,p], D, row data, box data
template <class SH>
MPCB_DEV void exp_vec(const Ctx<SH>& c, const TS<SH>& ts, int tid, double* qb) {
    typedef AL<SH> L;
    double* img = qb + L::G_IMG;
    if (tid < L::NV) {
        int k, comp;
        if (!L::decode(tid, k, comp)) {       // pads
            img[L::I_HD + tid] = 0; img[L::I_HA + tid] = 0;
            qb[L::G_VAR + tid] = 0; qb[L::G_VAR + L::NV + tid] = 0; qb[L::G_VAR + 2 * L::NV + tid] = 0;
        }
        if (tid >= SH::NA) { qb[L::G_ROW + tid] = 0; qb[L::G_ROW + L::NV + tid] = 0; qb[L::G_ROW + 2 * L::NV + tid] = 0; }
        if (tid >= SH::NEQ) qb[L::G_ROW + 3 * L::NV + tid] = 0;
    }
    if (tid < SH::NVAR) {                     // thread tid owns variable tid (kernel order)
        const int i = (ts.vk < 0) ? L::IP : L::vidx(ts.vk, ts.vcomp);
        img[L::I_HD + i] = ts.hd; img[L::I_HA + i] = ts.ha;
        qb[L::G_VAR + i] = ts.lb; qb[L::G_VAR + L::NV + i] = ts.ub; qb[L::G_VAR + 2 * L::NV + i] = ts.rhoB;
    }
    if (tid < SH::NA) {
        qb[L::G_ROW + tid] = ts.l; qb[L::G_ROW + L::NV + tid] = ts.u; qb[L::G_ROW + 2 * L::NV + tid] = ts.rho;
        if (tid < SH::NEQ) qb[L::G_ROW + 3 * L::NV + tid] = ts.pc;
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
template <class SH>
MPCB_DEV void export_qp(MPCB_DRIVER_PARAMS(SH), double* qb) {
    MPCB_PHASE({ exp_w(c, tid, qb); exp_fc(c, tid, qb); exp_vec(c, ts, tid, qb); });
}
// QP result (kernel order) -> the shared-memory slots the SQP stage reads (aliased onto the dead factor)
template <class SH>
MPCB_DEV void import_qp(Ctx<SH>& c, int tid, const double* qb) {
    typedef AL<SH> L;
    const double* res = qb + L::G_RES;
    if (tid < SH::NVAR) c.qp_p[tid] = res[tid];
    for (int i = tid; i < SH::NDUAL; i += NT) c.qp_lam[i] = res[L::R_LAM + i];
    if (tid == 0) { c.isc[IS_QP_STATUS] = (long long)res[L::R_INFO]; c.isc[IS_QP_ITERS] = (long long)res[L::R_INFO + 1]; }
}

}  // namespace mpcb
