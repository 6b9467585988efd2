// TEST-ONLY lockstep emulation of the OCP kernel on the host (see mpcb_ctx.h): the very same phase code the CUDA
// kernel runs, executed thread id by thread id with a loop per phase.  Built by tests/emul/Makefile into
// tests/emul/libmpcb_emul.so and loaded only by the CPU test-suite to check the kernel logic against the oracle
// without a GPU.  Never part of libmpcb.so.
#define MPCB_EMUL 1
#include <vector>
#include <cstring>
#include <type_traits>
#include "mpcb_sqp.h"
#include "mpcb_host_util.h"

using namespace mpcb;
typedef Shape<3, 6> SH36;

template <class SH>
struct Emu {
    std::vector<double> smem, gmat, z, lam, qb, asmem, jg, gj, gval;
    std::vector<TS<SH>> tsv;
    static constexpr bool FAST = SH::PO == 3 && SH::NS == 6;       // the (3,6) ADMM kernel exists for this shape
    typedef typename std::conditional<FAST, SH, SH36>::type ASH;
    std::vector<ATS<ASH>> atsv;
    Ctx<SH> c;
    ACtx<ASH> ac;
    ChebConst cheb;
    double limtab[40];
    int flags[2];
    Emu(const mpcb_settings* st, const mpcb_limits* lim)
        : smem(Layout<SH>::TOTAL, 0.0), gmat(Layout<SH>::CTA_BLOCK, 0.0), z(SH::VP, 0.0), lam(SH::NDUAL + 8, 0.0), qb(QBL<SH>::QPBLOCK, 0.0), jg(SH::NN * 176, 0.0),
          gj(SH::NINEQ, 0.0), gval(16 * SH::NINEQ, 0.0), tsv(SH::NT) {
        mpcb_fill_cheb(SH::PO, SH::NS, &cheb);
        fill_limtab(lim, limtab);
        c.bind(smem.data(), gmat.data());
        c.z = z.data(); c.lam = lam.data(); c.st = st; c.cheb = &cheb;
        c.gJg = jg.data(); c.gGj = gj.data(); c.gval = gval.data(); c.prof = nullptr;
        if constexpr (FAST) {
            asmem.assign(AL<SH>::TOTAL, 0.0);
            atsv.resize(NT);
            ac.bind(asmem.data());
            ac.st = st; ac.cheb = &cheb; ac.g = qb.data();
            std::memset(atsv.data(), 0, sizeof(ATS<SH>) * NT);
        }
        std::memset(tsv.data(), 0, sizeof(TS<SH>) * SH::NT);
        for (int tid = 0; tid < SH::NT; tid++) setup_roles(c, tsv[tid], tid);
        flags[0] = flags[1] = 0;
    }
    // host stand-ins of node_jac_kernel / node_val_kernel (one "thread" per node / per (trial, node))
    void node_jac_all() {
        rbd::DerivScratch S;
        for (int k = 0; k < SH::NN; k++) node_jac<SH>(z.data(), k, S, jg.data() + k * 176, gj.data() + k * 8);
    }
    void node_val_all() {
        const int ntrial = c.st->line_search_max_iter + 1;
        for (int t = 0; t < ntrial; t++)
            for (int k = 0; k < SH::NN; k++)
                node_val<SH>(z.data(), qb.data() + QBL<SH>::G_RES, t, c.st->sqp_tau, k, gval.data() + (t * SH::NN + k) * 8);
    }
};

template <class SH>
static int emul_solve_shape(const mpcb_settings* st, const mpcb_limits* lim, int n, const double* x0, const double* xT, const double* gx,
                            const double* gu, const double* gT, const double* glam, double* sx, double* su, double* sT, double* slam,
                            mpcb_stats* stats) {
    const bool generic = settings_generic_path(st) != 0;
    Emu<SH> e(st, lim);
    for (int i = 0; i < n; i++) {
        Problem pb;
        pb.limtab = e.limtab;
        pb.x0 = x0 + 14 * i; pb.xT = xT + 14 * i; pb.gx = gx + (size_t)14 * SH::NN * i; pb.gu = gu + (size_t)7 * SH::NN * i;
        pb.glam = glam ? glam + (size_t)SH::NDUAL * i : nullptr; pb.gT = gT[i];
        pb.sx = sx + (size_t)14 * SH::NN * i; pb.su = su + (size_t)7 * SH::NN * i; pb.sT = sT + i;
        pb.slam = slam ? slam + (size_t)SH::NDUAL * i : nullptr;
        pb.stats = stats ? stats + i : nullptr;
        // the same kernel sequence mpcb_solve_batch launches
        stage_load<SH>(e.c, e.tsv.data(), pb, e.flags);
        for (int it = 1; it <= st->sqp_max_iter; it++) {
            if (e.flags[0]) break;
            e.node_jac_all();
            if (generic) stage_qp_generic<SH>(e.c, e.tsv.data(), pb, e.qb.data());
            else if constexpr (SH::PO == 3 && SH::NS == 6) {
                stage_linfac<SH>(e.c, e.tsv.data(), pb, e.qb.data());
                admm_run<SH>(e.ac, e.atsv.data());
            }
            e.node_val_all();
            stage_lsstep<SH>(e.c, e.tsv.data(), pb, e.qb.data(), e.flags, it);
        }
    }
    return 0;
}

extern "C" {

int emul_smem_bytes() { return (int)Layout<SH36>::BYTES; }
int emul_admm_smem_bytes() { return (int)AL<SH36>::BYTES; }

int emul_solve(const mpcb_settings* st, const mpcb_limits* lim, int n, const double* x0, const double* xT, const double* gx,
               const double* gu, const double* gT, const double* glam, double* sx, double* su, double* sT, double* slam,
               mpcb_stats* stats) {
    const char* why;
    if (!settings_supported(st, &why)) return MPCB_ERR_UNSUPPORTED;
#define EMUL_ARGS st, lim, n, x0, xT, gx, gu, gT, glam, sx, su, sT, slam, stats
    if (st->poly_order == 3 && st->num_seg == 6) return emul_solve_shape<Shape<3, 6>>(EMUL_ARGS);
    if (st->poly_order == 3 && st->num_seg == 12) return emul_solve_shape<Shape<3, 12>>(EMUL_ARGS);
    if (st->poly_order == 5 && st->num_seg == 6) return emul_solve_shape<Shape<5, 6>>(EMUL_ARGS);
    if (st->poly_order == 7 && st->num_seg == 4) return emul_solve_shape<Shape<7, 4>>(EMUL_ARGS);
#undef EMUL_ARGS
    return MPCB_ERR_UNSUPPORTED;
}

// Debug entry: linearise at (z, lam) given in polympc order, factorise, export, and apply S^-1 to `rhs` (polympc order)
// with the ADMM kernel's own solve phases (B, B', C, D).
int emul_kkt_apply(const mpcb_settings* st, const mpcb_limits* lim, const double* x0, const double* xT, const double* zext,
                   const double* lamext, const double* rhs, double* out, double* rhoA, double* rhoB) {
    typedef SH36 SH;
    typedef AL<SH> L;
    Emu<SH> e(st, lim);
    std::vector<double> gx(zext, zext + 14 * SH::NN), gu(zext + 14 * SH::NN, zext + 21 * SH::NN), dummy(SH::NVAR + SH::NDUAL);
    Problem pb;
    pb.limtab = e.limtab; pb.x0 = x0; pb.xT = xT; pb.gx = gx.data(); pb.gu = gu.data(); pb.gT = zext[SH::NVAR - 1]; pb.glam = lamext;
    pb.sx = pb.su = pb.sT = dummy.data(); pb.slam = nullptr; pb.stats = nullptr;
    stage_load<SH>(e.c, e.tsv.data(), pb, e.flags);
    e.node_jac_all();
    stage_linfac<SH>(e.c, e.tsv.data(), pb, e.qb.data());
    for (int tid = 0; tid < SH::NVAR; tid++) rhoB[ext_index<SH>(e.tsv[tid].vk, e.tsv[tid].vcomp)] = e.tsv[tid].rhoB;
    for (int r = 0; r < SH::NA; r++) rhoA[r] = e.tsv[r].rho;
    ACtx<SH>& c = e.ac;
    ATS<SH>* tsv = e.atsv.data();
    MPCB_PHASE(a_load(c, ts, tid));
    for (int i = 0; i <= L::IP; i++) {
        int k, comp;
        if (L::decode(i, k, comp)) c.rhs[i] = rhs[ext_index<SH>(k, comp)];
    }
    c.sc[0] = 0;        // no A^T w gather in this test: rhs is given directly
    MPCB_PHASE_NS({ TmBuf<16> wch0; b_w<SH, false>(c, ts, tid, wch0); });
    MPCB_PHASE({ MPCB_WREDUCE_EMUL() const WRole r = wrole<SH>(tid); ts.t = ts.acc; if (r.valid && r.half == 0) c.xt[r.out] = ts.t; p_partial<SH>(tid, MPCB_PW, c.fpc, c.rhs, c.psum + 32); });
    for (int j = 0; j < 32; j++) c.psum[j] = 0;        // no A^T w gather in this test
    MPCB_PHASE_NS(b_sbi(c, ts, tid));
    MPCB_PHASE_NS({ ts.acc2 = ts.acc + MPCB_PEER(acc, 1); });
    MPCB_PHASE({ const double s4 = ts.acc2 + MPCB_PEER(acc2, 2); if (tid < 4 * (SH::NB - 1) && (tid & 3) == 0) c.rB[tid >> 2] = c.rhs[L::sepidx(tid >> 2)] - s4; });
    MPCB_PHASE_NS(c_separator(c, ts, tid));
    MPCB_PHASE_NS({ ts.acc2 = ts.acc + MPCB_PEER(acc, 1); });
    MPCB_PHASE({ const double s4 = ts.acc2 + MPCB_PEER(acc2, 2); const int cu = L::cunit(tid); if (cu >= 0 && (cu & 3) == 0) c.xt[L::sepidx(cu >> 2)] = s4; });
    MPCB_PHASE_NS(d_interior(c, ts, tid));
    MPCB_PHASE({
        const double o = MPCB_PEER(acc, 1);
        const WRole r = wrole<SH>(tid);
        const double fx = r.paired ? ts.acc + o : ts.acc;
        if (r.valid && r.half == 0) c.xt[r.out] = ts.t - fx;
    });
    for (int i = 0; i <= L::IP; i++) {
        int k, comp;
        if (L::decode(i, k, comp)) out[ext_index<SH>(k, comp)] = c.xt[i];
    }
    return 0;
}

}  // extern "C"
