// TEST-ONLY lockstep emulation of the OCP kernel on the host (see mpcb_ctx.h): the very same phase code the CUDA
// kernel runs, executed thread id by thread id with a loop per phase.  Built by tests/emul/Makefile into
// tests/emul/libmpcb_emul.so and loaded only by the CPU test-suite to check the kernel logic against the oracle
// without a GPU.  Never part of libmpcb.so.
#define MPCB_EMUL 1
#include <vector>
#include <cstring>
#include "mpcb_sqp.h"
#include "mpcb_host_util.h"

using namespace mpcb;
typedef Shape<3, 6> SH36;

template <class SH>
struct Emu {
    std::vector<double> smem, z, lam;
    std::vector<TS<SH>> tsv;
    Ctx<SH> c;
    ChebConst cheb;
    double limtab[40];
    Emu(const mpcb_settings* st, const mpcb_limits* lim) : smem(Layout<SH>::TOTAL, 0.0), z(SH::VP, 0.0), lam(SH::NDUAL + 8, 0.0), tsv(NT) {
        mpcb_fill_cheb(SH::PO, SH::NS, &cheb);
        fill_limtab(lim, limtab);
        c.bind(smem.data());
        c.z = z.data(); c.lam = lam.data(); c.st = st; c.cheb = &cheb;
        std::memset(tsv.data(), 0, sizeof(TS<SH>) * NT);
        for (int tid = 0; tid < NT; tid++) setup_roles(c, tsv[tid], tid);
    }
};

extern "C" {

int emul_smem_bytes() { return (int)Layout<SH36>::BYTES; }

int emul_solve(const mpcb_settings* st, const mpcb_limits* lim, int n, const double* x0, const double* xT, const double* gx,
               const double* gu, const double* gT, const double* glam, double* sx, double* su, double* sT, double* slam,
               mpcb_stats* stats) {
    typedef SH36 SH;
    const char* why;
    if (!settings_supported(st, &why)) return MPCB_ERR_UNSUPPORTED;
    Emu<SH> e(st, lim);
    for (int i = 0; i < n; i++) {
        Problem pb;
        pb.limtab = e.limtab;
        pb.x0 = x0 + 14 * i; pb.xT = xT + 14 * i; pb.gx = gx + (size_t)14 * SH::NN * i; pb.gu = gu + (size_t)7 * SH::NN * i;
        pb.glam = glam ? glam + (size_t)SH::NDUAL * i : nullptr; pb.gT = gT[i];
        pb.sx = sx + (size_t)14 * SH::NN * i; pb.su = su + (size_t)7 * SH::NN * i; pb.sT = sT + i;
        pb.slam = slam ? slam + (size_t)SH::NDUAL * i : nullptr;
        pb.stats = stats ? stats + i : nullptr;
        solve_ocp<SH>(e.c, e.tsv.data(), pb);
    }
    return 0;
}

// Debug entry: linearise at (z, lam) given in polympc order, factorise, and apply S^-1 to `rhs` (polympc order).
int emul_kkt_apply(const mpcb_settings* st, const mpcb_limits* lim, const double* x0, const double* xT, const double* zext,
                   const double* lamext, const double* rhs, double* out, double* rhoA, double* rhoB) {
    typedef SH36 SH;
    Emu<SH> e(st, lim);
    Ctx<SH>& c = e.c;
    TS<SH>* tsv = e.tsv.data();
    std::vector<double> gx(zext, zext + 14 * SH::NN), gu(zext + 14 * SH::NN, zext + 21 * SH::NN), dummy(SH::NVAR + SH::NDUAL);
    Problem pb;
    pb.limtab = e.limtab; pb.x0 = x0; pb.xT = xT; pb.gx = gx.data(); pb.gu = gu.data(); pb.gT = zext[SH::NVAR - 1]; pb.glam = lamext;
    pb.sx = pb.su = pb.sT = dummy.data(); pb.slam = nullptr; pb.stats = nullptr;
    MPCB_PHASE(sqp_load_a(c, tid, pb));
    MPCB_PHASE(sqp_load_b(c, ts, tid, pb));
    linearise<SH>(c, tsv);
    factorise<SH>(c, tsv);
    for (int tid = 0; tid < SH::NVAR; tid++) {
        c.rhs[tid] = rhs[ext_index<SH>(tsv[tid].vk, tsv[tid].vcomp)];
        rhoB[ext_index<SH>(tsv[tid].vk, tsv[tid].vcomp)] = tsv[tid].rhoB;
    }
    for (int r = 0; r < SH::NA; r++) rhoA[r] = tsv[r].rho;
    for (int j = 0; j < 32; j++) c.psum[j] = 0;
    MPCB_PHASE(admm_solve1(c, ts, tid));
    MPCB_PHASE(admm_solve2(c, tid));
    MPCB_PHASE(admm_solve3(c, ts, tid));
    for (int tid = 0; tid < SH::NVAR; tid++) out[ext_index<SH>(tsv[tid].vk, tsv[tid].vcomp)] = c.xt[tid];
    return 0;
}

}  // extern "C"
