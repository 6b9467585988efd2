// Interface between the C-ABI implementation (mpcb_capi.cpp, host C++) and the CUDA translation unit (mpcb_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include "mpcb.h"
#include "mpcb_shape.h"

namespace mpcb {

// Launch-invariant constants, passed as a __grid_constant__ kernel parameter.
struct KernelConst {
    mpcb_settings st;
    ChebConst cheb;
    double limtab[40];      // fill_limtab()
    double jmax[7];         // for the jerk flag of the accounting kernel
};
// Constants of the warm-start kernel (otg_guess_kernel): margins x limits per joint, the bound on T, the node times
struct OtgConst {
    double vl[7], al[7], jl[7];
    double p_ub;
    double tn[64];
    int N;
};
// Device pointers of one batch (problem-major arrays, layouts in mpcb.h).
struct BatchPtrs {
    const double *x0, *xT, *gx, *gu, *gT, *glam;
    double *sx, *su, *sT, *slam;
    mpcb_stats* stats;
};

size_t solve_smem_bytes();
size_t solve_workspace_doubles_per_ocp();
cudaError_t solve_prepare();
// kind 0: load the guess; kind 1: linearise + factorise + export; kind 2: line search / step of SQP iteration it
cudaError_t launch_sqp_stage(const KernelConst& kc, const BatchPtrs& bp, int n, int kind, int it, int grid, int* counter, double* workspace,
                             long long* prof, cudaStream_t stream);
// per-node rigid-body kernels over the whole batch: Jacobians at the iterate / values at the line-search trial points
cudaError_t launch_node_jac(int n, double* workspace, cudaStream_t stream);
cudaError_t launch_node_val(const KernelConst& kc, int n, double* workspace, cudaStream_t stream);
// prof: 16 cycle counters accumulated by CTA 0 in MPCB_PROFILE builds (may be null)
cudaError_t launch_admm(const KernelConst& kc, int n, int grid, int* counter, double* workspace, long long* prof, cudaStream_t stream);
cudaError_t launch_sample(const KernelConst& kc, const double* sx, const double* su, const double* sT, const double* target, int npts,
                          double* traj, mpcb_account* acc, int n, int grid, cudaStream_t stream);
// warm start: jerk-limited synchronised profiles sampled at the node times (8 lanes per problem, one joint per lane)
cudaError_t launch_otg_guess(const OtgConst& oc, int n, const double* x0, const double* xT, double* gx, double* gu, double* gT, cudaStream_t stream);
// problem generator: problem i of the launch draws from stream (seed, first + i)
cudaError_t launch_gen_problems(const mpcb_limits& lim, int kind, unsigned long long seed, unsigned long long first, int n, double* x0, double* xT,
                                cudaStream_t stream);
cudaError_t launch_fp64_peak(double* out, int blocks, int threads, int iters, cudaStream_t stream);

}  // namespace mpcb
