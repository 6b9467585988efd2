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

// Launch geometry and workspace sizes of one instantiated collocation shape (POLY_ORDER, NUM_SEG)
struct ShapeInfo {
    size_t stage_smem, admm_smem;   // dynamic shared memory of the stage kernel / of the (3,6) ADMM kernel (0: none)
    int stage_threads;
    size_t ws_doubles;              // per-OCP workspace
    size_t gmat_doubles;            // per-CTA matrix block in global memory (0: the matrices live in shared memory)
};
int shape_supported(int po, int ns);
cudaError_t shape_info(int po, int ns, ShapeInfo* si);
cudaError_t solve_prepare(int po, int ns);
// kind 0: load the guess; kind 1: linearise + factorise + export; kind 2: line search / step of SQP iteration it;
// kind 3: linearise + factorise + ADMM in one kernel (shape-generic / adaptive-rho path)
cudaError_t launch_sqp_stage(const KernelConst& kc, const BatchPtrs& bp, int n, int kind, int it, int grid, int* counter, double* workspace,
                             double* gmat, long long* prof, cudaStream_t stream);
// per-node rigid-body kernels over the whole batch: Jacobians at the iterate / values at the line-search trial points
cudaError_t launch_node_jac(const KernelConst& kc, int n, double* workspace, cudaStream_t stream);
cudaError_t launch_node_val(const KernelConst& kc, int n, double* workspace, cudaStream_t stream);
// prof: 16 cycle counters accumulated by CTA 0 in MPCB_PROFILE builds (may be null)
cudaError_t launch_admm(const KernelConst& kc, int n, int grid, int* counter, double* workspace, long long* prof, cudaStream_t stream);
cudaError_t launch_sample(const KernelConst& kc, const double* sx, const double* su, const double* sT, const double* target, int npts,
                          double* traj, mpcb_account* acc, int n, int grid, cudaStream_t stream);
// warm start: jerk-limited synchronised profiles sampled at the node times (8 lanes per problem, one joint per lane)
cudaError_t launch_otg_guess(const OtgConst& oc, int n, const double* x0, const double* xT, double* gx, double* gu, double* gT, cudaStream_t stream);
// warm-start profile read-out + benchmark accounting (the Ruckig half of examples/benchmark.cpp:72-114), one CTA per problem
cudaError_t launch_sample_profile(const OtgConst& oc, const KernelConst& kc, int n, const double* x0, const double* xT, int npts, double* traj,
                                  mpcb_account* acc, double* Tout, int grid, cudaStream_t stream);
// problem generator: problem i of the launch draws from stream (seed, first + i)
cudaError_t launch_gen_problems(const mpcb_limits& lim, int kind, unsigned long long seed, unsigned long long first, int n, double* x0, double* xT,
                                cudaStream_t stream);
// debug: node_constraints_value / node_constraints_jac of n states (q, v, a) on the device
cudaError_t launch_node_debug(int n, const double* qva, double* g_val, double* g_jac, double* Jg, cudaStream_t stream);
cudaError_t launch_fp64_peak(double* out, int blocks, int threads, int iters, cudaStream_t stream);
// scratch of mpcb_solve_batch_noalloc: device addresses of the file-scope static __device__ arrays of mpcb_kernels.cu on `device`
// (which must be current), looked up once per device
cudaError_t noalloc_scratch(int device, double** workspace, size_t* workspace_doubles, int** counters, int* n_counters);

}  // namespace mpcb
