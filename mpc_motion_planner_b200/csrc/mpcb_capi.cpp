// Implementation of the C ABI (include/mpcb.h): handle management, host<->device staging, kernel launches.
// There is deliberately no CPU path: every compute entry point needs a CUDA device.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <utility>
#include <vector>

#include "mpcb.h"
#include "mpcb_gen.h"
#include "mpcb_host_util.h"
#include "mpcb_kernels.h"

using namespace mpcb;

struct mpcb_handle {
    int device = 0;
    int sm_count = 0;
    mpcb_settings st;
    mpcb_limits lim;
    KernelConst kc;
    int grid_max = 0;
    ShapeInfo si;
    double* d_gmat = nullptr;            // per-CTA matrix blocks of the larger collocation shapes (ShapeInfo::gmat_doubles each)
    int* d_counter = nullptr;            // one work counter per kernel launch of a chunk
    int n_counter = 0;
    double* d_workspace = nullptr;       // per-OCP blocks (iterate, multipliers, QP block) of one chunk
    int ws_ocps = 0;
    long long* d_prof = nullptr;         // 16 per-phase cycle counters (filled only by -DMPCB_PROFILE builds)
    // per-launch timing of the last mpcb_solve_batch: event pairs and the kind of each launch (0 = SQP stage, 1 = ADMM)
    std::vector<cudaEvent_t> kev;
    std::vector<int> kkind;
    int kused = 0;
    // staging buffers for MPCB_MEM_HOST calls (grown on demand)
    void* d_stage = nullptr;
    size_t stage_bytes = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // one in-flight stream per handle: work enqueued on another stream first waits for the handle's last enqueue (the
    // workspace, counters and staging buffers are per handle, not per stream)
    cudaEvent_t ev_last = nullptr;
    cudaStream_t last_stream = nullptr;
    // a chunk of at least two waves is solved as up to MAX_PARTS parts on the caller's stream and on internal streams: the persistent
    // kernels of one part leave SMs idle in their last wave, the kernels of the other parts fill them (each part has its own work
    // counters, workspace slice and matrix blocks; fork / join with events, so the call still looks like work on one stream)
    static const int MAX_PARTS = 4;
    int parts = 1;
    cudaStream_t aux[MAX_PARTS - 1] = {nullptr, nullptr, nullptr};
    cudaEvent_t ev_fork = nullptr, ev_join[MAX_PARTS - 1] = {nullptr, nullptr, nullptr};
    bool busy = false;
    bool timed = false;
    cudaStream_t timed_stream = nullptr;
    long long launches = 0;
    std::string err;
};

static std::string g_create_error;

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            h->err = std::string(#call) + ": " + cudaGetErrorString(e_);                           \
            return MPCB_ERR_CUDA;                                                                  \
        }                                                                                          \
    } while (0)

static void refresh_const(mpcb_handle* h) {
    h->kc.st = h->st;
    mpcb_fill_cheb(h->st.poly_order, h->st.num_seg, &h->kc.cheb);
    fill_limtab(&h->lim, h->kc.limtab);
    for (int j = 0; j < 7; j++) h->kc.jmax[j] = h->lim.jmax[j];
}

extern "C" {

int mpcb_create(const mpcb_settings* s, int device, mpcb_handle** out) {
    if (!s || !out) { g_create_error = "null argument"; return MPCB_ERR_BAD_ARG; }
    *out = nullptr;
    const char* why = "";
    if (!settings_supported(s, &why)) { g_create_error = why; return MPCB_ERR_UNSUPPORTED; }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        g_create_error = std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                         " (libmpcb has no CPU path)";
        return MPCB_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= ndev) { g_create_error = "device index out of range"; return MPCB_ERR_BAD_ARG; }
    mpcb_handle* h = new mpcb_handle();
    h->device = device; h->st = *s;
    default_limits(&h->lim);
    refresh_const(h);
    auto fail = [&](const char* what, cudaError_t ce) {
        g_create_error = std::string(what) + ": " + cudaGetErrorString(ce);
        mpcb_destroy(h);
        return (int)MPCB_ERR_CUDA;
    };
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail("cudaSetDevice", e);
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return fail("cudaGetDeviceProperties", e);
    h->sm_count = prop.multiProcessorCount;
    if ((e = shape_info(s->poly_order, s->num_seg, &h->si)) != cudaSuccess) return fail("shape_info", e);
    if ((size_t)prop.sharedMemPerBlockOptin < (h->si.stage_smem > h->si.admm_smem ? h->si.stage_smem : h->si.admm_smem)) {
        g_create_error = "device offers too little opt-in shared memory per block for the OCP kernel";
        mpcb_destroy(h);
        return MPCB_ERR_UNSUPPORTED;
    }
    if ((e = solve_prepare(s->poly_order, s->num_seg)) != cudaSuccess) return fail("cudaFuncSetAttribute(solve kernels)", e);
    h->grid_max = h->sm_count;      // one CTA per SM (shared memory bound)
    {
        // MPCB_PARTS = 2..4 switches the splitting on (+2.5 % OCP/s on 1000 problems, profiles/r2zb_parts.txt).  Off by default: with
        // overlapping kernels the per-kernel times of mpcb_last_kernel_ms (start-to-end spans, which then include the wait for a free SM)
        // no longer measure a kernel that has the GPU to itself, which is what the roofline of bench.py is defined on.
        const char* pe = getenv("MPCB_PARTS");
        h->parts = pe ? atoi(pe) : 1;
        if (h->parts < 1) h->parts = 1;
        if (h->parts > mpcb_handle::MAX_PARTS) h->parts = mpcb_handle::MAX_PARTS;
        for (int i = 0; i + 1 < h->parts; i++) {
            if ((e = cudaStreamCreateWithFlags(&h->aux[i], cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
            if ((e = cudaEventCreateWithFlags(&h->ev_join[i], cudaEventDisableTiming)) != cudaSuccess) return fail("cudaEventCreate", e);
        }
        if ((e = cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming)) != cudaSuccess) return fail("cudaEventCreate", e);
    }
    if (h->si.gmat_doubles && (e = cudaMalloc(&h->d_gmat, sizeof(double) * h->si.gmat_doubles * (size_t)h->grid_max * h->parts)) != cudaSuccess) return fail("cudaMalloc(matrix blocks)", e);
    if ((e = cudaMalloc(&h->d_prof, 512 * sizeof(long long))) != cudaSuccess || (e = cudaMemset(h->d_prof, 0, 512 * sizeof(long long))) != cudaSuccess) return fail("cudaMalloc(prof)", e);
    if ((e = cudaEventCreate(&h->ev0)) != cudaSuccess || (e = cudaEventCreate(&h->ev1)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&h->ev_last, cudaEventDisableTiming)) != cudaSuccess) return fail("cudaEventCreate", e);
    *out = h;
    return MPCB_OK;
}

void mpcb_destroy(mpcb_handle* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->d_counter) cudaFree(h->d_counter);
    if (h->d_workspace) cudaFree(h->d_workspace);
    if (h->d_stage) cudaFree(h->d_stage);
    if (h->d_gmat) cudaFree(h->d_gmat);
    if (h->d_prof) cudaFree(h->d_prof);
    for (cudaEvent_t e : h->kev) cudaEventDestroy(e);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->ev_last) cudaEventDestroy(h->ev_last);
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    for (int i = 0; i < mpcb_handle::MAX_PARTS - 1; i++) {
        if (h->ev_join[i]) cudaEventDestroy(h->ev_join[i]);
        if (h->aux[i]) cudaStreamDestroy(h->aux[i]);
    }
    delete h;
}

const char* mpcb_last_error(const mpcb_handle* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int mpcb_set_limits(mpcb_handle* h, const mpcb_limits* lim) {
    if (!h || !lim) return MPCB_ERR_BAD_ARG;
    h->lim = *lim;
    refresh_const(h);
    return MPCB_OK;
}

// serialise the handle's work across streams (see mpcb_handle::ev_last)
static cudaError_t handle_enter(mpcb_handle* h, cudaStream_t stream) {
    if (h->busy && h->last_stream != stream) return cudaStreamWaitEvent(stream, h->ev_last, 0);
    return cudaSuccess;
}
static cudaError_t handle_leave(mpcb_handle* h, cudaStream_t stream) {
    h->busy = true; h->last_stream = stream;
    return cudaEventRecord(h->ev_last, stream);
}
// before a per-handle buffer is freed: everything enqueued so far has finished
static void handle_quiesce(mpcb_handle* h) { if (h->busy) cudaEventSynchronize(h->ev_last); }

static int ensure_stage(mpcb_handle* h, size_t bytes) {
    if (bytes <= h->stage_bytes) return MPCB_OK;
    handle_quiesce(h);
    if (h->d_stage) { cudaFree(h->d_stage); h->d_stage = nullptr; h->stage_bytes = 0; }
    cudaError_t e = cudaMalloc(&h->d_stage, bytes);
    if (e != cudaSuccess) { h->err = std::string("cudaMalloc(stage): ") + cudaGetErrorString(e); return MPCB_ERR_ALLOC; }
    h->stage_bytes = bytes;
    return MPCB_OK;
}
static size_t al(size_t b) { return (b + 255) / 256 * 256; }

// Problems are solved in chunks of at most this many OCPs: the per-OCP workspace (QP block) is ~350 KB.
static const int CHUNK_MAX = 4096;

static int ensure_workspace(mpcb_handle* h, int ocps, int counters) {
    if (ocps > h->ws_ocps || counters > h->n_counter) handle_quiesce(h);
    if (ocps > h->ws_ocps) {
        if (h->d_workspace) { cudaFree(h->d_workspace); h->d_workspace = nullptr; h->ws_ocps = 0; }
        cudaError_t e = cudaMalloc(&h->d_workspace, sizeof(double) * h->si.ws_doubles * (size_t)ocps);
        if (e != cudaSuccess) { h->err = std::string("cudaMalloc(workspace): ") + cudaGetErrorString(e); return MPCB_ERR_ALLOC; }
        h->ws_ocps = ocps;
    }
    if (counters > h->n_counter) {
        if (h->d_counter) { cudaFree(h->d_counter); h->d_counter = nullptr; h->n_counter = 0; }
        cudaError_t e = cudaMalloc(&h->d_counter, sizeof(int) * counters);
        if (e != cudaSuccess) { h->err = std::string("cudaMalloc(counters): ") + cudaGetErrorString(e); return MPCB_ERR_ALLOC; }
        h->n_counter = counters;
    }
    return MPCB_OK;
}
// record an event pair around one kernel launch
static int timed_launch_begin(mpcb_handle* h, int kind, cudaStream_t stream) {
    while ((int)h->kev.size() < 2 * (h->kused + 1)) {
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return -1;
        h->kev.push_back(e);
    }
    if ((int)h->kkind.size() < h->kused + 1) h->kkind.resize(h->kused + 1);
    h->kkind[h->kused] = kind;
    return cudaEventRecord(h->kev[2 * h->kused], stream) == cudaSuccess ? 0 : -1;
}
static int timed_launch_end(mpcb_handle* h, cudaStream_t stream) {
    const int rc = cudaEventRecord(h->kev[2 * h->kused + 1], stream) == cudaSuccess ? 0 : -1;
    h->kused++;
    h->launches++;
    return rc;
}

int mpcb_solve_batch(mpcb_handle* h, int n, const double* x0, const double* xT, const double* guess_x, const double* guess_u,
                     const double* guess_T, const double* guess_lam, double* sol_x, double* sol_u, double* sol_T, double* sol_lam,
                     mpcb_stats* stats, int mem, void* stream_) {
    if (!h) return MPCB_ERR_BAD_ARG;
    if (n < 0 || (n > 0 && (!x0 || !xT || !guess_x || !guess_u || !guess_T || !sol_x || !sol_u || !sol_T))) { h->err = "null or negative argument"; return MPCB_ERR_BAD_ARG; }
    if (mem != MPCB_MEM_HOST && mem != MPCB_MEM_DEVICE) { h->err = "mem must be MPCB_MEM_HOST or MPCB_MEM_DEVICE"; return MPCB_ERR_BAD_ARG; }
    if (n == 0) return MPCB_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    CK(cudaSetDevice(h->device));
    CK(handle_enter(h, stream));
    const int N = mpcb_num_nodes(&h->st), ND = mpcb_num_dual(&h->st);
    const size_t bx0 = sizeof(double) * 14 * n, bgx = sizeof(double) * 14 * N * n, bgu = sizeof(double) * 7 * N * n, bT = sizeof(double) * n,
                 blam = sizeof(double) * ND * n, bst = sizeof(mpcb_stats) * n;
    BatchPtrs bp;
    memset(&bp, 0, sizeof bp);
    if (mem == MPCB_MEM_DEVICE) {
        bp.x0 = x0; bp.xT = xT; bp.gx = guess_x; bp.gu = guess_u; bp.gT = guess_T; bp.glam = guess_lam;
        bp.sx = sol_x; bp.su = sol_u; bp.sT = sol_T; bp.slam = sol_lam; bp.stats = stats;
    } else {
        const size_t total = 2 * al(bx0) + 2 * al(bgx) + 2 * al(bgu) + 2 * al(bT) + (guess_lam ? al(blam) : 0) + (sol_lam ? al(blam) : 0) + (stats ? al(bst) : 0);
        int rc = ensure_stage(h, total);
        if (rc) return rc;
        char* p = (char*)h->d_stage;
        auto take = [&](size_t b) { char* r = p; p += al(b); return r; };
        double* dx0 = (double*)take(bx0); double* dxT = (double*)take(bx0);
        double* dgx = (double*)take(bgx); double* dgu = (double*)take(bgu); double* dgT = (double*)take(bT);
        double* dsx = (double*)take(bgx); double* dsu = (double*)take(bgu); double* dsT = (double*)take(bT);
        double* dgl = guess_lam ? (double*)take(blam) : nullptr;
        double* dsl = sol_lam ? (double*)take(blam) : nullptr;
        mpcb_stats* dst = stats ? (mpcb_stats*)take(bst) : nullptr;
        CK(cudaMemcpyAsync(dx0, x0, bx0, cudaMemcpyHostToDevice, stream));
        CK(cudaMemcpyAsync(dxT, xT, bx0, cudaMemcpyHostToDevice, stream));
        CK(cudaMemcpyAsync(dgx, guess_x, bgx, cudaMemcpyHostToDevice, stream));
        CK(cudaMemcpyAsync(dgu, guess_u, bgu, cudaMemcpyHostToDevice, stream));
        CK(cudaMemcpyAsync(dgT, guess_T, bT, cudaMemcpyHostToDevice, stream));
        if (dgl) CK(cudaMemcpyAsync(dgl, guess_lam, blam, cudaMemcpyHostToDevice, stream));
        bp.x0 = dx0; bp.xT = dxT; bp.gx = dgx; bp.gu = dgu; bp.gT = dgT; bp.glam = dgl;
        bp.sx = dsx; bp.su = dsu; bp.sT = dsT; bp.slam = dsl; bp.stats = dst;
    }
    const int chunk = n < CHUNK_MAX ? n : CHUNK_MAX;
    const int n_launch = 2 + 3 * h->st.sqp_max_iter;               // persistent kernels (one work counter each) per part of a chunk
    { int rc = ensure_workspace(h, chunk, n_launch * h->parts); if (rc) return rc; }
    h->kused = 0;
    const bool debug_sync = getenv("MPCB_DEBUG_SYNC") != nullptr;      // debugging aid: synchronise after every launch and name the failing kernel
    CK(cudaEventRecord(h->ev0, stream));
#define TIMED(s_, kind_, call_)                                                                                    \
    do {                                                                                                           \
        if (timed_launch_begin(h, kind_, s_)) { h->err = "cudaEventRecord failed"; return MPCB_ERR_CUDA; }         \
        CK(call_);                                                                                                 \
        timed_launch_end(h, s_);                                                                                   \
        if (debug_sync) {                                                                                          \
            cudaError_t se_ = cudaStreamSynchronize(s_);                                                           \
            if (se_ != cudaSuccess) {                                                                              \
                h->err = std::string(#call_) + " [kernel]: " + cudaGetErrorString(se_);                            \
                return MPCB_ERR_CUDA;                                                                              \
            }                                                                                                      \
        }                                                                                                          \
    } while (0)
    struct Part { int nn, grid; BatchPtrs cp; int* ctr; double* ws; double* gm; cudaStream_t s; };
    const bool generic = settings_generic_path(&h->st) != 0;
    for (int c0 = 0; c0 < n; c0 += chunk) {
        const int nn = (n - c0) < chunk ? (n - c0) : chunk;
        // parts of at least one wave each (a chunk below two waves stays whole: the latency path)
        int P = (h->parts > 1 && nn >= 2 * h->grid_max) ? nn / h->grid_max : 1;
        if (P > h->parts) P = h->parts;
        Part part[mpcb_handle::MAX_PARTS];
        for (int p = 0; p < P; p++) {
            const int a = (int)((long long)nn * p / P), b = (int)((long long)nn * (p + 1) / P), o = c0 + a;
            Part& pt = part[p];
            pt.nn = b - a; pt.grid = pt.nn < h->grid_max ? pt.nn : h->grid_max;
            pt.cp = bp;
            pt.cp.x0 += (size_t)14 * o; pt.cp.xT += (size_t)14 * o; pt.cp.gx += (size_t)14 * N * o; pt.cp.gu += (size_t)7 * N * o; pt.cp.gT += o;
            if (pt.cp.glam) pt.cp.glam += (size_t)ND * o;
            pt.cp.sx += (size_t)14 * N * o; pt.cp.su += (size_t)7 * N * o; pt.cp.sT += o;
            if (pt.cp.slam) pt.cp.slam += (size_t)ND * o;
            if (pt.cp.stats) pt.cp.stats += o;
            pt.ctr = h->d_counter + p * n_launch;
            pt.ws = h->d_workspace + h->si.ws_doubles * (size_t)a;
            pt.gm = h->d_gmat ? h->d_gmat + h->si.gmat_doubles * (size_t)h->grid_max * p : nullptr;
            pt.s = p == 0 ? stream : h->aux[p - 1];
        }
        CK(cudaMemsetAsync(h->d_counter, 0, sizeof(int) * n_launch * P, stream));
        if (P > 1) {
            CK(cudaEventRecord(h->ev_fork, stream));
            for (int p = 1; p < P; p++) CK(cudaStreamWaitEvent(part[p].s, h->ev_fork, 0));
        }
        int li = 0;
        // load -> [node Jacobians -> linearise/factorise -> ADMM -> node trial values -> line search/step] x sqp_max_iter, stage by
        // stage over the parts (the shape-generic / adaptive-rho path runs linearise + factorise + ADMM as one kernel, accounted as
        // the ADMM kind)
        for (int p = 0; p < P; p++) {
            Part& pt = part[p];
            TIMED(pt.s, MPCB_KERNEL_SQP_STAGE, launch_sqp_stage(h->kc, pt.cp, pt.nn, 0, 0, pt.grid, pt.ctr + li, pt.ws, pt.gm, h->d_prof, pt.s));
        }
        li++;
        for (int it = 1; it <= h->st.sqp_max_iter; it++) {
            for (int p = 0; p < P; p++) { Part& pt = part[p]; TIMED(pt.s, MPCB_KERNEL_NODE, launch_node_jac(h->kc, pt.nn, pt.ws, pt.s)); }
            if (generic) {
                for (int p = 0; p < P; p++) {
                    Part& pt = part[p];
                    TIMED(pt.s, MPCB_KERNEL_ADMM, launch_sqp_stage(h->kc, pt.cp, pt.nn, 3, it, pt.grid, pt.ctr + li, pt.ws, pt.gm, h->d_prof, pt.s));
                }
                li++;
            } else {
                for (int p = 0; p < P; p++) {
                    Part& pt = part[p];
                    TIMED(pt.s, MPCB_KERNEL_SQP_STAGE, launch_sqp_stage(h->kc, pt.cp, pt.nn, 1, it, pt.grid, pt.ctr + li, pt.ws, pt.gm, h->d_prof, pt.s));
                }
                li++;
                for (int p = 0; p < P; p++) {
                    Part& pt = part[p];
                    TIMED(pt.s, MPCB_KERNEL_ADMM, launch_admm(h->kc, pt.nn, pt.grid, pt.ctr + li, pt.ws, h->d_prof, pt.s));
                }
                li++;
            }
            for (int p = 0; p < P; p++) { Part& pt = part[p]; TIMED(pt.s, MPCB_KERNEL_NODE, launch_node_val(h->kc, pt.nn, pt.ws, pt.s)); }
            for (int p = 0; p < P; p++) {
                Part& pt = part[p];
                TIMED(pt.s, MPCB_KERNEL_SQP_STAGE, launch_sqp_stage(h->kc, pt.cp, pt.nn, 2, it, pt.grid, pt.ctr + li, pt.ws, pt.gm, h->d_prof, pt.s));
            }
            li++;
        }
        for (int p = 1; p < P; p++) {
            CK(cudaEventRecord(h->ev_join[p - 1], part[p].s));
            CK(cudaStreamWaitEvent(stream, h->ev_join[p - 1], 0));
        }
    }
#undef TIMED
    CK(cudaEventRecord(h->ev1, stream));
    h->timed = true; h->timed_stream = stream;
    if (mem == MPCB_MEM_HOST) {
        CK(cudaMemcpyAsync(sol_x, bp.sx, bgx, cudaMemcpyDeviceToHost, stream));
        CK(cudaMemcpyAsync(sol_u, bp.su, bgu, cudaMemcpyDeviceToHost, stream));
        CK(cudaMemcpyAsync(sol_T, bp.sT, bT, cudaMemcpyDeviceToHost, stream));
        if (sol_lam) CK(cudaMemcpyAsync(sol_lam, bp.slam, blam, cudaMemcpyDeviceToHost, stream));
        if (stats) CK(cudaMemcpyAsync(stats, bp.stats, bst, cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
    }
    CK(handle_leave(h, stream));
    return MPCB_OK;
}

// No-allocation variant of mpcb_solve_batch (device pointers only; no counterpart in the reference): the same launch sequence over
// the static __device__ scratch of mpcb_kernels.cu (noalloc_scratch), without the handle's lazily grown buffers and event list.
int mpcb_solve_batch_noalloc(mpcb_handle* h, int n, const double* x0, const double* xT, const double* guess_x, const double* guess_u,
                             const double* guess_T, const double* guess_lam, double* sol_x, double* sol_u, double* sol_T, double* sol_lam,
                             mpcb_stats* stats, void* stream_) {
    if (!h) return MPCB_ERR_BAD_ARG;
    if (n < 0 || (n > 0 && (!x0 || !xT || !guess_x || !guess_u || !guess_T || !sol_x || !sol_u || !sol_T))) { h->err = "null or negative argument"; return MPCB_ERR_BAD_ARG; }
    if (n == 0) return MPCB_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    CK(cudaSetDevice(h->device));
    double* ws = nullptr; int* counters = nullptr; size_t ws_doubles = 0; int n_counters = 0;
    CK(noalloc_scratch(h->device, &ws, &ws_doubles, &counters, &n_counters));      // cudaGetSymbolAddress, first call per device only
    size_t limit = ws_doubles / h->si.ws_doubles;                 // the base shape has the largest per-OCP workspace
    if (limit > (size_t)MPCB_NOALLOC_MAX_OCPS) limit = MPCB_NOALLOC_MAX_OCPS;
    if ((size_t)n > limit) {
        h->err = "mpcb_solve_batch_noalloc: n = " + std::to_string(n) + " exceeds the static scratch, which holds at most " + std::to_string(limit) +
                 " OCPs (MPCB_NOALLOC_MAX_OCPS); use mpcb_solve_batch or split the batch";
        return MPCB_ERR_BAD_ARG;
    }
    const int n_launch = 2 + 3 * h->st.sqp_max_iter;
    if (n_launch > n_counters) { h->err = "mpcb_solve_batch_noalloc: sqp_max_iter too large for the static work counters"; return MPCB_ERR_BAD_ARG; }
    CK(handle_enter(h, stream));
    BatchPtrs bp;
    memset(&bp, 0, sizeof bp);
    bp.x0 = x0; bp.xT = xT; bp.gx = guess_x; bp.gu = guess_u; bp.gT = guess_T; bp.glam = guess_lam;
    bp.sx = sol_x; bp.su = sol_u; bp.sT = sol_T; bp.slam = sol_lam; bp.stats = stats;
    const int grid = n < h->grid_max ? n : h->grid_max;
    CK(cudaMemsetAsync(counters, 0, sizeof(int) * n_launch, stream));      // the only part of the scratch that must start at zero
    int li = 0;
    const bool generic = settings_generic_path(&h->st) != 0;
    CK(launch_sqp_stage(h->kc, bp, n, 0, 0, grid, counters + li++, ws, h->d_gmat, h->d_prof, stream));
    long long launched = 1;
    for (int it = 1; it <= h->st.sqp_max_iter; it++) {
        CK(launch_node_jac(h->kc, n, ws, stream));
        if (generic) {
            CK(launch_sqp_stage(h->kc, bp, n, 3, it, grid, counters + li++, ws, h->d_gmat, h->d_prof, stream));
            launched += 4;
        } else {
            CK(launch_sqp_stage(h->kc, bp, n, 1, it, grid, counters + li++, ws, h->d_gmat, h->d_prof, stream));
            CK(launch_admm(h->kc, n, grid, counters + li++, ws, h->d_prof, stream));
            launched += 5;
        }
        CK(launch_node_val(h->kc, n, ws, stream));
        CK(launch_sqp_stage(h->kc, bp, n, 2, it, grid, counters + li++, ws, h->d_gmat, h->d_prof, stream));
    }
    h->launches += launched;
    CK(handle_leave(h, stream));
    return MPCB_OK;
}

int mpcb_sample_batch(mpcb_handle* h, int n, const double* sol_x, const double* sol_u, const double* sol_T, const double* target,
                      int npts, double* traj, mpcb_account* acc, int mem, void* stream_) {
    if (!h) return MPCB_ERR_BAD_ARG;
    if (n < 0 || npts < 1 || npts > 1000 || (n > 0 && (!sol_x || !sol_u || !sol_T)) || (acc && !target)) { h->err = "bad argument"; return MPCB_ERR_BAD_ARG; }
    if (mem != MPCB_MEM_HOST && mem != MPCB_MEM_DEVICE) { h->err = "mem must be MPCB_MEM_HOST or MPCB_MEM_DEVICE"; return MPCB_ERR_BAD_ARG; }
    if (n == 0 || (!traj && !acc)) return MPCB_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    CK(cudaSetDevice(h->device));
    CK(handle_enter(h, stream));
    const int N = mpcb_num_nodes(&h->st);
    const size_t bsx = sizeof(double) * 14 * N * n, bsu = sizeof(double) * 7 * N * n, bT = sizeof(double) * n, btg = sizeof(double) * 14 * n,
                 btraj = sizeof(double) * 29 * (npts + 1) * n, bacc = sizeof(mpcb_account) * n;
    const double *dsx = sol_x, *dsu = sol_u, *dsT = sol_T, *dtg = target;
    double* dtraj = traj;
    mpcb_account* dacc = acc;
    if (mem == MPCB_MEM_HOST) {
        const size_t total = al(bsx) + al(bsu) + al(bT) + (acc ? al(btg) + al(bacc) : 0) + (traj ? al(btraj) : 0);
        int rc = ensure_stage(h, total);
        if (rc) return rc;
        char* p = (char*)h->d_stage;
        auto take = [&](size_t b) { char* r = p; p += al(b); return r; };
        double* a = (double*)take(bsx); double* b = (double*)take(bsu); double* c = (double*)take(bT);
        CK(cudaMemcpyAsync(a, sol_x, bsx, cudaMemcpyHostToDevice, stream));
        CK(cudaMemcpyAsync(b, sol_u, bsu, cudaMemcpyHostToDevice, stream));
        CK(cudaMemcpyAsync(c, sol_T, bT, cudaMemcpyHostToDevice, stream));
        dsx = a; dsu = b; dsT = c;
        if (acc) {
            double* t = (double*)take(btg);
            CK(cudaMemcpyAsync(t, target, btg, cudaMemcpyHostToDevice, stream));
            dtg = t; dacc = (mpcb_account*)take(bacc);
        }
        if (traj) dtraj = (double*)take(btraj);
    }
    const int grid = n < 8 * h->sm_count ? n : 8 * h->sm_count;
    CK(launch_sample(h->kc, dsx, dsu, dsT, dtg, npts, dtraj, dacc, n, grid, stream));
    h->launches += 1;
    if (mem == MPCB_MEM_HOST) {
        if (traj) CK(cudaMemcpyAsync(traj, dtraj, btraj, cudaMemcpyDeviceToHost, stream));
        if (acc) CK(cudaMemcpyAsync(acc, dacc, bacc, cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
    }
    CK(handle_leave(h, stream));
    return MPCB_OK;
}

static int fill_otg_const(mpcb_handle* h, OtgConst* oc) {
    oc->N = mpcb_num_nodes(&h->st);
    if (oc->N > 64 || mpcb_time_nodes(&h->st, oc->tn) != MPCB_OK) { h->err = "unsupported number of nodes"; return MPCB_ERR_UNSUPPORTED; }
    for (int d = 0; d < 7; d++) {
        oc->vl[d] = h->lim.margins[1] * h->lim.vmax[d]; oc->al[d] = h->lim.margins[2] * h->lim.amax[d]; oc->jl[d] = h->lim.margins[4] * h->lim.jmax[d];
    }
    oc->p_ub = h->lim.p_ub;
    return MPCB_OK;
}

int mpcb_sample_profile_batch(mpcb_handle* h, int n, const double* x0, const double* xT, int npts, double* traj, mpcb_account* acc, double* T_out,
                              int mem, void* stream_) {
    if (!h) return MPCB_ERR_BAD_ARG;
    if (n < 0 || npts < 1 || npts > 1000 || (n > 0 && (!x0 || !xT))) { h->err = "bad argument"; return MPCB_ERR_BAD_ARG; }
    if (mem != MPCB_MEM_HOST && mem != MPCB_MEM_DEVICE) { h->err = "mem must be MPCB_MEM_HOST or MPCB_MEM_DEVICE"; return MPCB_ERR_BAD_ARG; }
    if (n == 0 || (!traj && !acc && !T_out)) return MPCB_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    CK(cudaSetDevice(h->device));
    CK(handle_enter(h, stream));
    OtgConst oc;
    { int rc = fill_otg_const(h, &oc); if (rc) return rc; }
    const size_t b14 = sizeof(double) * 14 * n, btraj = sizeof(double) * 29 * (npts + 1) * n, bacc = sizeof(mpcb_account) * n, bT = sizeof(double) * n;
    const double *dx0 = x0, *dxT = xT;
    double *dtraj = traj, *dT = T_out;
    mpcb_account* dacc = acc;
    if (mem == MPCB_MEM_HOST) {
        int rc = ensure_stage(h, 2 * al(b14) + (traj ? al(btraj) : 0) + (acc ? al(bacc) : 0) + (T_out ? al(bT) : 0));
        if (rc) return rc;
        char* p = (char*)h->d_stage;
        auto take = [&](size_t b) { char* r = p; p += al(b); return r; };
        double* a = (double*)take(b14); double* b = (double*)take(b14);
        CK(cudaMemcpyAsync(a, x0, b14, cudaMemcpyHostToDevice, stream));
        CK(cudaMemcpyAsync(b, xT, b14, cudaMemcpyHostToDevice, stream));
        dx0 = a; dxT = b;
        if (traj) dtraj = (double*)take(btraj);
        if (acc) dacc = (mpcb_account*)take(bacc);
        if (T_out) dT = (double*)take(bT);
    }
    const int grid = n < 8 * h->sm_count ? n : 8 * h->sm_count;
    CK(launch_sample_profile(oc, h->kc, n, dx0, dxT, npts, dtraj, dacc, dT, grid, stream));
    h->launches += 1;
    if (mem == MPCB_MEM_HOST) {
        if (traj) CK(cudaMemcpyAsync(traj, dtraj, btraj, cudaMemcpyDeviceToHost, stream));
        if (acc) CK(cudaMemcpyAsync(acc, dacc, bacc, cudaMemcpyDeviceToHost, stream));
        if (T_out) CK(cudaMemcpyAsync(T_out, dT, bT, cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
    }
    CK(handle_leave(h, stream));
    return MPCB_OK;
}

int mpcb_guess_batch_gpu(mpcb_handle* h, int n, const double* x0, const double* xT, double* gx, double* gu, double* gT, int mem, void* stream_) {
    if (!h) return MPCB_ERR_BAD_ARG;
    if (n < 0 || (n > 0 && (!x0 || !xT || !gx || !gu || !gT))) { h->err = "bad argument"; return MPCB_ERR_BAD_ARG; }
    if (mem != MPCB_MEM_HOST && mem != MPCB_MEM_DEVICE) { h->err = "mem must be MPCB_MEM_HOST or MPCB_MEM_DEVICE"; return MPCB_ERR_BAD_ARG; }
    if (n == 0) return MPCB_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    CK(cudaSetDevice(h->device));
    CK(handle_enter(h, stream));
    OtgConst oc;
    { int rc = fill_otg_const(h, &oc); if (rc) return rc; }
    const int N = oc.N;
    const size_t b14 = sizeof(double) * 14 * n, bgx = sizeof(double) * 14 * N * n, bgu = sizeof(double) * 7 * N * n, bT = sizeof(double) * n;
    const double *dx0 = x0, *dxT = xT;
    double *dgx = gx, *dgu = gu, *dgT = gT;
    if (mem == MPCB_MEM_HOST) {
        int rc = ensure_stage(h, 2 * al(b14) + al(bgx) + al(bgu) + al(bT));
        if (rc) return rc;
        char* p = (char*)h->d_stage;
        auto take = [&](size_t b) { char* r = p; p += al(b); return r; };
        double* a = (double*)take(b14); double* b = (double*)take(b14);
        dgx = (double*)take(bgx); dgu = (double*)take(bgu); dgT = (double*)take(bT);
        CK(cudaMemcpyAsync(a, x0, b14, cudaMemcpyHostToDevice, stream));
        CK(cudaMemcpyAsync(b, xT, b14, cudaMemcpyHostToDevice, stream));
        dx0 = a; dxT = b;
    }
    CK(launch_otg_guess(oc, n, dx0, dxT, dgx, dgu, dgT, stream));
    h->launches += 1;
    if (mem == MPCB_MEM_HOST) {
        CK(cudaMemcpyAsync(gx, dgx, bgx, cudaMemcpyDeviceToHost, stream));
        CK(cudaMemcpyAsync(gu, dgu, bgu, cudaMemcpyDeviceToHost, stream));
        CK(cudaMemcpyAsync(gT, dgT, bT, cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
    }
    CK(handle_leave(h, stream));
    return MPCB_OK;
}

int mpcb_generate_problems_gpu(mpcb_handle* h, int kind, unsigned long long seed, unsigned long long first, int n, double* x0, double* xT, int mem,
                               void* stream_) {
    if (!h) return MPCB_ERR_BAD_ARG;
    if (n < 0 || (n > 0 && (!x0 || !xT)) || (kind != 0 && kind != 1)) { h->err = "bad argument"; return MPCB_ERR_BAD_ARG; }
    if (!(h->lim.min_height <= MPCB_GEN_MAX_HEIGHT)) { h->err = "min_height is out of the robot's reach: the rejection sampler would not terminate"; return MPCB_ERR_BAD_ARG; }
    if (mem != MPCB_MEM_HOST && mem != MPCB_MEM_DEVICE) { h->err = "mem must be MPCB_MEM_HOST or MPCB_MEM_DEVICE"; return MPCB_ERR_BAD_ARG; }
    if (n == 0) return MPCB_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    CK(cudaSetDevice(h->device));
    CK(handle_enter(h, stream));
    const size_t b14 = sizeof(double) * 14 * n;
    double *da = x0, *db = xT;
    if (mem == MPCB_MEM_HOST) {
        int rc = ensure_stage(h, 2 * al(b14));
        if (rc) return rc;
        da = (double*)h->d_stage; db = (double*)((char*)h->d_stage + al(b14));
    }
    CK(launch_gen_problems(h->lim, kind, seed, first, n, da, db, stream));
    h->launches += 1;
    if (mem == MPCB_MEM_HOST) {
        CK(cudaMemcpyAsync(x0, da, b14, cudaMemcpyDeviceToHost, stream));
        CK(cudaMemcpyAsync(xT, db, b14, cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
    }
    CK(handle_leave(h, stream));
    return MPCB_OK;
}

double mpcb_last_solve_ms(mpcb_handle* h) {
    if (!h || !h->timed) return -1.0;
    cudaSetDevice(h->device);
    if (cudaEventSynchronize(h->ev1) != cudaSuccess) return -1.0;
    float ms = -1.f;
    if (cudaEventElapsedTime(&ms, h->ev0, h->ev1) != cudaSuccess) return -1.0;
    return (double)ms;
}
long long mpcb_launch_count(const mpcb_handle* h) { return h ? h->launches : 0; }

double mpcb_last_kernel_ms(mpcb_handle* h, int kind) {
    if (!h || !h->timed) return -1.0;
    cudaSetDevice(h->device);
    if (cudaEventSynchronize(h->ev1) != cudaSuccess) return -1.0;
    // the launches of one kind may overlap (parts on different streams): the length of the UNION of their intervals, measured from the
    // start event of the call (equal to the sum of the durations when nothing overlaps)
    std::vector<std::pair<float, float>> iv;
    for (int i = 0; i < h->kused; i++) {
        if (h->kkind[i] != kind) continue;
        float a = 0.f, b = 0.f;
        if (cudaEventElapsedTime(&a, h->ev0, h->kev[2 * i]) != cudaSuccess || cudaEventElapsedTime(&b, h->ev0, h->kev[2 * i + 1]) != cudaSuccess) return -1.0;
        iv.push_back(std::make_pair(a, b));
    }
    std::sort(iv.begin(), iv.end());
    double total = 0;
    float lo = 0.f, hi = -1.f;
    for (size_t i = 0; i < iv.size(); i++) {
        if (hi < lo || iv[i].first > hi) { if (hi >= lo) total += hi - lo; lo = iv[i].first; hi = iv[i].second; }
        else if (iv[i].second > hi) hi = iv[i].second;
    }
    if (hi >= lo) total += hi - lo;
    return total;
}
int mpcb_debug_phase_cycles(mpcb_handle* h, long long* out32, int reset) {   // 512 slots: 0..31 phases, 64.. per-warp stamps
    if (!h || !out32) return MPCB_ERR_BAD_ARG;
    CK(cudaSetDevice(h->device));
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(out32, h->d_prof, 512 * sizeof(long long), cudaMemcpyDeviceToHost));
    if (reset) CK(cudaMemset(h->d_prof, 0, 512 * sizeof(long long)));
    return MPCB_OK;
}
int mpcb_last_kernel_launches(const mpcb_handle* h, int kind) {
    if (!h) return 0;
    int c = 0;
    for (int i = 0; i < h->kused; i++) c += (h->kkind[i] == kind);
    return c;
}

int mpcb_debug_node_constraints(mpcb_handle* h, int n, const double* qva, double* g_val, double* g_jac, double* Jg) {
    if (!h || n < 0 || (n > 0 && (!qva || !g_val || !g_jac || !Jg))) return MPCB_ERR_BAD_ARG;
    if (n == 0) return MPCB_OK;
    CK(cudaSetDevice(h->device));
    const size_t bin = sizeof(double) * 21 * n, bg = sizeof(double) * 8 * n, bj = sizeof(double) * 176 * n;
    int rc = ensure_stage(h, al(bin) + 2 * al(bg) + al(bj));
    if (rc) return rc;
    char* p = (char*)h->d_stage;
    double* din = (double*)p; p += al(bin);
    double* dv = (double*)p; p += al(bg);
    double* dg = (double*)p; p += al(bg);
    double* dj = (double*)p;
    CK(cudaMemcpy(din, qva, bin, cudaMemcpyHostToDevice));
    CK(launch_node_debug(n, din, dv, dg, dj, 0));
    h->launches += 1;
    CK(cudaMemcpy(g_val, dv, bg, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(g_jac, dg, bg, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(Jg, dj, bj, cudaMemcpyDeviceToHost));
    return MPCB_OK;
}

double mpcb_measure_fp64_tflops(mpcb_handle* h) {
    if (!h) return -1.0;
    cudaSetDevice(h->device);
    const int blocks = h->sm_count * 8, threads = 256, iters = 1 << 16;
    double* out = nullptr;
    if (cudaMalloc(&out, sizeof(double) * blocks * threads) != cudaSuccess) return -1.0;
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    double best = 0;
    for (int rep = 0; rep < 4; rep++) {
        cudaEventRecord(a, 0);
        launch_fp64_peak(out, blocks, threads, iters, 0);
        cudaEventRecord(b, 0);
        if (cudaEventSynchronize(b) != cudaSuccess) { best = -1.0; break; }
        float ms = 0;
        cudaEventElapsedTime(&ms, a, b);
        h->launches += 1;
        const double fl = 2.0 * 8.0 * (double)iters * blocks * threads;
        const double tf = fl / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(a); cudaEventDestroy(b); cudaFree(out);
    return best;
}

}  // extern "C"
