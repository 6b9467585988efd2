// Tensor memory (TMEM, 256 KB per SM on sm_100a) used as THREAD-PRIVATE operand storage of the ADMM kernel.
//
// Every per-OCP operator of the ADMM iteration (overflow columns of W, F rows, F^T column chunks, C^-1 rows, row / box
// constants) is read by exactly one thread per iteration.  Kept in shared memory such an operand costs one LDS wavefront
// pair per 32 FMAs, and the shared-memory return path (128 B/clk/SM) -- not the FP64 pipe -- bounds the kernel
// (profiles/r1h_admm_kernel_raw.csv: 59 % LSU wavefront utilisation at 13 % FP64 utilisation).  tcgen05.ld with the
// 32x32b shape hands thread t of warp w the columns [c, c + N) of TMEM lane 32 (w % 4) + t, i.e. a private strip, at
// 690-910 B/clk/SM measured (profiles/r1h_tmem_microbench.txt) on a path separate from the LSU.
//
// A "slot" is one double = two 32-bit TMEM columns.  Warps w, w + 4, w + 8, w + 12 share one lane quadrant and split its
// 256 slots (tm_warp_base).  Loads are asynchronous: tm_ld<N>() issues, tm_wait() orders the registers behind
// tcgen05.wait::ld (the registers are passed through the wait as in/out operands so that the compiler cannot move their
// uses above it).  In the host emulation (MPCB_EMUL, tests only) a strip is a plain array in the per-thread state.
#pragma once
#include <stdint.h>
#include "mpcb_common.h"

namespace mpcb {

constexpr int TM_SLOTS = 72;          // largest per-thread strip (doubles); see AL<SH>::tm_size

#ifdef MPCB_EMUL
template <int N>
struct TmBuf {
    double d[N];
    double get(int j) const { return d[j]; }
};
template <int N>
inline void tm_ld(const double* strip, int slot, TmBuf<N>& b) { for (int j = 0; j < N; j++) b.d[j] = strip[slot + j]; }
template <int N>
inline void tm_ld_after(const double* strip, int slot, TmBuf<16>& b, double (&)[N]) { tm_ld(strip, slot, b); }
template <int N>
inline void tm_wait(TmBuf<N>&) {}
template <int N, int M>
inline void tm_wait(TmBuf<N>&, TmBuf<M>&) {}
template <int N, int M, int K>
inline void tm_wait(TmBuf<N>&, TmBuf<M>&, TmBuf<K>&) {}
#else
template <int N>
struct TmBuf {
    uint32_t r[2 * N];
    __device__ __forceinline__ double get(int j) const { return __hiloint2double((int)r[2 * j + 1], (int)r[2 * j]); }
};

__device__ __forceinline__ void tm_alloc512(uint32_t* smem_slot) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;\n" ::"l"((uint64_t)__cvta_generic_to_shared(smem_slot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
}
__device__ __forceinline__ void tm_free512(uint32_t base) { asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;\n" ::"r"(base)); }
__device__ __forceinline__ void tm_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tm_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }

#define MPCB_TM_O2(b, o) "=r"(b.r[o]), "=r"(b.r[o + 1])
#define MPCB_TM_O8(b, o) MPCB_TM_O2(b, o), MPCB_TM_O2(b, o + 2), MPCB_TM_O2(b, o + 4), MPCB_TM_O2(b, o + 6)
#define MPCB_TM_P2(b, o) "+r"(b.r[o]), "+r"(b.r[o + 1])
#define MPCB_TM_P8(b, o) MPCB_TM_P2(b, o), MPCB_TM_P2(b, o + 2), MPCB_TM_P2(b, o + 4), MPCB_TM_P2(b, o + 6)

__device__ __forceinline__ void tm_ld(uint32_t addr, int slot, TmBuf<1>& b) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0,%1}, [%2];\n" : MPCB_TM_O2(b, 0) : "r"(addr + 2 * slot));
}
__device__ __forceinline__ void tm_ld(uint32_t addr, int slot, TmBuf<2>& b) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];\n" : MPCB_TM_O2(b, 0), MPCB_TM_O2(b, 2) : "r"(addr + 2 * slot));
}
__device__ __forceinline__ void tm_ld(uint32_t addr, int slot, TmBuf<4>& b) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];\n" : MPCB_TM_O8(b, 0) : "r"(addr + 2 * slot));
}
__device__ __forceinline__ void tm_ld(uint32_t addr, int slot, TmBuf<8>& b) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
                 : MPCB_TM_O8(b, 0), MPCB_TM_O8(b, 8)
                 : "r"(addr + 2 * slot));
}
__device__ __forceinline__ void tm_ld(uint32_t addr, int slot, TmBuf<16>& b) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];\n"
        : MPCB_TM_O8(b, 0), MPCB_TM_O8(b, 8), MPCB_TM_O8(b, 16), MPCB_TM_O8(b, 24)
        : "r"(addr + 2 * slot));
}
// the same load with an ordering handle: `dep` (register values the following arithmetic reads) passes through the statement, so the
// compiler cannot schedule that arithmetic in front of the load (it otherwise sinks the load behind every independent multiply-add
// to shorten the live range of the 32 destination registers, and the tensor-memory latency is exposed)
template <int N>
__device__ __forceinline__ void tm_ld_after(uint32_t addr, int slot, TmBuf<16>& b, double (&dep)[N]) {
    static_assert(N >= 7, "at least seven values order the load");
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%39];\n"
        : MPCB_TM_O8(b, 0), MPCB_TM_O8(b, 8), MPCB_TM_O8(b, 16), MPCB_TM_O8(b, 24), "+d"(dep[0]), "+d"(dep[1]), "+d"(dep[2]), "+d"(dep[3]), "+d"(dep[4]),
          "+d"(dep[5]), "+d"(dep[6])
        : "r"(addr + 2 * slot));
}
// dependency-only pass of a buffer's registers through an (empty) volatile asm: keeps their uses behind the wait
__device__ __forceinline__ void tm_dep(TmBuf<1>& b) { asm volatile("" : MPCB_TM_P2(b, 0)); }
__device__ __forceinline__ void tm_dep(TmBuf<2>& b) { asm volatile("" : MPCB_TM_P2(b, 0), MPCB_TM_P2(b, 2)); }
__device__ __forceinline__ void tm_dep(TmBuf<4>& b) { asm volatile("" : MPCB_TM_P8(b, 0)); }
__device__ __forceinline__ void tm_dep(TmBuf<8>& b) { asm volatile("" : MPCB_TM_P8(b, 0), MPCB_TM_P8(b, 8)); }
__device__ __forceinline__ void tm_dep(TmBuf<16>& b) { asm volatile("" : MPCB_TM_P8(b, 0), MPCB_TM_P8(b, 8), MPCB_TM_P8(b, 16), MPCB_TM_P8(b, 24)); }
__device__ __forceinline__ void tm_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tm_wait(TmBuf<N>& a) { tm_wait_ld(); tm_dep(a); }
template <int N, int M>
__device__ __forceinline__ void tm_wait(TmBuf<N>& a, TmBuf<M>& b) { tm_wait_ld(); tm_dep(a); tm_dep(b); }
template <int N, int M, int K>
__device__ __forceinline__ void tm_wait(TmBuf<N>& a, TmBuf<M>& b, TmBuf<K>& c) { tm_wait_ld(); tm_dep(a); tm_dep(b); tm_dep(c); }

// 8 doubles -> the thread's strip
__device__ __forceinline__ void tm_st8(uint32_t addr, int slot, const double* v) {
    uint32_t r[16];
#pragma unroll
    for (int j = 0; j < 8; j++) { r[2 * j] = (uint32_t)__double2loint(v[j]); r[2 * j + 1] = (uint32_t)__double2hiint(v[j]); }
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};\n" ::"r"(addr + 2 * slot),
                 "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]),
                 "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
                 : "memory");
}
__device__ __forceinline__ void tm_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory"); }
#endif

// first slot of warp w's strip inside its lane quadrant (strip sizes: 56 | 72 x 7 | 56 x 5, see AL<SH>::tm_size)
MPCB_HD int tm_warp_base(int w) { return w < 4 ? 0 : (w < 8 ? 56 : (w < 12 ? 128 : (w == 15 ? 184 : 200))); }

}  // namespace mpcb
