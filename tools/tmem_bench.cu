// Micro-benchmark: tensor memory (TMEM) as thread-private operand storage for FP64 mat-vecs.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tmem_bench tools/tmem_bench.cu && /tmp/tmem_bench
// Measures tcgen05.ld (LDTM) throughput with 4/8/16 warps, alone and while the LSU streams shared memory, and the
// rate of a TMEM-fed DFMA loop.  Output feeds DESIGN.md; nothing here is product code.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define REP 512

__device__ __forceinline__ void tmem_alloc(uint32_t* slot) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;\n" ::"l"((uint64_t)__cvta_generic_to_shared(slot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
}
__device__ __forceinline__ void tmem_free(uint32_t addr) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;\n" ::"r"(addr));
}
#define R8(v, o) "=r"(v[o]), "=r"(v[o + 1]), "=r"(v[o + 2]), "=r"(v[o + 3]), "=r"(v[o + 4]), "=r"(v[o + 5]), "=r"(v[o + 6]), "=r"(v[o + 7])
#define W8(v, o) "r"(v[o]), "r"(v[o + 1]), "r"(v[o + 2]), "r"(v[o + 3]), "r"(v[o + 4]), "r"(v[o + 5]), "r"(v[o + 6]), "r"(v[o + 7])
__device__ __forceinline__ void ldtm32(uint32_t* v, uint32_t a) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];\n"
        : R8(v, 0), R8(v, 8), R8(v, 16), R8(v, 24)
        : "r"(a));
}
__device__ __forceinline__ void ldtm16(uint32_t* v, uint32_t a) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
                 : R8(v, 0), R8(v, 8)
                 : "r"(a));
}
__device__ __forceinline__ void sttm32(const uint32_t* v, uint32_t a) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};\n" ::"r"(a),
        W8(v, 0), W8(v, 8), W8(v, 16), W8(v, 24));
}
__device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }
__device__ __forceinline__ void wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory"); }

// MODE 0: LDTM x32 only; 1: LDTM x32 + 16 DFMA per load (the mat-vec inner loop, vector operand from registers);
// 2: like 1 plus LDS.128 broadcast of the vector operand (8 loads per 16 FMAs); 3: LDS-only reference (matrix from smem, LDS.64
// consecutive + LDS.128 broadcast vector); 4: LDTM x16
template <int MODE>
__global__ void k_tmem(double* out, long long* cyc, int* ok) {
    __shared__ uint32_t slot;
    __shared__ __align__(16) double sm[4096];
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
    for (int i = tid; i < 4096; i += blockDim.x) sm[i] = 1.0 + 1e-9 * i;
    if (w == 0) tmem_alloc(&slot);
    asm volatile("tcgen05.fence::before_thread_sync;\n");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;\n");
    const uint32_t base = slot;
    // warp w owns lanes 32 (w % 4) .. +31 and columns 128 (w / 4) .. +127  (64 doubles per thread)
    const uint32_t mine = base + ((uint32_t)(32 * (w & 3)) << 16) + 128 * (w >> 2);
    uint32_t v[32];
    for (int q = 0; q < 4; q++) {
        for (int j = 0; j < 16; j++) {
            const double d = 1.0 + 1e-6 * (tid * 64 + q * 16 + j);
            v[2 * j] = (uint32_t)__double_as_longlong(d);
            v[2 * j + 1] = (uint32_t)(__double_as_longlong(d) >> 32);
        }
        sttm32(v, mine + 32 * q);
    }
    wait_st();
    // read back check
    {
        ldtm32(v, mine + 32);
        wait_ld();
        const double d = __longlong_as_double(((long long)v[3] << 32) | v[2]);
        const double e = 1.0 + 1e-6 * (tid * 64 + 16 + 1);
        if (d != e) atomicAdd(ok, 1);
    }
    // loads at column offsets that are not multiples of the load width (the ADMM kernel packs uneven per-thread pieces)
    for (int off = 1; off < 40; off += 3) {         // offset in doubles (2 columns each)
        ldtm16(v, mine + 2 * off);
        wait_ld();
        for (int j = 0; j < 8; j++) {
            const double d = __longlong_as_double(((long long)v[2 * j + 1] << 32) | v[2 * j]);
            if (d != 1.0 + 1e-6 * (tid * 64 + off + j)) atomicAdd(ok, 1);
        }
    }
    __syncthreads();
    double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    double x[16];
#pragma unroll
    for (int j = 0; j < 16; j++) x[j] = 1.0 + 1e-3 * j + 1e-6 * lane;
    long long t0 = clock64();
#pragma unroll 2
    for (int i = 0; i < REP; i++) {
        if (MODE == 3) {
            const double* m = sm + ((i & 3) * 1024) + lane;
            const double* xv = sm + 2048 + (i & 7) * 16;
#pragma unroll
            for (int j = 0; j < 16; j += 4) {
                const double2 p = *reinterpret_cast<const double2*>(xv + j), q2 = *reinterpret_cast<const double2*>(xv + j + 2);
                a0 = fma(m[32 * j], p.x, a0); a1 = fma(m[32 * (j + 1)], p.y, a1); a2 = fma(m[32 * (j + 2)], q2.x, a2); a3 = fma(m[32 * (j + 3)], q2.y, a3);
            }
            continue;
        }
        if (MODE == 4) {
            ldtm16(v, mine + 16 * (i & 7));
            wait_ld();
            a0 += __longlong_as_double(((long long)v[1] << 32) | v[0]);
            continue;
        }
        ldtm32(v, mine + 32 * (i & 3));
        wait_ld();
        if (MODE == 0) {
            a0 += __longlong_as_double(((long long)v[1] << 32) | v[0]);
        } else {
            if (MODE == 2) {
                const double* xv = sm + 2048 + (i & 7) * 16;
#pragma unroll
                for (int j = 0; j < 16; j += 2) { const double2 p = *reinterpret_cast<const double2*>(xv + j); x[j] = p.x; x[j + 1] = p.y; }
            }
#pragma unroll
            for (int j = 0; j < 16; j += 4) {
                a0 = fma(__longlong_as_double(((long long)v[2 * j + 1] << 32) | v[2 * j]), x[j], a0);
                a1 = fma(__longlong_as_double(((long long)v[2 * j + 3] << 32) | v[2 * j + 2]), x[j + 1], a1);
                a2 = fma(__longlong_as_double(((long long)v[2 * j + 5] << 32) | v[2 * j + 4]), x[j + 2], a2);
                a3 = fma(__longlong_as_double(((long long)v[2 * j + 7] << 32) | v[2 * j + 6]), x[j + 3], a3);
            }
        }
    }
    long long t1 = clock64();
    __syncthreads();
    out[blockIdx.x * blockDim.x + tid] = (a0 + a1) + (a2 + a3);
    if (tid == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
    __syncthreads();
    if (w == 0) tmem_free(base);
}

template <int MODE>
static void run(const char* name, int warps, double* out, long long* cyc, int* ok) {
    cudaMemset(ok, 0, sizeof(int));
    k_tmem<MODE><<<148, 32 * warps>>>(out, cyc, ok);
    k_tmem<MODE><<<148, 32 * warps>>>(out, cyc, ok);
    cudaError_t e = cudaDeviceSynchronize();
    long long c; int bad;
    cudaMemcpy(&c, cyc, sizeof(c), cudaMemcpyDeviceToHost);
    cudaMemcpy(&bad, ok, sizeof(bad), cudaMemcpyDeviceToHost);
    printf("%-64s %2d warps: %8.2f cycles per loop iteration per warp-slot, %7.3f cycles per warp-iteration per SM  (err=%s, readback mismatches=%d)\n", name, warps,
           (double)c / REP, (double)c / REP / warps, cudaGetErrorString(e), bad);
}

int main() {
    double* out; long long* cyc; int* ok;
    cudaMalloc(&out, 148 * 1024 * sizeof(double));
    cudaMalloc(&cyc, 64);
    cudaMalloc(&ok, 64);
    for (int warps = 4; warps <= 16; warps *= 2) {
        run<0>("LDTM 32x32b.x32 (16 doubles per thread)", warps, out, cyc, ok);
        run<4>("LDTM 32x32b.x16 (8 doubles per thread)", warps, out, cyc, ok);
        run<1>("LDTM x32 + 16 DFMA, vector in registers", warps, out, cyc, ok);
        run<2>("LDTM x32 + 8 LDS.128 broadcast + 16 DFMA", warps, out, cyc, ok);
        run<3>("16 LDS.64 consecutive + 8 LDS.128 broadcast + 16 DFMA (smem only)", warps, out, cyc, ok);
    }
    return 0;
}
