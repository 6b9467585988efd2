// Micro-benchmarks of the SM resources the ADMM kernel is bound by (FP64 pipe, shared-memory return path, shuffles,
// barriers).  Build + run on a B200:   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/mb tools/microbench.cu && /tmp/mb
// Output feeds DESIGN.md section 5 (measured per-SM rates); nothing here is product code.
#include <cuda_runtime.h>
#include <stdio.h>

#define REP 2048

__global__ void k_dfma_latency(double* out, long long* cyc) {          // one warp, one dependent chain
    double a = threadIdx.x * 1e-9, b = 0.999999, c = 1e-7;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < REP; i++) a = fma(a, b, c);
    long long t1 = clock64();
    out[threadIdx.x] = a;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
template <int CH>
__global__ void k_dfma_tput(double* out, long long* cyc) {             // blockDim warps, CH independent chains each
    double a[CH];
#pragma unroll
    for (int j = 0; j < CH; j++) a[j] = threadIdx.x * 1e-9 + j;
    const double b = 0.999999, c = 1e-7;
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < REP; i++) {
#pragma unroll
        for (int j = 0; j < CH; j++) a[j] = fma(a[j], b, c);
    }
    __syncthreads();
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int j = 0; j < CH; j++) s += a[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
// shared-memory loads: mode 0 = LDS.64 all lanes same address (broadcast), 1 = LDS.64 consecutive (conflict-free),
// 2 = LDS.128 broadcast, 3 = LDS.128 consecutive
template <int MODE>
__global__ void k_lds(double* out, long long* cyc) {
    __shared__ __align__(16) double sm[4096];
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = i;
    __syncthreads();
    double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
    const int lane = threadIdx.x & 31;
    long long t0 = clock64();
#pragma unroll 8
    for (int i = 0; i < REP; i++) {
        const int base = (i * 8) & 2047;
        if (MODE == 0) { s0 += sm[base]; }
        if (MODE == 1) { s0 += sm[base + lane]; }
        if (MODE == 2) { const double2 v = *reinterpret_cast<const double2*>(sm + base); s0 += v.x; s1 += v.y; }
        if (MODE == 3) { const double2 v = *reinterpret_cast<const double2*>(sm + base + 2 * lane); s0 += v.x; s1 += v.y; }
    }
    __syncthreads();
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = s0 + s1 + s2 + s3;
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_shfl(double* out, long long* cyc) {
    double a = threadIdx.x;
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 8
    for (int i = 0; i < REP; i++) a += __shfl_xor_sync(0xffffffffu, a, 1 + (i & 15));
    __syncthreads();
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = a;
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_shfl_indep(double* out, long long* cyc) {            // 4 independent shuffles per step (throughput)
    double a = threadIdx.x, b = a + 1, c = a + 2, d = a + 3;
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < REP; i++) {
        const double ta = __shfl_xor_sync(0xffffffffu, a, 1), tb = __shfl_xor_sync(0xffffffffu, b, 2);
        const double tc = __shfl_xor_sync(0xffffffffu, c, 4), td = __shfl_xor_sync(0xffffffffu, d, 8);
        a = tb; b = tc; c = td; d = ta;
    }
    __syncthreads();
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = a + b + c + d;
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_bar(double* out, long long* cyc) {
    __shared__ double sm[1024];
    sm[threadIdx.x] = threadIdx.x;
    __syncthreads();
    double a = 0;
    long long t0 = clock64();
    for (int i = 0; i < REP; i++) {
        a += sm[(threadIdx.x + i) & 1023];
        __syncthreads();
        sm[threadIdx.x] = a;
        __syncthreads();
    }
    long long t1 = clock64();
    out[threadIdx.x] = a;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_ddiv(double* out, long long* cyc) {
    double a = 1.0 + threadIdx.x * 1e-3, b = 1.000001;
    long long t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < 256; i++) a = a / b + 1e-9;
    long long t1 = clock64();
    out[threadIdx.x] = a;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}

int main() {
    double* out; long long* cyc; long long h;
    cudaMalloc(&out, sizeof(double) * 1 << 20); cudaMalloc(&cyc, 64);
    auto get = [&]() { cudaDeviceSynchronize(); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); return (double)h; };
    k_dfma_latency<<<1, 32>>>(out, cyc); k_dfma_latency<<<1, 32>>>(out, cyc);
    printf("DFMA dependent latency            : %.2f cycles\n", get() / REP);
    for (int w = 4; w <= 16; w *= 2) {
        k_dfma_tput<1><<<1, 32 * w>>>(out, cyc); printf("DFMA %2d warps x 1 chain  : %.3f cycles per warp-DFMA per SM\n", w, get() / (REP * 1.0 * w));
        k_dfma_tput<2><<<1, 32 * w>>>(out, cyc); printf("DFMA %2d warps x 2 chains : %.3f\n", w, get() / (REP * 2.0 * w));
        k_dfma_tput<4><<<1, 32 * w>>>(out, cyc); printf("DFMA %2d warps x 4 chains : %.3f\n", w, get() / (REP * 4.0 * w));
        k_dfma_tput<8><<<1, 32 * w>>>(out, cyc); printf("DFMA %2d warps x 8 chains : %.3f\n", w, get() / (REP * 8.0 * w));
    }
    const int W = 16;
    k_lds<0><<<1, 32 * W>>>(out, cyc); printf("LDS.64 broadcast   16 warps: %.3f cycles per warp-load per SM\n", get() / (REP * 1.0 * W));
    k_lds<1><<<1, 32 * W>>>(out, cyc); printf("LDS.64 consecutive 16 warps: %.3f\n", get() / (REP * 1.0 * W));
    k_lds<2><<<1, 32 * W>>>(out, cyc); printf("LDS.128 broadcast  16 warps: %.3f\n", get() / (REP * 1.0 * W));
    k_lds<3><<<1, 32 * W>>>(out, cyc); printf("LDS.128 consecutive16 warps: %.3f\n", get() / (REP * 1.0 * W));
    k_lds<0><<<1, 32>>>(out, cyc); printf("LDS.64 broadcast   1 warp (latency-ish): %.3f\n", get() / REP);
    k_shfl<<<1, 32>>>(out, cyc); printf("SHFL.64 dependent 1 warp (latency incl. DADD): %.2f\n", get() / REP);
    k_shfl_indep<<<1, 32 * W>>>(out, cyc); printf("SHFL.64 throughput 16 warps: %.3f cycles per warp-64bit-shuffle per SM\n", get() / (REP * 4.0 * W));
    k_bar<<<1, 512>>>(out, cyc); printf("LDS + BAR + STS + BAR, 16 warps: %.1f cycles per round (2 barriers)\n", get() / REP);
    k_bar<<<1, 256>>>(out, cyc); printf("LDS + BAR + STS + BAR,  8 warps: %.1f\n", get() / REP);
    k_ddiv<<<1, 32>>>(out, cyc); printf("DDIV dependent (a/b + c): %.1f cycles\n", get() / 256);
    return 0;
}
