// Micro-benchmark: non-tensor FP32 throughput on this GPU, for the field-evaluation roofline.
//   ffma      : dependent FFMA chains               (2 flop / instr / lane)   -- "fast" mode ceiling
//   fmul_fadd : FMUL then FADD, separately rounded  (1 flop / instr / lane)   -- "exact" (parity) mode ceiling
//   f32x2     : packed FFMA2(x,a,-0)+FADD2 (sm_100 f32x2), separately rounded   -- exact mode with packed issue (half the instructions)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o fp32_peak fp32_peak.cu
// Prints one JSON line.
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ILP = 8;
constexpr int ITERS = 4096;

__global__ void k_ffma(float* out, float a, float b) {
    float v[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) v[i] = threadIdx.x * 1e-3f + i;
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) v[i] = __fmaf_rn(v[i], a, b);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_fmul_fadd(float* out, float a, float b) {
    float v[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) v[i] = threadIdx.x * 1e-3f + i;
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) v[i] = __fadd_rn(__fmul_rn(v[i], a), b);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// Exact mode on register pairs, as the library issues it (csg_eval.cuh mul2 / add2): the product is fma(x, a, nz) with nz = -0.0 read
// from constant memory, so that ptxas cannot see a plain multiplication and contract it with the addition that follows (it does
// that to mul.rn.f32x2 + add.rn.f32x2, and to fma(x,a,-0) + fma(p,1,b) with literal constants: round 1 of this tool measured that
// FUSED form by mistake and reported it as an "exact packed peak" twice the real one).  FFMA2 / FADD2 issue at half the rate of their
// scalar forms, so packing saves issue slots, not FP32-pipe cycles: the exact-mode ceiling stays that of FMUL+FADD.
__constant__ float c_negzero = -0.0f;
__device__ __forceinline__ float2 mul2(float2 a, float2 b, float nz) {
    unsigned long long r, x = *reinterpret_cast<unsigned long long*>(&a), y = *reinterpret_cast<unsigned long long*>(&b);
    float2 z2 = make_float2(nz, nz);
    unsigned long long z = *reinterpret_cast<unsigned long long*>(&z2);
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(x), "l"(y), "l"(z));
    return *reinterpret_cast<float2*>(&r);
}
__device__ __forceinline__ float2 add2(float2 a, float2 b) {
    unsigned long long r, x = *reinterpret_cast<unsigned long long*>(&a), y = *reinterpret_cast<unsigned long long*>(&b);
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(x), "l"(y));
    return *reinterpret_cast<float2*>(&r);
}
__global__ void k_f32x2(float* out, float a, float b) {
    float2 v[ILP];
    const float2 aa = make_float2(a, a), bb = make_float2(b, b);
    const float nz = c_negzero;
#pragma unroll
    for (int i = 0; i < ILP; i++) v[i] = make_float2(threadIdx.x * 1e-3f + i, threadIdx.x * 2e-3f + i);
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) v[i] = add2(mul2(v[i], aa, nz), bb);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += v[i].x + v[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_ffma2(float* out, float a, float b) {
    float2 v[ILP];
    const float2 aa = make_float2(a, a), bb = make_float2(b, b);
#pragma unroll
    for (int i = 0; i < ILP; i++) v[i] = make_float2(threadIdx.x * 1e-3f + i, threadIdx.x * 2e-3f + i);
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) v[i] = __ffma2_rn(v[i], aa, bb);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += v[i].x + v[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class K>
static double run(K kern, float* d, int blocks, int threads, double flop_per_thread) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int w = 0; w < 3; w++) kern<<<blocks, threads>>>(d, 0.999f, 0.001f);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0);
        kern<<<blocks, threads>>>(d, 0.999f, 0.001f);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return flop_per_thread * blocks * threads / (best * 1e-3) / 1e12;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int blocks = p.multiProcessorCount * 8, threads = 256;
    float* d; cudaMalloc(&d, sizeof(float) * blocks * threads);
    double per = (double)ILP * ITERS;
    double t_ffma = run(k_ffma, d, blocks, threads, per * 2);
    double t_ma = run(k_fmul_fadd, d, blocks, threads, per * 2);
    double t_x2 = run(k_f32x2, d, blocks, threads, per * 4);
    double t_fma2 = run(k_ffma2, d, blocks, threads, per * 4);
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"ffma_tflops\": %.2f, \"fmul_fadd_tflops\": %.2f, \"fmul2_fadd2_tflops\": %.2f, \"ffma2_tflops\": %.2f}\n",
           p.name, p.multiProcessorCount, t_ffma, t_ma, t_x2, t_fma2);
    return 0;
}
