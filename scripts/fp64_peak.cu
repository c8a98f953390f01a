// Measures the FP64 peaks the rooflines of DESIGN.md are quoted against ("FP64 FMA and DMMA peaks are not in
// MEASURED_PEAKS.json - the builder must measure them on the box"):
//   dfma   dependent-free DFMA chains, 8 accumulators per thread, all SMs
//   dmma   mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64, 4 independent accumulator tiles per warp
//   copy   device-to-device double2 copy (read + write), a second opinion on MEASURED_PEAKS.json hbm_gbs
//   read   read-only stream with the same load instruction as the cached apply (ld.global.nc.L1::no_allocate)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/build/fp64_peak scripts/fp64_peak.cu
// Prints one JSON line.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

#define CK(x)                                                                         \
    do {                                                                              \
        cudaError_t e_ = (x);                                                         \
        if (e_ != cudaSuccess) {                                                      \
            fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_));                  \
            exit(1);                                                                  \
        }                                                                             \
    } while (0)

__global__ void __launch_bounds__(256) dfma_kernel(double* out, int iters, double a, double b) {
    double acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = threadIdx.x + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = fma(acc[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += acc[i];
    if (s == 12345.678) out[0] = s;  // keep the chain alive
}

__global__ void __launch_bounds__(256) dmma_kernel(double* out, int iters, double a, double b) {
    double c[4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i) c[i][0] = c[i][1] = threadIdx.x * 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c[i][0]), "+d"(c[i][1])
                         : "d"(a), "d"(b));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) s += c[i][0] + c[i][1];
    if (s == 12345.678) out[0] = s;
}

__global__ void copy_kernel(const double2* __restrict__ in, double2* __restrict__ out, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) out[i] = in[i];
}

__global__ void __launch_bounds__(256) read_kernel(const double2* __restrict__ in, double* out, size_t n) {
    double s0 = 0, s1 = 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < n; i += 4 * stride) {
        double2 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(v[u].x), "=d"(v[u].y) : "l"(in + i + u * stride));
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            s0 += v[u].x;
            s1 += v[u].y;
        }
    }
    for (; i < n; i += stride) {
        s0 += in[i].x;
        s1 += in[i].y;
    }
    if (s0 + s1 == 12345.678) out[0] = s0;
}

template <class F>
static float time_ms(F f, int reps) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    f();
    f();
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        CK(cudaEventRecord(e0));
        f();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float t;
        CK(cudaEventElapsedTime(&t, e0, e1));
        if (t < best) best = t;
    }
    return best;
}

int main() {
    cudaDeviceProp p;
    CK(cudaGetDeviceProperties(&p, 0));
    const int sms = p.multiProcessorCount;
    double* d;
    CK(cudaMalloc(&d, 1 << 20));
    const int iters = 1 << 14, ctas = sms * 8;
    const float t_fma = time_ms([&] { dfma_kernel<<<ctas, 256>>>(d, iters, 1.0000001, 1e-9); }, 5);
    const double fma_flops = 2.0 * 8 * iters * 256.0 * ctas;
    const float t_mma = time_ms([&] { dmma_kernel<<<ctas, 256>>>(d, iters, 1.0000001, 1e-9); }, 5);
    const double mma_flops = 2.0 * 8 * 8 * 4 * 4.0 * iters * (256 / 32) * ctas;  // m8n8k4 = 512 flops per warp instruction
    const size_t n = (size_t)1 << 28;  // 4 GiB of double2 per buffer
    double2 *a, *b;
    CK(cudaMalloc(&a, n * sizeof(double2)));
    CK(cudaMalloc(&b, n * sizeof(double2)));
    CK(cudaMemset(a, 0, n * sizeof(double2)));
    CK(cudaMemset(b, 0, n * sizeof(double2)));
    const float t_copy = time_ms([&] { copy_kernel<<<sms * 16, 512>>>(a, b, n); }, 5);
    const float t_read = time_ms([&] { read_kernel<<<sms * 16, 256>>>(a, d, n); }, 5);
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"dfma_tflops\": %.2f, \"dmma_tflops\": %.2f, \"copy_gbs\": %.0f, \"read_gbs\": %.0f}\n",
           p.name, sms, fma_flops / t_fma / 1e9, mma_flops / t_mma / 1e9, 2.0 * n * 16 / t_copy / 1e6, 1.0 * n * 16 / t_read / 1e6);
    return 0;
}
