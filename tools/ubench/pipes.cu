// Micro-benchmark: which pipe executes packed 16-bit min/max on sm_100a?  (tools only; not part of the product)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu && ./pipes
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t hmin2u(uint32_t a, uint32_t b) { uint32_t r; asm volatile("min.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
__device__ __forceinline__ uint32_t hmax2u(uint32_t a, uint32_t b) { uint32_t r; asm volatile("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
__device__ __forceinline__ uint32_t vmin2u(uint32_t a, uint32_t b) { uint32_t r; asm volatile("min.u16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
__device__ __forceinline__ uint32_t vmax2u(uint32_t a, uint32_t b) { uint32_t r; asm volatile("max.u16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }

template <int MODE>
__global__ void __launch_bounds__(512, 1) k(uint32_t *out, int iters, uint32_t seed)
{
    uint32_t v[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = ((seed * (threadIdx.x + 1) * (i + 3)) & 0x3fff3fffu);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int s = 1; s < 16; s <<= 1) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                if ((i & s) == 0) {
                    const uint32_t a = v[i], b = v[i | s];
                    if (MODE == 0) { v[i] = vmin2u(a, b); v[i | s] = vmax2u(a, b); }
                    if (MODE == 1) { v[i] = hmin2u(a, b); v[i | s] = hmax2u(a, b); }
                    if (MODE == 2) { v[i] = vmin2u(a, b); v[i | s] = hmax2u(a, b); }
                }
            }
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = (v[i] + 0x00010001u * (uint32_t)it) & 0x3fff3fffu;   // keep it from collapsing
    }
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) x ^= v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = x;
}

// FP64 add chain latency/throughput and F2F conversions
template <int MODE>
__global__ void __launch_bounds__(512, 1) kd(double *out, int iters, double seed)
{
    double a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = seed * (threadIdx.x + i + 1);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) a[i] = __dadd_rn(a[i], 1.25);                       // DADD
            if (MODE == 1) a[i] = (double)__double2float_rn(__dadd_rn(a[i], 1.25));   // DADD + F2F + F2F
            if (MODE == 2) { const double m = __hiloint2double((__double2hiint(a[i]) & 0x7ff00000) + 0x01d00000, 0);
                             a[i] = __dsub_rn(__dadd_rn(__dadd_rn(a[i], 1.25), m), m); }
        }
    }
    double x = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) x += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = x;
}

template <typename F> float timeit(F f)
{
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}

int main()
{
    uint32_t *o; cudaMalloc(&o, 148 * 512 * 8);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const int iters = 20000;
    const double inst = 64.0 * iters;       // min/max instructions per thread (4 stages x 8 pairs x 2), + 32 fixups not counted
    const char *names[3] = {"VIMNMX.U16x2 min+max", "HMNMX2 min+max", "min=VIMNMX max=HMNMX2"};
    for (int m = 0; m < 3; ++m) {
        float ms = m == 0 ? timeit([&] { k<0><<<148, 512>>>(o, iters, 12345u); })
                 : m == 1 ? timeit([&] { k<1><<<148, 512>>>(o, iters, 12345u); })
                          : timeit([&] { k<2><<<148, 512>>>(o, iters, 12345u); });
        // per SMSP: 4 warps x inst instructions in ms
        const double cyc = ms * 1e-3 * clk * 1e3;
        printf("%-24s %.3f ms  -> %.3f cycles per warp-instruction per SMSP (incl. 32 fix-up ALU ops per 64)\n", names[m], ms, cyc / (4.0 * inst));
    }
    const char *dn[3] = {"DADD x8 indep", "DADD+F2F+F2F x8", "DADD+magic round x8"};
    for (int m = 0; m < 3; ++m) {
        float ms = m == 0 ? timeit([&] { kd<0><<<148, 512>>>((double*)o, 5000, 1.5); })
                 : m == 1 ? timeit([&] { kd<1><<<148, 512>>>((double*)o, 5000, 1.5); })
                          : timeit([&] { kd<2><<<148, 512>>>((double*)o, 5000, 1.5); });
        const double cyc = ms * 1e-3 * clk * 1e3;
        printf("%-24s %.3f ms  -> %.2f cycles per chain step per SMSP (4 warps x 8 chains)\n", dn[m], ms, cyc / (4.0 * 8 * 5000));
    }
    // single warp per SMSP latency
    for (int m = 0; m < 3; ++m) {
        float ms = m == 0 ? timeit([&] { kd<0><<<148, 128>>>((double*)o, 5000, 1.5); })
                 : m == 1 ? timeit([&] { kd<1><<<148, 128>>>((double*)o, 5000, 1.5); })
                          : timeit([&] { kd<2><<<148, 128>>>((double*)o, 5000, 1.5); });
        const double cyc = ms * 1e-3 * clk * 1e3;
        printf("1 warp/SMSP %-20s -> %.2f cycles per 8-chain round\n", dn[m], cyc / 5000);
    }
    return 0;
}
