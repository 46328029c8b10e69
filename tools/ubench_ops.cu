// ubench_ops.cu -- throughput of the integer / packed / half2 min-max instructions the aggregation recurrence can be
// built from, alone and mixed, in warp-instructions per cycle and SM.   nvcc -arch=sm_100a -O3 -o ubench_ops ubench_ops.cu
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

template <int OP> __device__ __forceinline__ uint32_t op(uint32_t a, uint32_t b, uint32_t c)
{
    if (OP == 0) return __vimin3_u16x2(a, b, c);                       // VIMNMX3.U16x2
    if (OP == 1) return __viaddmin_u16x2(a, b, c);                     // VIADDMNMX.U16x2
    if (OP == 2) return __vminu2(a, b);                                // VIMNMX.U16x2
    if (OP == 3) return a + b + c;                                     // IADD3
    if (OP == 4) { __half2 r = __hmin2(*reinterpret_cast<__half2*>(&a), *reinterpret_cast<__half2*>(&b)); return *reinterpret_cast<uint32_t*>(&r); }   // HMNMX2
    if (OP == 5) return min(a, b);                                     // VIMNMX.U32
    if (OP == 6) return a * b + c;                                     // IMAD
    if (OP == 7) return __byte_perm(a, b, 0x5432);                     // PRMT
    return a;
}

// 16 independent chains per thread, N iterations; MIX: chain i uses OPA if i is even, OPB if odd
template <int OPA, int OPB>
__global__ void __launch_bounds__(256) k(uint32_t* out, int n, uint32_t seed)
{
    uint32_t v[16];
#pragma unroll
    for (int i = 0; i < 16; i++) v[i] = seed * (i + 1) + threadIdx.x;
    const uint32_t b = seed ^ 0x01230123u, c = seed + 0x00110011u;
    for (int it = 0; it < n; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++) v[i] = (i & 1) ? op<OPB>(v[i], b, c) : op<OPA>(v[i], b, c);
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) s ^= v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OPA, int OPB> void run(const char* name, uint32_t* out)
{
    const int n = 4096, blocks = 148 * 8, threads = 256;
    k<OPA, OPB><<<blocks, threads>>>(out, 16, 12345u);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<OPA, OPB><<<blocks, threads>>>(out, n, 12345u);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    int clk;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double winst = (double)blocks * (threads / 32) * 16.0 * n;
    printf("%-34s %7.3f ms  %6.2f warp-inst/clk/SM (at %d MHz)\n", name, ms, winst / (ms * 1e-3 * clk * 1e3) / 148.0, clk / 1000);
}

int main()
{
    uint32_t* out;
    cudaMalloc(&out, 148 * 8 * 256 * 4);
    run<0, 0>("VIMNMX3.U16x2", out);
    run<1, 1>("VIADDMNMX.U16x2", out);
    run<2, 2>("VIMNMX.U16x2 (2 inputs)", out);
    run<5, 5>("VIMNMX.U32", out);
    run<3, 3>("IADD3", out);
    run<7, 7>("PRMT", out);
    run<6, 6>("IMAD", out);
    run<4, 4>("HMNMX2 (min.f16x2)", out);
    run<0, 4>("VIMNMX3.U16x2 + HMNMX2", out);
    run<0, 3>("VIMNMX3.U16x2 + IADD3", out);
    run<0, 6>("VIMNMX3.U16x2 + IMAD", out);
    run<4, 6>("HMNMX2 + IMAD", out);
    run<4, 3>("HMNMX2 + IADD3", out);
    return 0;
}
