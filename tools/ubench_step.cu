// Micro-benchmark of the SGBM recurrence step (register only): cycles per (pixel, direction) step per warp.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_step ubench_step.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int G>
__device__ __forceinline__ uint32_t group_min(uint32_t v)
{
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// variant A: as in sgbm_agg.cu today
template <int R, int G>
__device__ __forceinline__ uint32_t stepA(uint32_t (&L)[R], const uint32_t (&C)[R], int lane, uint32_t p1x2,
                                          uint32_t deltax2, uint32_t minx2)
{
    uint32_t up = __shfl_up_sync(0xffffffffu, L[R - 1], 1, G);
    uint32_t dn = __shfl_down_sync(0xffffffffu, L[0], 1, G);
    if (lane == 0) up = 0x7fff0000u;
    if (lane == G - 1) dn = 0x00007fffu;
    uint32_t mn = 0xffffffffu, prev = up;
#pragma unroll
    for (int i = 0; i < R; i++) {
        uint32_t cur = L[i];
        uint32_t next = (i + 1 < R) ? L[i + 1] : dn;
        uint32_t dm1 = __byte_perm(prev, cur, 0x5432);
        uint32_t dp1 = __byte_perm(cur, next, 0x5432);
        uint32_t m = __vminu2(dm1, dp1);
        uint32_t t = __viaddmin_u16x2(m, p1x2, deltax2);
        uint32_t u = __vminu2(cur, t);
        uint32_t l = C[i] + u - minx2;
        prev = cur;
        L[i] = l;
        mn = __vminu2(mn, l);
    }
    return mn;
}

// variant B: one PRMT per register (dp1(i) == dm1(i+1)), L+P1 precomputed, 3-input min
template <int R, int G>
__device__ __forceinline__ uint32_t stepB(uint32_t (&L)[R], const uint32_t (&C)[R], int lane, uint32_t p1x2,
                                          uint32_t deltax2, uint32_t minx2)
{
    uint32_t up = __shfl_up_sync(0xffffffffu, L[R - 1], 1, G);
    uint32_t dn = __shfl_down_sync(0xffffffffu, L[0], 1, G);
    if (lane == 0) up = 0x7fff0000u;
    if (lane == G - 1) dn = 0x00007fffu;
    // sh[i] = (d-1, d) pair for register i, sh[R] = (d+1,d+2) of the last one ; all + P1
    uint32_t sh[R + 1];
    sh[0] = __byte_perm(up, L[0], 0x5432);
#pragma unroll
    for (int i = 1; i < R; i++) sh[i] = __byte_perm(L[i - 1], L[i], 0x5432);
    sh[R] = __byte_perm(L[R - 1], dn, 0x5432);
    uint32_t mn = 0xffffffffu;
    const uint32_t dmp = deltax2 - p1x2;       // delta - P1 >= 0 since P2 > P1
#pragma unroll
    for (int i = 0; i < R; i++) {
        uint32_t m = __vimin3_u16x2(sh[i], sh[i + 1], dmp);     // min(L(d-1), L(d+1), delta-P1)
        uint32_t u = __viaddmin_u16x2(m, p1x2, L[i]);           // min(m + P1, L(d))
        uint32_t l = C[i] + u - minx2;
        L[i] = l;
        mn = __vminu2(mn, l);
    }
    return mn;
}

template <int VAR, int NDIR>
__global__ void __launch_bounds__(256) k(const uint32_t* __restrict__ cin, uint32_t* out, int nsteps, long long* cyc)
{
    constexpr int R = 8, G = 8;
    const int lane = threadIdx.x % G;
    uint32_t L[NDIR][R], C[R], S[R];
    uint32_t minx2[NDIR];
#pragma unroll
    for (int k2 = 0; k2 < NDIR; k2++) { minx2[k2] = 0;
#pragma unroll
        for (int i = 0; i < R; i++) L[k2][i] = 0; }
#pragma unroll
    for (int i = 0; i < R; i++) C[i] = cin[(threadIdx.x * R + i) & 1023];
    const uint32_t p1 = 600 * 0x10001u, p2 = 2400 * 0x10001u;
    long long t0 = clock64();
    for (int s = 0; s < nsteps; s++) {
#pragma unroll
        for (int i = 0; i < R; i++) { C[i] = (C[i] * 5 + 0x00030007u) & 0x0fff0fffu; S[i] = 0; }   // stand-in for the C load
#pragma unroll
        for (int k2 = 0; k2 < NDIR; k2++) {
            uint32_t mn = VAR == 0 ? stepA<R, G>(L[k2], C, lane, p1, minx2[k2] + p2, minx2[k2])
                                   : stepB<R, G>(L[k2], C, lane, p1, minx2[k2] + p2, minx2[k2]);
#pragma unroll
            for (int i = 0; i < R; i++) S[i] = __viaddmin_u16x2(S[i], L[k2][i], 0x7fff7fffu);
            uint32_t m16 = min(mn & 0xffffu, mn >> 16);
            minx2[k2] = group_min<G>(m16) * 0x10001u;
        }
        if (s == nsteps - 1) {
#pragma unroll
            for (int i = 0; i < R; i++) out[(blockIdx.x * 256 + threadIdx.x) * R + i] = S[i];
        }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int VAR, int NDIR>
void run(const char* name, int blocks_per_sm)
{
    uint32_t *cin, *out; long long* cyc;
    int nb = 148 * blocks_per_sm, nsteps = 2000;
    cudaMalloc(&cin, 4096); cudaMemset(cin, 1, 4096);
    cudaMalloc(&out, (size_t)nb * 256 * 8 * 4); cudaMalloc(&cyc, 8);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<VAR, NDIR><<<nb, 256>>>(cin, out, 10, cyc);
    cudaEventRecord(e0);
    k<VAR, NDIR><<<nb, 256>>>(cin, out, nsteps, cyc);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    // cells processed: nb*256 threads * 16 cells * NDIR * nsteps
    double cells = (double)nb * 256 * 16 * NDIR * nsteps;
    printf("%-10s ndir=%d blocks/SM=%d: %.3f ms, %.1f Gcell-dir/s, %.1f cycles/step(block0), => %.1f us per KITTI pair for 8 dirs\n",
           name, NDIR, blocks_per_sm, ms, cells / ms / 1e6, (double)c / nsteps, 53.5e6 * 8 / (cells / ms / 1e3) );
    cudaFree(cin); cudaFree(out); cudaFree(cyc);
}

int main()
{
    for (int bps : {1, 2, 4}) {
        run<0, 1>("A", bps); run<1, 1>("B", bps);
        run<0, 4>("A", bps); run<1, 4>("B", bps);
    }
    return 0;
}
