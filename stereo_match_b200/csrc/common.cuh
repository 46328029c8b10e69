// common.cuh -- shared declarations of libb2stereo (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/b2stereo.h"

// Effective SGBM parameters after cv2's silent fallbacks plus the
// volume geometry.  Passed by value to kernels.
struct SgbmEff {
    int H, W, batch;
    int minD, D, Dp;          // Dp: padded disparity stride (u16 cells) of cost/S volumes
    int SW2;                  // block half size
    int P1, P2, uniq, lrmax, ftzero;
    int minX1, maxX1, width1;
    int invalid;              // (minD-1)*16
    int ndirs;                // 5 (MODE_SGBM), 8 (MODE_HH) or 4 (MODE_HH4)
    int G;                    // lanes per pixel in the aggregation/WTA kernels (8, 16 or 32)
    int R;                    // packed u16x2 registers per lane: Dp = 2*R*G
    int swz;                  // 1: [Dp] records of the cost/S volumes use the permuted cell order of b2s_cell_pos
    // --- where the cost kernels read their descriptors (defaults: cn = 1, Hsrc = H, nst = 1) ---
    int cn;                   // colour channels of the input (1 or 3): the descriptor arrays hold (pairs * cn) gray
                              // images, channel c of pair p at index p * cn + c; the BT costs of the channels add up
    int Hsrc;                 // rows of the source images
    int nst, sz, ov;          // MODE_SGBM_3WAY: the volumes hold batch = pairs * nst "stripe images" of H rows; stripe
                              // st of a pair starts at source row b2s_stripe_start() and clamps its block sum there
    int uqle;                 // 1: the uniqueness test rejects on equality as well (MODE_SGBM_3WAY with a ratio > 0), in
                              // 16 bits: S < (int16)(floor(100 * minS / (100 - ratio)) + 1), the threshold wrapping
    int simdD;                // MODE_SGBM_3WAY with D % 8 != 0: cv2 runs d < simdD = 8 * (D / 8) on SIMD lanes and the rest in
                              // a scalar tail with its own rules (sgbm_agg.cu TM, k_wta); D otherwise
    int ov_raw;               // MODE_SGBM_3WAY's stripe overlap takes blockSize AS GIVEN: blockSize / 2 + 1 (C division)
    int way3;                 // 1: MODE_SGBM_3WAY's winner rule (b2s_tie8_*) ; 0: first minimum, and a pixel whose S are
                              // all saturated (minS = 32767) has no winner at all (cv2 starts from bestDisp = -1)
};

#ifdef __CUDACC__
// MODE_SGBM_3WAY of cv2 4.13.0 searches the minimum on 8 int16 SIMD lanes: lane j keeps the LAST minimum among its
// disparities d = j (mod 8), the lanes are then merged to the SMALLEST position (probed; oracle/stereo_oracle.c
// wta_rows_ex).  One byte per lane class: (d >> 3) + 1 of the last tied minimum (0: none), d < 1024.
__device__ __forceinline__ unsigned long long b2s_tie8_one(int d)
{
    return (unsigned long long)((d >> 3) + 1) << ((d & 7) * 8);
}
__device__ __forceinline__ unsigned long long b2s_tie8_merge(unsigned long long a, unsigned long long b)
{
    const unsigned lo = __vmaxu4((unsigned)a, (unsigned)b), hi = __vmaxu4((unsigned)(a >> 32), (unsigned)(b >> 32));
    return ((unsigned long long)hi << 32) | lo;
}
__device__ __forceinline__ int b2s_tie8_best(unsigned long long w)
{
    int best = 1 << 30;
#pragma unroll
    for (int c = 0; c < 8; c++) {
        const int v = (int)((w >> (8 * c)) & 0xffull);
        if (v) best = min(best, ((v - 1) << 3) | c);
    }
    return best;
}
#endif

// first source row of stripe st (cv2 4.13.0 computeDisparity3WAY: range.start * stripe_sz - stripe_overlap, clamped)
__host__ __device__ inline int b2s_stripe_start(const SgbmEff& e, int st)
{
    if (e.nst == 1) return 0;
    int s = st * e.sz - e.ov;
    s = s < e.Hsrc ? s : e.Hsrc;
    return s > 0 ? s : 0;
}

// aggregation direction table (step from predecessor to pixel), same order as the
// oracle's DIRS: -> , down-right, down, down-left, <- , up-right, up, up-left
__host__ __device__ inline void b2s_dir(int k, int& dx, int& dy)
{
    const int DX[8] = { 1, 1, 0, -1, -1, 1, 0, -1 };
    const int DY[8] = { 0, 1, 1, 1, 0, -1, -1, -1 };
    dx = DX[k]; dy = DY[k];
}

// Position (u16 index) of disparity cell d inside a [Dp] record.  Natural order unless `swz` (R = 8: lane j of the
// wavefront kernel owns cells 16j .. 16j+15).  Then
//   * register i of the lane packs the cells (16j + i, 16j + 8 + i): the neighbours d-1 / d+1 of BOTH halves are the
//     neighbouring registers, so the recurrence needs no shifted copies of L (sgbm_wave_impl.cuh, step());
//   * registers 0-3 and 4-7 form two 16-byte chunks, and chunk c of lane j is stored at chunk slot c*G + j, which
//     makes every group-wide vector access contiguous.
__host__ __device__ inline int b2s_cell_pos(int d, int R, int G, int swz)
{
    if (!swz) return d;
    (void)R;
    const int j = d >> 4, k = d & 15, h = k >> 3, i = k & 7, c = i >> 2, w = i & 3;
    return ((((c * G + j) << 2) + w) << 1) | h;
}

void b2s_set_error(const char* fmt, ...);
void b2s_count_launch(int n = 1);
int b2s_poison(void* ws, size_t bytes, cudaStream_t st);   // debug aid: B2S_POISON (api.cu)

#define B2S_CUDA_CHECK(expr)                                                          \
    do {                                                                              \
        cudaError_t _e = (expr);                                                      \
        if (_e != cudaSuccess) {                                                      \
            b2s_set_error("CUDA error %s at %s:%d: %s", cudaGetErrorName(_e),        \
                          __FILE__, __LINE__, cudaGetErrorString(_e));                \
            return B2S_E_CUDA;                                                        \
        }                                                                             \
    } while (0)

#define B2S_LAUNCH_CHECK() do { b2s_count_launch(); B2S_CUDA_CHECK(cudaGetLastError()); } while (0)

static inline int b2s_cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }

// ---- stage launchers (each in its own .cu) ---------------------------------
// descriptors: per image pixel 2 words: (u, lo, hi, 0) for the Sobel channel and the raw channel
int b2s_launch_descriptors(const uint8_t* img, const SgbmEff& e, uint2* desc, cudaStream_t st);
size_t b2s_desc_right_bytes(const SgbmEff& e);
int b2s_launch_descriptors_right(const uint8_t* img, const SgbmEff& e, void* desc, cudaStream_t st);
int b2s_launch_cost(const uint2* descL, const void* descR, const SgbmEff& e, uint16_t* cost,
                    int* overflow_flag, cudaStream_t st);
// first!=0: S = L_r (plain store); otherwise S = min(32767, S + L_r)
int b2s_launch_aggregate(const uint16_t* cost, uint16_t* S, const SgbmEff& e, int dir, int first,
                         cudaStream_t st);
int b2s_launch_wta(const uint16_t* S, const SgbmEff& e, int16_t* disp_raw, uint32_t* disp2key,
                   cudaStream_t st);
int b2s_launch_finish_rec(const SgbmEff& e, const uint2* rec, const uint32_t* disp2key, int16_t* disp_raw, cudaStream_t st);
// MODE_SGBM_3WAY: rows of the stripe images (v = volume geometry) -> rows of the image, cv2's buffer offsets included
int b2s_launch_assemble_stripes(const SgbmEff& v, const int16_t* raw_stripes, int16_t* raw, cudaStream_t st);
int b2s_launch_median3(const int16_t* src, int H, int W, int batch, int16_t* dst, cudaStream_t st);
int b2s_launch_speckles(int16_t* img, int H, int W, int batch, int newVal, int maxSize, int maxDiff,
                        int32_t* labels, int32_t* counts, cudaStream_t st);
