// sgbm_post.cu -- winner-take-all, uniqueness, sub-pixel, LR check,
// and the 3x3 median (A.4): the tail of cv2.StereoSGBM.compute
// (reference call site stereo_vision/stereo_vision.py:178).
#include "common.cuh"

template <int R>
__device__ __forceinline__ void load_regs_s(const uint16_t* p, uint32_t (&r)[R])
{
    if constexpr (R % 4 == 0) {
#pragma unroll
        for (int i = 0; i < R / 4; i++) {
            uint4 v = __ldg(reinterpret_cast<const uint4*>(p) + i);
            r[4 * i] = v.x; r[4 * i + 1] = v.y; r[4 * i + 2] = v.z; r[4 * i + 3] = v.w;
        }
    } else if constexpr (R % 2 == 0) {
#pragma unroll
        for (int i = 0; i < R / 2; i++) {
            uint2 v = __ldg(reinterpret_cast<const uint2*>(p) + i);
            r[2 * i] = v.x; r[2 * i + 1] = v.y;
        }
    } else {
#pragma unroll
        for (int i = 0; i < R; i++) r[i] = __ldg(reinterpret_cast<const uint32_t*>(p) + i);
    }
}

// ---------------------------------------------------------------------------
// WTA: one G-lane group per computed pixel.  Writes the pre-LR-check disparity and
// scatters (cost, x) into the per-row disp2 key array with atomicMin:
//   key = (minS + 0x8000) << 16 | (0xffff - x)   ->  lowest (signed) cost wins, ties go to the largest x,
// which is cv2's "strict > while x runs downward" rule, independent of thread order.
// ---------------------------------------------------------------------------
#define WTA_NT 256
template <int R, int G>
__global__ void __launch_bounds__(WTA_NT)
k_wta(const uint16_t* __restrict__ S, SgbmEff e, int16_t* __restrict__ disp, uint32_t* __restrict__ disp2key)
{
    const int lane = threadIdx.x % G;
    const long long gid = ((long long)blockIdx.x * WTA_NT + threadIdx.x) / G;
    const long long total = (long long)e.batch * e.H * e.width1;
    const bool active = gid < total;
    const long long pid = active ? gid : 0;
    const int x = (int)(pid % e.width1);
    const long long by = pid / e.width1;               // b*H + y

    uint32_t v[R];
    load_regs_s<R>(S + (size_t)pid * e.Dp + (size_t)lane * 2 * R, v);

    const int dbase = lane * 2 * R;
    // S cells are SIGNED int16: they are >= 0 everywhere except in MODE_SGBM_3WAY's scalar-tail case (sgbm_agg.cu TM = 1),
    // where one cell of the down path can go negative, as it does in cv2.  Keys order signed values through the 0x8000 bias.
    // MODE_SGBM_3WAY: disparities below simdD are searched on cv2's 8 SIMD lanes (b2s_tie8_*), the rest (numDisparities % 8
    // != 0 only) by a scalar tail that continues with a strict compare; every other mode: one strict search from d = 0
    const int simdD = e.way3 ? e.simdD : e.D;
    uint32_t key = 0xffffffffu, keyt = 0xffffffffu;
#pragma unroll
    for (int i = 0; i < R; i++) {
        int d0 = dbase + 2 * i;
        uint32_t k0 = (((v[i] & 0xffffu) ^ 0x8000u) << 16) | (uint32_t)d0;
        uint32_t k1 = ((v[i] & 0xffff0000u) ^ 0x80000000u) | (uint32_t)(d0 + 1);
        if (d0 < simdD) key = min(key, k0); else if (d0 < e.D) keyt = min(keyt, k0);
        if (d0 + 1 < simdD) key = min(key, k1); else if (d0 + 1 < e.D) keyt = min(keyt, k1);
    }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) {
        key = min(key, __shfl_xor_sync(0xffffffffu, key, o));
        keyt = min(keyt, __shfl_xor_sync(0xffffffffu, keyt, o));
    }
    if (key == 0xffffffffu) key = 0xffff0000u;         // no SIMD part (numDisparities < 8): cv2 starts from (SHRT_MAX, 0)
    int best = (int)(key & 0xffffu);
    int minS = (int)(key >> 16) - 0x8000;
    if (e.way3) {                              // MODE_SGBM_3WAY: a tied minimum goes to cv2's 8-lane winner (common.cuh)
        bool tie = false;
#pragma unroll
        for (int i = 0; i < R; i++) {
            const int d0 = dbase + 2 * i;
            tie = tie || ((int)(int16_t)(v[i] & 0xffffu) == minS && d0 != best && d0 < simdD) ||
                  ((int)(int16_t)(v[i] >> 16) == minS && d0 + 1 != best && d0 + 1 < simdD);
        }
        if (__any_sync(0xffffffffu, tie)) {
            unsigned long long w8 = 0ull;
#pragma unroll
            for (int i = 0; i < R; i++) {
                const int d0 = dbase + 2 * i;
                if ((int)(int16_t)(v[i] & 0xffffu) == minS && d0 < simdD) w8 = b2s_tie8_merge(w8, b2s_tie8_one(d0));
                if ((int)(int16_t)(v[i] >> 16) == minS && d0 + 1 < simdD) w8 = b2s_tie8_merge(w8, b2s_tie8_one(d0 + 1));
            }
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) w8 = b2s_tie8_merge(w8, __shfl_xor_sync(0xffffffffu, w8, o));
            if (w8) best = b2s_tie8_best(w8);
        }
    }
    if (keyt != 0xffffffffu && (int)(keyt >> 16) - 0x8000 < minS) {       // the tail wins only strictly
        minS = (int)(keyt >> 16) - 0x8000;
        best = (int)(keyt & 0xffffu);
    }

    // uniqueness + neighbours of the winner
    const int mul = 100 - e.uniq;
    // S * mul < minS * 100 ; MODE_SGBM_3WAY (uqle): S < (int16)(floor(minS * 100 / mul) + 1), cv2's 16-bit compare, wrapping
    const int thr = minS * 100;
    const int thr16 = (e.uqle && mul > 0) ? (int)(int16_t)(uint16_t)((thr / mul + 1) & 0xffff) : 0;
    uint32_t bad = (minS >= 32767 && !e.way3) ? 1u : 0u;       // every S saturated: cv2 finds no winner (bestDisp = -1 -> INVALID)
    int sm1 = 0, sp1 = 0;
#pragma unroll
    for (int i = 0; i < R; i++) {
#pragma unroll
        for (int h = 0; h < 2; h++) {
            int d = dbase + 2 * i + h;
            int val = h ? (int)(int16_t)(v[i] >> 16) : (int)(int16_t)(v[i] & 0xffffu);
            if (d < e.D) {
                int ad = d - best;
                // (3WAY's scalar tail d >= simdD uses the exact 32-bit rule of the other modes)
                if ((ad > 1 || ad < -1) && ((e.uqle && d < simdD) ? val < thr16 : (val * mul < thr && (!e.way3 || e.uniq > 0)))) bad = 1;
                if (ad == -1) sm1 = val;
                if (ad == 1) sp1 = val;
            }
        }
    }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) {
        bad |= __shfl_xor_sync(0xffffffffu, bad, o);
        sm1 += __shfl_xor_sync(0xffffffffu, sm1, o);       // exactly one lane holds each neighbour, the others 0
        sp1 += __shfl_xor_sync(0xffffffffu, sp1, o);
    }
    if (!active || lane != 0) return;
    int16_t* drow = disp + (size_t)by * e.W;
    if (bad) { drow[x + e.minX1] = (int16_t)e.invalid; return; }
    int x2 = x + e.minX1 - best - e.minD;
    // cv2: disp2cost starts at SHRT_MAX and is only replaced by a strictly smaller cost
    if (minS < 32767) atomicMin(disp2key + (size_t)by * e.W + x2, ((uint32_t)(minS + 0x8000) << 16) | (uint32_t)(0xffff - x));
    int dq = best * 16;
    if (best > 0 && best < e.D - 1) {
        int denom2 = max(sm1 + sp1 - 2 * minS, 1);
        dq += ((sm1 - sp1) * 16 + denom2) / (denom2 * 2);
    }
    drow[x + e.minX1] = (int16_t)(dq + e.minD * 16);
}

// LR consistency check of one pixel (second loop): true = keep d1.
__device__ __forceinline__ bool lr_consistent(const SgbmEff& e, const uint32_t* __restrict__ keyrow, int x, int d1)
{
    const int _d = d1 >> 4, d_ = (d1 + 15) >> 4;
    const int _x = x - _d, x_ = x - d_;
    if (!(0 <= _x && _x < e.W)) return true;
    {
        uint32_t k = keyrow[_x];
        // disp2 = best + minD of the winning x;  empty slots read as the scaled INVALID (cv2 quirk)
        int v = (k == 0xffffffffu) ? e.invalid : ((0xffff - (int)(k & 0xffffu)) + e.minX1 - _x);
        if (!(v >= e.minD && abs(v - _d) > e.lrmax)) return true;
    }
    if (!(0 <= x_ && x_ < e.W)) return true;
    {
        uint32_t k = keyrow[x_];
        int v = (k == 0xffffffffu) ? e.invalid : ((0xffff - (int)(k & 0xffffu)) + e.minX1 - x_);
        if (!(v >= e.minD && abs(v - d_) > e.lrmax)) return true;
    }
    return false;
}

// in place on the pre-check disparity; also fills the never-computed columns with INVALID.
__global__ void __launch_bounds__(256)
k_lrcheck(SgbmEff e, int16_t* __restrict__ disp, const uint32_t* __restrict__ disp2key)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= e.W) return;
    const size_t row = ((size_t)blockIdx.z * e.H + blockIdx.y) * e.W;
    if (x < e.minX1 || x >= e.maxX1) { disp[row + x] = (int16_t)e.invalid; return; }
    const int d1 = disp[row + x];
    if (d1 == e.invalid) return;
    if (!lr_consistent(e, disp2key + row, x, d1)) disp[row + x] = (int16_t)e.invalid;
}

// Tail of the fused WTA (sgbm_wave): record -> sub-pixel disparity -> LR check -> raw map.
// rec.x = minS | best << 16 (best = 0xffff: rejected by the uniqueness test), rec.y = S(best-1) | S(best+1) << 16
__device__ __forceinline__ int finish_pixel(const SgbmEff& e, const uint2 r, const uint32_t* __restrict__ keyrow, int x)
{
    if (x < e.minX1 || x >= e.maxX1) return e.invalid;
    const int best = (int)(r.x >> 16);
    if (best == 0xffff) return e.invalid;
    const int minS = (int)(r.x & 0xffffu), sm1 = (int)(r.y & 0xffffu), sp1 = (int)(r.y >> 16);
    int dq = best * 16;
    if (best > 0 && best < e.D - 1) {
        int denom2 = max(sm1 + sp1 - 2 * minS, 1);
        dq += ((sm1 - sp1) * 16 + denom2) / (denom2 * 2);
    }
    const int d1 = dq + e.minD * 16;
    return lr_consistent(e, keyrow, x, d1) ? d1 : e.invalid;
}

__global__ void __launch_bounds__(256)
k_finish_rec(SgbmEff e, const uint2* __restrict__ rec, const uint32_t* __restrict__ disp2key, int16_t* __restrict__ disp)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= e.W) return;
    const size_t row = ((size_t)blockIdx.z * e.H + blockIdx.y) * e.W;
    uint2 r = make_uint2(0xffff0000u, 0u);
    if (x >= e.minX1 && x < e.maxX1) r = rec[row + x];
    disp[row + x] = (int16_t)finish_pixel(e, r, disp2key + row, x);
}

// two pixels per thread: one 16-byte record load, one 32-bit store (even W)
__global__ void __launch_bounds__(256)
k_finish_rec_x2(SgbmEff e, const uint2* __restrict__ rec, const uint32_t* __restrict__ disp2key, int16_t* __restrict__ disp)
{
    const int x = 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    if (x >= e.W) return;
    const size_t row = ((size_t)blockIdx.z * e.H + blockIdx.y) * e.W;
    uint4 r = make_uint4(0xffff0000u, 0u, 0xffff0000u, 0u);
    if (x + 1 >= e.minX1 && x < e.maxX1) r = *reinterpret_cast<const uint4*>(rec + row + x);
    const int o0 = finish_pixel(e, make_uint2(r.x, r.y), disp2key + row, x);
    const int o1 = finish_pixel(e, make_uint2(r.z, r.w), disp2key + row, x + 1);
    *reinterpret_cast<uint32_t*>(disp + row + x) = ((uint32_t)o0 & 0xffffu) | ((uint32_t)o1 << 16);
}

int b2s_launch_finish_rec(const SgbmEff& e, const uint2* rec, const uint32_t* disp2key, int16_t* disp_raw, cudaStream_t st)
{
    if ((e.W & 1) == 0) {
        dim3 grid(b2s_cdiv(e.W / 2, 256), e.H, e.batch);
        k_finish_rec_x2<<<grid, 256, 0, st>>>(e, rec, disp2key, disp_raw);
    } else {
        dim3 grid(b2s_cdiv(e.W, 256), e.H, e.batch);
        k_finish_rec<<<grid, 256, 0, st>>>(e, rec, disp2key, disp_raw);
    }
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

// MODE_SGBM_3WAY (cv2 4.13.0 computeDisparity3WAY): stripe st writes its source row y to row (st ? 0 : ov) + y - s0 of a
// buffer of sz + ov rows preset to INVALID, and image row i is copied from row ov + i % sz of stripe i / sz.  Unless the
// early start of a stripe was clamped at the image top (tiny images) that is the stripe's own row i.
__global__ void __launch_bounds__(256)
k_assemble_stripes(SgbmEff v, const int16_t* __restrict__ src, int16_t* __restrict__ dst)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= v.W) return;
    const int i = blockIdx.y, b = blockIdx.z;
    const int st = i / v.sz;
    const int s0 = b2s_stripe_start(v, st);
    const int s1 = min((st + 1) * v.sz, v.Hsrc);
    const int y = s0 + v.ov + (i - st * v.sz) - (st == 0 ? v.ov : 0);
    int val = v.invalid;
    if (y >= s0 && y < s1 && y - s0 < v.H) val = src[(((size_t)b * v.nst + st) * v.H + (y - s0)) * v.W + x];
    dst[((size_t)b * v.Hsrc + i) * v.W + x] = (int16_t)val;
}

int b2s_launch_assemble_stripes(const SgbmEff& v, const int16_t* raw_stripes, int16_t* raw, cudaStream_t st)
{
    dim3 grid(b2s_cdiv(v.W, 256), v.Hsrc, v.batch / v.nst);
    k_assemble_stripes<<<grid, 256, 0, st>>>(v, raw_stripes, raw);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

template <int R, int G>
static int launch_wta_rg(const uint16_t* S, const SgbmEff& e, int16_t* disp, uint32_t* disp2key, cudaStream_t st)
{
    long long groups = (long long)e.batch * e.H * e.width1;
    k_wta<R, G><<<b2s_cdiv(groups * G, WTA_NT), WTA_NT, 0, st>>>(S, e, disp, disp2key);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

int b2s_launch_wta(const uint16_t* S, const SgbmEff& e, int16_t* disp_raw, uint32_t* disp2key, cudaStream_t st)
{
    B2S_CUDA_CHECK(cudaMemsetAsync(disp2key, 0xff, (size_t)e.batch * e.H * e.W * sizeof(uint32_t), st));
    int rc = B2S_E_UNSUPPORTED;
#define WTA_CASE(RR, GG) if (e.R == RR && e.G == GG) rc = launch_wta_rg<RR, GG>(S, e, disp_raw, disp2key, st);
    WTA_CASE(12, 32) WTA_CASE(16, 32)          // D > 512; smaller D runs the fused wavefront WTA, except
    WTA_CASE(2, 32) WTA_CASE(4, 32) WTA_CASE(8, 32)   // MODE_SGBM_3WAY with numDisparities % 8 != 0
#undef WTA_CASE
    if (rc != B2S_OK) {
        if (rc == B2S_E_UNSUPPORTED) b2s_set_error("no WTA kernel for R=%d G=%d", e.R, e.G);
        return rc;
    }
    dim3 grid(b2s_cdiv(e.W, 256), e.H, e.batch);
    k_lrcheck<<<grid, 256, 0, st>>>(e, disp_raw, disp2key);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

// ---------------------------------------------------------------------------
// 3x3 median on int16, replicate border, out of place (cv2.medianBlur(disp, 3)).
// 19-exchange median-of-9 network.
// ---------------------------------------------------------------------------
#define MSORT(a, b) { int _t = min(a, b); b = max(a, b); a = _t; }
__global__ void __launch_bounds__(256)
k_median3(const int16_t* __restrict__ src, int H, int W, int16_t* __restrict__ dst)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    if (x >= W) return;
    const size_t base = (size_t)blockIdx.z * H * W;
    const int xm = max(x - 1, 0), xp = min(x + 1, W - 1);
    const int16_t* r0 = src + base + (size_t)max(y - 1, 0) * W;
    const int16_t* r1 = src + base + (size_t)y * W;
    const int16_t* r2 = src + base + (size_t)min(y + 1, H - 1) * W;
    int p0 = r0[xm], p1 = r0[x], p2 = r0[xp];
    int p3 = r1[xm], p4 = r1[x], p5 = r1[xp];
    int p6 = r2[xm], p7 = r2[x], p8 = r2[xp];
    MSORT(p1, p2) MSORT(p4, p5) MSORT(p7, p8) MSORT(p0, p1) MSORT(p3, p4) MSORT(p6, p7)
    MSORT(p1, p2) MSORT(p4, p5) MSORT(p7, p8) MSORT(p0, p3) MSORT(p5, p8) MSORT(p4, p7)
    MSORT(p3, p6) MSORT(p1, p4) MSORT(p2, p5) MSORT(p4, p7) MSORT(p4, p2) MSORT(p6, p4)
    MSORT(p4, p2)
    dst[base + (size_t)y * W + x] = (int16_t)p4;
}

// two horizontally adjacent pixels per thread as packed s16x2 (VIMNMX.S16x2): half the exchanges, 32-bit
// loads and stores.  Needs 4-byte aligned rows (even W).
#define MSORT2(a, b) { uint32_t _t = __vmins2(a, b); b = __vmaxs2(a, b); a = _t; }
__global__ void __launch_bounds__(256)
k_median3_x2(const int16_t* __restrict__ src, int H, int W, int16_t* __restrict__ dst)
{
    const int x = 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    const int y = blockIdx.y;
    if (x >= W) return;
    const size_t base = (size_t)blockIdx.z * H * W;
    const uint32_t* r[3] = { reinterpret_cast<const uint32_t*>(src + base + (size_t)max(y - 1, 0) * W),
                             reinterpret_cast<const uint32_t*>(src + base + (size_t)y * W),
                             reinterpret_cast<const uint32_t*>(src + base + (size_t)min(y + 1, H - 1) * W) };
    const int k = x >> 1;
    uint32_t p[9];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const uint32_t a = r[i][k];                                   // (x, x+1)
        const uint32_t l = x >= 2 ? r[i][k - 1] : __byte_perm(a, a, 0x1010);      // (.., x-1), replicate at the border
        const uint32_t q = x + 2 < W ? r[i][k + 1] : __byte_perm(a, a, 0x3232);   // (x+2, ..)
        p[3 * i] = __byte_perm(l, a, 0x5432);                         // (x-1, x)
        p[3 * i + 1] = a;
        p[3 * i + 2] = __byte_perm(a, q, 0x5432);                     // (x+1, x+2)
    }
    MSORT2(p[1], p[2]) MSORT2(p[4], p[5]) MSORT2(p[7], p[8]) MSORT2(p[0], p[1]) MSORT2(p[3], p[4]) MSORT2(p[6], p[7])
    MSORT2(p[1], p[2]) MSORT2(p[4], p[5]) MSORT2(p[7], p[8]) MSORT2(p[0], p[3]) MSORT2(p[5], p[8]) MSORT2(p[4], p[7])
    MSORT2(p[3], p[6]) MSORT2(p[1], p[4]) MSORT2(p[2], p[5]) MSORT2(p[4], p[7]) MSORT2(p[4], p[2]) MSORT2(p[6], p[4])
    MSORT2(p[4], p[2])
    reinterpret_cast<uint32_t*>(dst + base + (size_t)y * W)[k] = p[4];
}

int b2s_launch_median3(const int16_t* src, int H, int W, int batch, int16_t* dst, cudaStream_t st)
{
    if ((W & 1) == 0 && (((uintptr_t)src | (uintptr_t)dst) & 3) == 0) {
        dim3 grid(b2s_cdiv(W / 2, 256), H, batch);
        k_median3_x2<<<grid, 256, 0, st>>>(src, H, W, dst);
        B2S_LAUNCH_CHECK();
        return B2S_OK;
    }
    dim3 grid(b2s_cdiv(W, 256), H, batch);
    k_median3<<<grid, 256, 0, st>>>(src, H, W, dst);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}


