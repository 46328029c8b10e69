// sgbm_cost.cu -- pixel descriptors, Birchfield-Tomasi cost and block sum.
// Replaces the calcPixelCostBT / block-sum stage inside cv2.StereoSGBM.compute
// (reference call site: stereo_vision/stereo_vision.py:178).
//
// Layout in HBM
//   desc   [batch][H][W] uint2 : .x = Sobel channel word, .y = raw channel word,
//                                 word = value | lo<<8 | hi<<16 (half-pixel interval), top byte 0
//   cost   [batch][H][width1][Dp] uint16 : raw block cost (cv2's C minus P2); padded cells
//                                 d >= D hold 0 (the aggregation kernel pins their L_r at 0x7fff)
#include "common.cuh"
#include <cuda.h>
#include <cstdlib>
#include <type_traits>

// ---------------------------------------------------------------------------
// descriptors
// ---------------------------------------------------------------------------
__device__ __forceinline__ int sobel_clip(const uint8_t* c, const uint8_t* n, const uint8_t* s,
                                          int x, int W, int ftzero)
{
    if (x <= 0 || x >= W - 1) return ftzero;
    int v = 2 * ((int)c[x + 1] - (int)c[x - 1]) + ((int)n[x + 1] - (int)n[x - 1]) +
            ((int)s[x + 1] - (int)s[x - 1]);
    return min(max(v, -ftzero), ftzero) + ftzero;
}

__device__ __forceinline__ uint32_t pack_desc(int pm, int p, int pp, bool has_l, bool has_r)
{
    int pl = has_l ? (p + pm) >> 1 : p;
    int pr = has_r ? (p + pp) >> 1 : p;
    int lo = min(min(pl, pr), p), hi = max(max(pl, pr), p);
    return (uint32_t)p | ((uint32_t)lo << 8) | ((uint32_t)hi << 16);
}

// (Sobel value, raw value) of the pixels x0 + dir*i, i in [-1, n+1], into sv[i + 1] as g | r << 8: the pixel's own
// two values are computed once and its neighbours' come from shared memory.  x outside the image -> 0.
__device__ __forceinline__ void row_values(const uint8_t* base, int x0, int dir, int n, int y, int H, int W, int ftzero, uint16_t* sv)
{
    const uint8_t* c = base + (size_t)y * W;
    const uint8_t* nn = base + (size_t)(y > 0 ? y - 1 : y) * W;
    const uint8_t* ss = base + (size_t)(y < H - 1 ? y + 1 : y) * W;
    for (int i = (int)threadIdx.x - 1; i <= n + 1; i += (int)blockDim.x) {
        const int x = x0 + dir * i;
        uint32_t v = 0;
        if (x >= 0 && x < W) {
            const int g = sobel_clip(c, nn, ss, x, W, ftzero);
            const int r = (x <= 0 || x >= W - 1) ? ftzero : c[x];   // raw channel: the two border columns are the constant ftzero
            v = (uint32_t)g | ((uint32_t)r << 8);
        }
        sv[i + 1] = (uint16_t)v;
    }
}

// descriptor of pixel x from its own values and those of x-1 (vm) and x+1 (vp)
__device__ __forceinline__ uint2 desc_from(uint32_t vm, uint32_t v, uint32_t vp, int x, int W)
{
    uint2 o;
    o.x = pack_desc((int)(vm & 0xffu), (int)(v & 0xffu), (int)(vp & 0xffu), x > 0, x < W - 1);
    o.y = pack_desc((int)(vm >> 8), (int)(v >> 8), (int)(vp >> 8), x > 0, x < W - 1);
    return o;
}

__global__ void __launch_bounds__(256)
k_descriptors(const uint8_t* __restrict__ img, int H, int W, int ftzero, uint2* __restrict__ desc)
{
    __shared__ uint16_t sv[256 + 4];
    const int x0 = blockIdx.x * 256;
    const int y = blockIdx.y;
    const size_t b = blockIdx.z;
    row_values(img + b * (size_t)H * W, x0, 1, 256, y, H, W, ftzero, sv);
    __syncthreads();
    const int x = x0 + threadIdx.x;
    if (x >= W) return;
    desc[b * (size_t)H * W + (size_t)y * W + x] = desc_from(sv[threadIdx.x], sv[threadIdx.x + 1], sv[threadIdx.x + 2], x, W);
}

int b2s_launch_descriptors(const uint8_t* img, const SgbmEff& e, uint2* desc, cudaStream_t st)
{
    dim3 grid(b2s_cdiv(e.W, 256), e.H, e.batch);
    k_descriptors<<<grid, 256, 0, st>>>(img, e.H, e.W, e.ftzero, desc);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

// Right-image descriptors for the fast cost kernel: per image row twelve u16 planes of PW elements,
//   plane 2p   [ri]     = component p of pixel x   (p: Sobel value, lo, hi, raw value, lo, hi)
//   plane 2p+1 [ri - 1] = the same                 (copy shifted by one element)
// in REVERSED x order, ri = B2S_RPAD + (W - 1 - x), zero outside the image.  The operand pair (x-d, x-d-1)
// of a packed disparity pair is then one aligned 32-bit word of one of the two copies, and the window of a
// CTA is twelve contiguous 16-byte aligned runs: one bulk copy (TMA) each, no unpacking in the cost kernel.
#define B2S_RPAD 8
static inline int plane_pitch(int W) { return (W + B2S_RPAD + 328 + 7) & ~7; }

#define B2S_PLANE_EPB 512      // plane elements per block
__global__ void __launch_bounds__(256)
k_descriptor_planes(const uint8_t* __restrict__ img, int H, int W, int ftzero, int PW, uint16_t* __restrict__ planes)
{
    // a block = 512 consecutive plane elements ri0 .. ri0+511 (x descending) of one image row
    constexpr int N = B2S_PLANE_EPB;
    __shared__ uint16_t sv[N + 6];
    __shared__ uint2 sd[N + 2];
    const int ri0 = blockIdx.x * N;
    const int y = blockIdx.y;
    const size_t b = blockIdx.z;
    const int xfirst = W - 1 - (ri0 - B2S_RPAD);
    row_values(img + b * (size_t)H * W, xfirst, -1, N + 1, y, H, W, ftzero, sv);     // elements -1 .. N+2
    __syncthreads();
    for (int i = threadIdx.x; i <= N; i += 256) {                                    // element ri0+N feeds the shifted copy
        const int x = xfirst - i;
        // in reversed order the element before is x+1 and the one after is x-1
        sd[i] = (x >= 0 && x < W && ri0 + i < PW) ? desc_from(sv[i + 2], sv[i + 1], sv[i], x, W) : make_uint2(0u, 0u);
    }
    __syncthreads();
    // thread k owns word k of all twelve plane copies: copy A = elements (2k, 2k+1), copy B = elements (2k+1, 2k+2).
    // The top byte of a descriptor word is 0, so one PRMT builds a packed pair of u16 values.
    const int k = threadIdx.x;
    if (ri0 + 2 * k >= PW) return;
    const uint2 a = sd[2 * k], m = sd[2 * k + 1], c = sd[2 * k + 2];
    uint32_t* row = reinterpret_cast<uint32_t*>(planes + ((b * H + y) * 12) * (size_t)PW) + (ri0 >> 1) + k;
    const size_t pw2 = (size_t)(PW >> 1);
    row[0] = __byte_perm(a.x, m.x, 0x7430);          row[pw2] = __byte_perm(m.x, c.x, 0x7430);
    row[2 * pw2] = __byte_perm(a.x, m.x, 0x7531);    row[3 * pw2] = __byte_perm(m.x, c.x, 0x7531);
    row[4 * pw2] = __byte_perm(a.x, m.x, 0x7632);    row[5 * pw2] = __byte_perm(m.x, c.x, 0x7632);
    row[6 * pw2] = __byte_perm(a.y, m.y, 0x7430);    row[7 * pw2] = __byte_perm(m.y, c.y, 0x7430);
    row[8 * pw2] = __byte_perm(a.y, m.y, 0x7531);    row[9 * pw2] = __byte_perm(m.y, c.y, 0x7531);
    row[10 * pw2] = __byte_perm(a.y, m.y, 0x7632);   row[11 * pw2] = __byte_perm(m.y, c.y, 0x7632);
}

// The same for the interleaved records (b2s_cell_pos, swz): a packed cell pair is (d, d + 8), so its right-image
// operands are the pixels x - d and x - d - 8: six u32 planes, element ri = component(ri) | component(ri + 8) << 16.
__global__ void __launch_bounds__(256)
k_descriptor_planes8(const uint8_t* __restrict__ img, int H, int W, int ftzero, int PW, uint32_t* __restrict__ planes)
{
    constexpr int N = B2S_PLANE_EPB;
    __shared__ uint16_t sv[N + 8 + 4];
    __shared__ uint2 sd[N + 8];
    const int ri0 = blockIdx.x * N;
    const int y = blockIdx.y;
    const size_t b = blockIdx.z;
    const int xfirst = W - 1 - (ri0 - B2S_RPAD);
    row_values(img + b * (size_t)H * W, xfirst, -1, N + 8, y, H, W, ftzero, sv);     // elements -1 .. N+9
    __syncthreads();
    for (int i = threadIdx.x; i < N + 8; i += 256) {
        const int x = xfirst - i;
        sd[i] = (x >= 0 && x < W && ri0 + i < PW) ? desc_from(sv[i + 2], sv[i + 1], sv[i], x, W) : make_uint2(0u, 0u);
    }
    __syncthreads();
    uint32_t* row = planes + ((b * H + y) * 6) * (size_t)PW + ri0;
#pragma unroll
    for (int j = 0; j < N / 256; j++) {
        const int k = threadIdx.x + 256 * j;
        if (ri0 + k >= PW) break;
        const uint2 a = sd[k], c = sd[k + 8];                 // the top byte of a descriptor word is 0
        row[k] = __byte_perm(a.x, c.x, 0x7430);               row[(size_t)PW + k] = __byte_perm(a.x, c.x, 0x7531);
        row[2 * (size_t)PW + k] = __byte_perm(a.x, c.x, 0x7632);
        row[3 * (size_t)PW + k] = __byte_perm(a.y, c.y, 0x7430);   row[4 * (size_t)PW + k] = __byte_perm(a.y, c.y, 0x7531);
        row[5 * (size_t)PW + k] = __byte_perm(a.y, c.y, 0x7632);
    }
}

// The same once more for k_cost_fast2, PACKED: every descriptor component is 8 bits wide, so one u32 holds two components
// of the pixel pair (ri, ri + 8) -- three planes instead of six, half the shared-memory wavefronts in the cost kernel,
// which unpacks a word into two u16x2 operands with one PRMT each:
//   plane 0 = (g.v, g.v8, g.lo, g.lo8)   plane 1 = (g.hi, g.hi8, r.v, r.v8)   plane 2 = (r.lo, r.lo8, r.hi, r.hi8)
// (g: Sobel channel, r: raw channel, v/lo/hi: value and half-pixel interval, "8": the pixel eight columns to the left)
__global__ void __launch_bounds__(256)
k_descriptor_planes8b(const uint8_t* __restrict__ img, int H, int W, int ftzero, int PW, uint32_t* __restrict__ planes)
{
    constexpr int N = B2S_PLANE_EPB;
    __shared__ uint16_t sv[N + 8 + 4];
    __shared__ uint2 sd[N + 8];
    const int ri0 = blockIdx.x * N;
    const int y = blockIdx.y;
    const size_t b = blockIdx.z;
    const int xfirst = W - 1 - (ri0 - B2S_RPAD);
    row_values(img + b * (size_t)H * W, xfirst, -1, N + 8, y, H, W, ftzero, sv);
    __syncthreads();
    for (int i = threadIdx.x; i < N + 8; i += 256) {
        const int x = xfirst - i;
        sd[i] = (x >= 0 && x < W && ri0 + i < PW) ? desc_from(sv[i + 2], sv[i + 1], sv[i], x, W) : make_uint2(0u, 0u);
    }
    __syncthreads();
    uint32_t* row = planes + ((b * H + y) * 3) * (size_t)PW + ri0;
#pragma unroll
    for (int j = 0; j < N / 256; j++) {
        const int k = threadIdx.x + 256 * j;
        if (ri0 + k >= PW) break;
        const uint2 a = sd[k], c = sd[k + 8];                 // descriptor word = value | lo << 8 | hi << 16
        row[k] = __byte_perm(a.x, c.x, 0x5140);
        row[(size_t)PW + k] = __byte_perm(__byte_perm(a.x, c.x, 0x0062), __byte_perm(a.y, c.y, 0x0040), 0x5410);
        row[2 * (size_t)PW + k] = __byte_perm(a.y, c.y, 0x6251);
    }
}

// ---------------------------------------------------------------------------
// BT cost of a disparity pair (d0, d0+1) in packed u16x2, biased by 0x0100 per half so
// that plain 32-bit adds/subs never borrow across halves.
//   wl : left descriptor word at x            (same for both halves)
//   wa : right descriptor word at x - d0      (low half)
//   wb : right descriptor word at x - d0 - 1  (high half)
// returns packed cost (0..255 per half)
// ---------------------------------------------------------------------------
#define B2S_BIAS 0x01000100u
__device__ __forceinline__ uint32_t bt_pair(uint32_t wl, uint32_t wa, uint32_t wb)
{
    uint32_t u = __byte_perm(wl, 0, 0x4040), lo1 = __byte_perm(wl, 0, 0x4141), hi1 = __byte_perm(wl, 0, 0x4242);
    uint32_t v = __byte_perm(wa, wb, 0x3430), lo2 = __byte_perm(wa, wb, 0x3531), hi2 = __byte_perm(wa, wb, 0x3632);
    uint32_t A = u + B2S_BIAS - hi2;       // u - hi2   (biased)
    uint32_t Bv = lo2 + B2S_BIAS - u;      // lo2 - u
    uint32_t Cv = v + B2S_BIAS - hi1;      // v - hi1
    uint32_t Dv = lo1 + B2S_BIAS - v;      // lo1 - v
    uint32_t c0 = __vimax3_u16x2(A, Bv, B2S_BIAS);
    uint32_t c1 = __vimax3_u16x2(Cv, Dv, B2S_BIAS);
    return __vminu2(c0, c1) - B2S_BIAS;
}

// Generic cost kernel (any blockSize, Dp <= 2048).
// One CTA: column strip [x0, x0+TW) of the cropped axis, row band [yb, yb+TR), all Dp.
// Marches down the band; per row: pixel costs for TW+2*SW2 columns -> smem, horizontal
// sum, vertical running sum through a ring of blockSize horizontal sums in smem.
#define COST_NT 512
#define COST_IPT 2   // items (x, dpair) per thread
__global__ void __launch_bounds__(COST_NT)
k_cost(const uint2* __restrict__ descL, const uint2* __restrict__ descR, SgbmEff e, int TW, int TR,
       uint16_t* __restrict__ cost, int* __restrict__ overflow_flag)
{
    extern __shared__ uint32_t smem[];
    const int DP2 = e.Dp >> 1;
    const int bs = 2 * e.SW2 + 1;
    const int PW = TW + 2 * e.SW2;                 // columns of pixel costs per row
    uint32_t* pix = smem;                          // [PW][DP2]
    uint32_t* ring = smem + (size_t)PW * DP2;      // [bs][TW][DP2]
    const int x0 = blockIdx.x * TW;
    const int yb = blockIdx.y * TR;
    const size_t b = blockIdx.z;
    const int yend = min(yb + TR, e.H);
    // source images of this (stripe) image: channel c of pair bsrc, rows from s0 on
    const size_t bsrc = b / e.nst;
    const int s0 = b2s_stripe_start(e, (int)(b % e.nst));
    const size_t plane = (size_t)e.Hsrc * e.W;
    const uint2* dL = descL + bsrc * e.cn * plane;
    const uint2* dR = descR + bsrc * e.cn * plane;
    uint16_t* out = cost + b * (size_t)e.H * e.width1 * e.Dp;
    const int nitems = TW * DP2;

    uint32_t acc[COST_IPT];
#pragma unroll
    for (int i = 0; i < COST_IPT; i++) acc[i] = 0;
    for (int i = threadIdx.x; i < bs * nitems; i += COST_NT) ring[i] = 0;
    uint32_t vmax = 0;

    int slot = 0;
    for (int yv = yb - e.SW2; yv < yend + e.SW2; yv++) {
        const int yr = min(s0 + max(yv, 0), e.Hsrc - 1);     // clamped at the stripe's first row and at the image's last
        __syncthreads();                           // previous row's pix fully consumed
        for (int it = threadIdx.x; it < PW * DP2; it += COST_NT) {
            int xh = it / DP2, dp = it - xh * DP2;
            int xc = min(max(x0 - e.SW2 + xh, 0), e.width1 - 1);   // clamp on the cropped axis
            int xi = xc + e.minX1;
            int d0 = 2 * dp;
            uint32_t pc = 0;
            if (d0 < e.D) {
                for (int c = 0; c < e.cn; c++) {
                    const uint2* rowL = dL + c * plane + (size_t)yr * e.W;
                    const uint2* rowR = dR + c * plane + (size_t)yr * e.W;
                    uint2 wl = __ldg(rowL + xi);
                    int xr = xi - d0 - e.minD;
                    uint2 wa = __ldg(rowR + xr);
                    uint2 wb = (d0 + 1 < e.D) ? __ldg(rowR + xr - 1) : wa;
                    uint32_t cg = bt_pair(wl.x, wa.x, wb.x);
                    uint32_t cr = bt_pair(wl.y, wa.y, wb.y);
                    pc += cg + ((cr >> 2) & 0x003f003fu);
                }
            }
            pix[it] = pc;
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < COST_IPT; i++) {
            int it = threadIdx.x + i * COST_NT;
            if (it < nitems) {
                int xl = it / DP2, dp = it - xl * DP2;
                uint32_t hs = 0;
                for (int k = 0; k < bs; k++) hs += pix[(xl + k) * DP2 + dp];
                uint32_t* rp = ring + (size_t)slot * nitems + it;
                acc[i] += hs - *rp;                // halves never borrow: acc >= old contribution
                *rp = hs;
                // a half can only wrap past 0xffff after first being seen >= 0x8000 (one row
                // changes it by < 0x8000), so tracking the running max catches every overflow
                vmax = __vmaxu2(vmax, acc[i]);
                int yo = yv - e.SW2;
                int x = x0 + xl;
                if (yo >= yb && x < e.width1) {
                    uint32_t v = acc[i];
                    int d0 = 2 * dp;
                    if (d0 >= e.D) v = 0u;
                    else if (d0 + 1 >= e.D) v = v & 0xffffu;
                    uint16_t* rec = out + ((size_t)yo * e.width1 + x) * e.Dp;
                    if (e.swz) {                      // interleaved records: d0 and d0 + 1 are not neighbours in memory
                        rec[b2s_cell_pos(d0, e.R, e.G, 1)] = (uint16_t)(v & 0xffffu);
                        rec[b2s_cell_pos(d0 + 1, e.R, e.G, 1)] = (uint16_t)(v >> 16);
                    } else {
                        *reinterpret_cast<uint32_t*>(rec + d0) = v;
                    }
                }
            }
        }
        slot = slot + 1 == bs ? 0 : slot + 1;
    }
    uint32_t m = max(vmax & 0xffffu, vmax >> 16);
    // Guaranteed-parity domain: cv2's SIMD lanes evaluate C + min(...) with saturating int16
    // adds before subtracting delta, its scalar tail does not.  Both agree (and agree with this
    // engine) while (raw + P2) + (minL + P2) <= 32767; minL <= raw_max + P2 bounds it.
    if (2 * (int)m + 3 * e.P2 > 32767) atomicOr(overflow_flag, 1);
}

// ---------------------------------------------------------------------------
// Fast cost kernel (blockSize <= 11, Dp <= 256).
// CTA = 32 columns of pixel costs (32 - (BS-1) output columns) x all Dp disparities, marching
// down a band of rows.  Warp w owns disparities [8w, 8w+8); lane l owns column l.
//   stage   : left descriptor words of the row -> smem; right descriptors -> six u16 planes
//             (value, lo, hi of the Sobel and of the raw channel) in REVERSED x order, each in
//             two copies offset by one element, so that the pair (x-d, x-d-1) is one aligned
//             32-bit word: packed u16x2 operands come straight out of LDS with no byte shuffles
//   phase A : BT cost of (column, 8 disparities) -> pix[column][d] (smem, u16x2)
//   phase B : horizontal BS-sum from smem, vertical running sum through a ring of BS rows held
//             in registers (the row loop is unrolled by BS), one 16-byte store per thread
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t bt2(uint32_t u, uint32_t lo1, uint32_t hi1, uint32_t v, uint32_t lo2, uint32_t hi2)
{
    uint32_t A = u + B2S_BIAS - hi2;
    uint32_t Bv = lo2 + B2S_BIAS - u;
    uint32_t Cv = v + B2S_BIAS - hi1;
    uint32_t Dv = lo1 + B2S_BIAS - v;
    return __vminu2(__vimax3_u16x2(A, Bv, B2S_BIAS), __vimax3_u16x2(Cv, Dv, B2S_BIAS));   // biased by 0x0100
}

// sum of v over lanes [l, l + BS) (packed u16x2, no overflow: BS * 317 < 65536); lanes whose window
// leaves the warp return garbage (they are the halo columns and never store)
template <int BS>
__device__ __forceinline__ uint32_t window_sum(uint32_t v)
{
    uint32_t sum = 0, p = v;              // p = sum over a window of `len` lanes starting at this lane
    int off = 0;
#pragma unroll
    for (int len = 1; len <= BS; len <<= 1) {
        if (BS & len) {                   // take this power-of-two block at lane offset `off`
            sum += off ? __shfl_down_sync(0xffffffffu, p, off) : p;
            off += len;
        }
        if (2 * len <= BS) p += __shfl_down_sync(0xffffffffu, p, len);
    }
    return sum;
}

// ---- mbarrier / bulk-copy helpers (shared-memory addresses as 32-bit .shared offsets) ----
__device__ __forceinline__ void cmbar_init(unsigned bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void cmbar_arrive(unsigned bar)
{
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void cmbar_expect_tx(unsigned bar, unsigned bytes)
{
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cmbar_wait(unsigned bar, unsigned parity)
{
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, 0x989680;\n\t"
                 "@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void lds64(unsigned addr, uint32_t& a, uint32_t& b)
{
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(a), "=r"(b) : "r"(addr));
}
__device__ __forceinline__ uint32_t lds32(unsigned addr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}

#ifndef B2S_COST_NSTG
#define B2S_COST_NSTG 8
#endif
#ifndef B2S_COST_PRODUCER_LAST
#define B2S_COST_PRODUCER_LAST 1
#endif
template <int MAXW, int WIN = 32> struct CostPlan {
    static constexpr int NSTG = B2S_COST_NSTG;                            // rows in flight
    static constexpr int NELP = (WIN + 8 * MAXW + 16 + 63) / 64 * 64 + 32; // plane pitch in smem: copies land 64 B apart mod 128
    static constexpr int STAGE = 12 * NELP * 2;
    static constexpr int bars = NSTG * STAGE;                             // full[NSTG], empty[NSTG]
    static constexpr int total = bars + 2 * NSTG * 8;
};

// k_cost_fast2: 64 window columns, three packed u32 planes per image row
template <int MAXW> struct CostPlan2 {
    static constexpr int NSTG = B2S_COST_NSTG;
    static constexpr int NELP = (64 + 8 * MAXW + 16 + 63) / 64 * 64 + 32;   // u32 elements per staged plane
    static constexpr int STAGE = 3 * NELP * 4;
    static constexpr int bars = NSTG * STAGE;
    static constexpr int total = bars + 2 * NSTG * 8;
};

// P8: the records are the wavefront kernel's interleaved ones (b2s_cell_pos, swz): a packed pair is (d, d + 8), warp
// o owns the cells {16q + 4s + w, 16q + 4s + w + 8 : w < 4} with q = o / 2, s = o % 2, and the staged planes hold one
// u32 (value(x), value(x - 8)) per element instead of two shifted u16 copies.
// CN colour channels: the ring carries one stage per (row, channel), the BT costs of the channels add up before the
// box sum.  Stripe images (MODE_SGBM_3WAY, e.nst > 1) read their rows from the pair's source image, see SgbmEff.
template <int BS, int MAXW, bool PAD, bool P8, int CN>
__global__ void __launch_bounds__(32 * MAXW, (MAXW == 16 && BS <= 7) ? 2 : 1)
k_cost_fast(const uint2* __restrict__ descL, const __grid_constant__ CUtensorMap tmap, SgbmEff e, int TR,
            uint16_t* __restrict__ cost, int* __restrict__ overflow_flag)
{
    using P = CostPlan<MAXW>;
    constexpr int SW2 = BS / 2, TW = 32 - (BS - 1), NSTG = P::NSTG, NELP = P::NELP;
    extern __shared__ __align__(128) unsigned char smem_c[];
    const unsigned sm0 = (unsigned)__cvta_generic_to_shared(smem_c);
    const int Dp = e.Dp;
    const int lane = threadIdx.x & 31, oct = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int x0 = blockIdx.x * TW, yb = blockIdx.y * TR;
    const unsigned b = blockIdx.z;
    const int yend = min(yb + TR, e.H);
    const int nrows = yend - yb + 2 * SW2;                                    // rows yb-SW2 .. yend+SW2-1 enter the window
    // phase A geometry of this thread (row independent)
    const int xiA = min(max(x0 - SW2 + lane, 0), e.width1 - 1) + e.minX1;     // clamp on the cropped axis
    const int XRM = min(max(x0 - SW2 + 31, 0), e.width1 - 1) + e.minX1 - e.minD;   // image column of reversed element 0
    const int gbase = B2S_RPAD + (e.W - 1 - XRM);                             // its index in the plane rows
    const int shift = gbase & (P8 ? 3 : 7), gal = gbase - shift;              // bulk copies start 16-byte aligned
    const int dfirst = P8 ? 16 * (oct >> 1) + 4 * (oct & 1) : 8 * oct;        // first cell of this warp
    const int e0 = XRM - (xiA - e.minD - dfirst) + shift;                     // first element of this thread in a staged plane
    const int par = e0 & 1;
    const unsigned roff = P8 ? (unsigned)e0 * 4u : (unsigned)((par * NELP) * 2 + ((e0 - par) >> 1) * 4);
    // phase B geometry
    const int opos = b2s_cell_pos(dfirst, e.R, e.G, e.swz);                   // where this warp's four words live in the record
    const int xB = x0 + lane;
    const bool outB = lane < TW && xB < e.width1;
    const size_t ostride = (size_t)e.width1 * Dp;
    uint16_t* optr = cost + (((size_t)b * e.H + yb) * e.width1 + min(xB, e.width1 - 1)) * Dp + opos;   // running store pointer
    uint32_t pmask[4];
#pragma unroll
    for (int w = 0; w < 4; w++) {
        const int dlo = P8 ? dfirst + w : dfirst + 2 * w, dhi = P8 ? dlo + 8 : dlo + 1;
        pmask[w] = (dlo < e.D ? 0xffffu : 0u) | (dhi < e.D ? 0xffff0000u : 0u);
    }
    const unsigned bsrc = b / (unsigned)e.nst;
    const int s0 = b2s_stripe_start(e, (int)(b % (unsigned)e.nst));
    const unsigned plane = (unsigned)(e.Hsrc * e.W);
    // the stripe's first row is folded into the bases, so rows are clamped to [0, hrel] like those of a whole image
    const unsigned bimg = bsrc * CN * plane + (unsigned)(s0 * e.W);           // pairs*cn*H*W < 2^31 is checked on the host
    const unsigned rowbase = bsrc * CN * (unsigned)e.Hsrc + (unsigned)s0;
    const int hrel = e.Hsrc - 1 - s0;
    const int nitems = nrows * CN;                                            // ring items: (row, channel)
    auto src_row = [&](int r) { return min(max(yb - SW2 + r, 0), hrel); };    // relative to the stripe's first row
    auto left_word = [&](int q) {
        const int r = q / CN, c = q - r * CN;
        return __ldg(descL + (bimg + (unsigned)c * plane + (unsigned)(src_row(r) * e.W + xiA)));
    };
    // producer (thread 0): ONE 3-d tensor copy (TMA) per row -- box {8 elements, NELP/8, 12 planes} of the view
    // {8, PW/8, batch*H*12} of the plane rows -> stage r % NSTG, dense [12][NELP].  Before a stage is refilled every
    // warp must have handed it back (empty barrier of the row that used it last).
    const int gal8 = P8 ? gal >> 2 : gal >> 3;                                // in 16-byte units
    constexpr unsigned NPL = P8 ? 6u : 12u;                                   // plane rows per image row
    auto issue = [&](int q) {
        const int r = q / CN, c = q - r * CN;
        const int yr = src_row(r);
        const unsigned st = (unsigned)(q & (NSTG - 1));
        const unsigned full = sm0 + P::bars + st * 8, dst = sm0 + st * P::STAGE;
        if (q >= NSTG) cmbar_wait(sm0 + P::bars + (NSTG + st) * 8, (unsigned)((q / NSTG) - 1) & 1u);
        cmbar_expect_tx(full, (unsigned)P::STAGE);
        const int row = (int)((rowbase + (unsigned)(c * e.Hsrc + yr)) * NPL);
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(dst), "l"(&tmap), "r"(0), "r"(gal8), "r"(row), "r"(full) : "memory");
    };

    if (threadIdx.x == 0) {
        for (int i = 0; i < NSTG; i++) {
            cmbar_init(sm0 + P::bars + i * 8, 1);
            cmbar_init(sm0 + P::bars + (NSTG + i) * 8, nwarps);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const bool producer = threadIdx.x == (B2S_COST_PRODUCER_LAST ? blockDim.x - 32 : 0u);
    if (producer)
        for (int q = 0; q < NSTG - 1 && q < nitems; q++) issue(q);

    uint32_t acc[4] = { 0, 0, 0, 0 }, ring[BS][4], vmax = 0;
#pragma unroll
    for (int s = 0; s < BS; s++) { ring[s][0] = ring[s][1] = ring[s][2] = ring[s][3] = 0; }
    uint2 wl_next = left_word(0);

    for (int r0 = 0; r0 < nrows; r0 += BS) {
#pragma unroll
        for (int s = 0; s < BS; s++) {
            const int r = r0 + s;
            if (r >= nrows) break;                                            // uniform
            // ---- phase A: pixel costs of (column, 8 disparities), summed over the colour channels ----
            uint32_t hs[4], pc[4] = { 0, 0, 0, 0 };
#pragma unroll
            for (int c = 0; c < CN; c++) {
                const int q = r * CN + c;
                if (producer && q + NSTG - 1 < nitems) issue(q + NSTG - 1);      // keep NSTG-1 items ahead
                const uint2 wl = wl_next;
                if (q + 1 < nitems) wl_next = left_word(q + 1);
                const unsigned st = (unsigned)(q & (NSTG - 1));
                cmbar_wait(sm0 + P::bars + st * 8, (unsigned)(q / NSTG) & 1u);
                const unsigned rb = sm0 + st * P::STAGE + roff;
                const uint32_t ug = __byte_perm(wl.x, 0, 0x4040), log = __byte_perm(wl.x, 0, 0x4141), hig = __byte_perm(wl.x, 0, 0x4242);
                const uint32_t ur = __byte_perm(wl.y, 0, 0x4040), lor = __byte_perm(wl.y, 0, 0x4141), hir = __byte_perm(wl.y, 0, 0x4242);
#pragma unroll
                for (int w = 0; w < 4; w++) {
                    const uint32_t cg = bt2(ug, log, hig, lds32(rb + 4 * w), lds32(rb + NELP * 4 + 4 * w), lds32(rb + 2 * NELP * 4 + 4 * w));
                    const uint32_t cr = bt2(ur, lor, hir, lds32(rb + 3 * NELP * 4 + 4 * w), lds32(rb + 4 * NELP * 4 + 4 * w),
                                            lds32(rb + 5 * NELP * 4 + 4 * w));
                    // (cr - bias) >> 2 == ((cr >> 2) & 0x7f) - 0x40 per half
                    pc[w] += cg + ((cr >> 2) & 0x007f007fu) - (B2S_BIAS + 0x00400040u);
                }
                // this warp is done with the stage: hand it back to the producer
                __syncwarp();
                if (lane == 0) cmbar_arrive(sm0 + P::bars + (NSTG + st) * 8);
            }
            // ---- phase B: box sums.  The BS columns are the next BS-1 lanes of this warp: window sum
            // by doubling (p2 = x + x+1, p4 = p2 + p2(x+2), ...) through shuffles, no shared memory.
#pragma unroll
            for (int w = 0; w < 4; w++) hs[w] = window_sum<BS>(pc[w]);
            {
#pragma unroll
                for (int w = 0; w < 4; w++) {
                    acc[w] += hs[w] - ring[s][w];          // halves never borrow: acc >= old contribution
                    ring[s][w] = hs[w];
                }
                // a half can only wrap past 0xffff after first being seen >= 0x8000
                vmax = __vimax3_u16x2(vmax, __vmaxu2(acc[0], acc[1]), __vmaxu2(acc[2], acc[3]));
                if (r >= 2 * SW2) {                        // output row yb + r - 2*SW2
                    if (outB) {
                        uint4 o = make_uint4(acc[0], acc[1], acc[2], acc[3]);
                        if (PAD) { o.x &= pmask[0]; o.y &= pmask[1]; o.z &= pmask[2]; o.w &= pmask[3]; }
                        *reinterpret_cast<uint4*>(optr) = o;
                    }
                    optr += ostride;
                }
            }
        }
    }
    if (PAD) vmax &= (pmask[0] | pmask[1] | pmask[2] | pmask[3]);   // coarse: padded lanes never decide
    uint32_t m = max(vmax & 0xffffu, vmax >> 16);
    if (lane >= TW) m = 0;
    if (2 * (int)m + 3 * e.P2 > 32767) atomicOr(overflow_flag, 1);
}

// ---------------------------------------------------------------------------
// k_cost_fast2: the same kernel with TWO adjacent columns per thread (interleaved records only: P8, every cell real).
// A warp covers 64 window columns, lane l the columns 2l and 2l + 1.  What that buys (measured: three quarters of
// the 24 LDS.32 per thread and row cost 1.5 ms of 8.75, the window-sum shuffles 0.7 ms, per 128 KITTI pairs):
//   * the right-image operands of the two columns overlap in all but one word: 5 words per plane for 16 cell pairs
//     instead of 8, fetched as three aligned LDS.64 (lanes are 8 bytes apart: 256 contiguous bytes, no conflicts);
//   * the horizontal window sum works on column pairs q = p(2l) + p(2l+1):
//         s(2l) = q(l) + ... + q(l+h-1) + p(2(l+h)),  s(2l+1) = s(2l) - p(2l) + p(2(l+h)+1)      (h = BS/2)
//     h + 1 shuffles for two columns instead of 3 each (BS = 5);
//   * BS - 1 halo columns per 64 instead of per 32.
// Columns outside the cropped image (cv2 clamps the window on the cropped axis) are computed on whatever lies there
// and then replaced by the border column's pixel cost with shuffles, in the first and the last strip only.
// ---------------------------------------------------------------------------
// NH > 1 (Dp = 128 * NH): a strip's disparities are split over NH CTAs of 16 warps, blockIdx.y = band * NH + part
template <int BS, int MAXW, int CN, int NH>
__global__ void __launch_bounds__(32 * MAXW, 1)
k_cost_fast2(const uint2* __restrict__ descL, const __grid_constant__ CUtensorMap tmap, SgbmEff e, int TR,
             uint16_t* __restrict__ cost, int* __restrict__ overflow_flag)
{
    using P = CostPlan2<MAXW * NH>;
    constexpr int SW2 = BS / 2, TW = 64 - (BS - 1), NSTG = P::NSTG, NELP = P::NELP;
    extern __shared__ __align__(128) unsigned char smem_c[];
    const unsigned sm0 = (unsigned)__cvta_generic_to_shared(smem_c);
    const int Dp = e.Dp;
    const int lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    const int oct = (int)(threadIdx.x >> 5) + MAXW * (int)(blockIdx.y % NH);      // this warp's slot among the Dp / 8 of a strip
    const int x0 = blockIdx.x * TW, yb = (int)(blockIdx.y / NH) * TR;
    const unsigned b = blockIdx.z;
    const int yend = min(yb + TR, e.H);
    const int nrows = yend - yb + 2 * SW2;
    // window column c = 2*lane + s (s = 0, 1) is the cropped-axis column x0 - SW2 + c
    const int xa = x0 - SW2 + 2 * lane;                                       // slot 0's column (slot 1: xa + 1)
    const int clo = max(0, SW2 - x0), chi = min(63, e.width1 - 1 - (x0 - SW2));   // window columns inside the cropped image
    const bool border = clo > 0 || chi < 63;                                  // CTA-uniform
    // image column of reversed element 0: the last window column inside the image, plus a margin of two elements.  The
    // lane whose slot 0 is that column has its slot 1 one column outside: its e1 - par is then still >= 0, so the
    // clamp of roff below only ever touches lanes with BOTH columns outside (margin <= B2S_RPAD: the planes are padded)
    const int XRM = min(x0 - SW2 + 63, e.width1 - 1) + e.minX1 - e.minD + 2;
    const int gbase = B2S_RPAD + (e.W - 1 - XRM);
    const int shift = gbase & 3, gal = gbase - shift;                         // bulk copies start 16-byte aligned
    const int dfirst = 16 * (oct >> 1) + 4 * (oct & 1);                       // first cell of this warp
    // staged element of (column, cell dfirst + w): e(c) + w, e(c + 1) = e(c) - 1.  The two columns need e1 .. e1 + 4
    // (e1 = slot 1's); three LDS.64 fetch the six words from the even index e1 - par on.  par is the same for all lanes.
    const int e1 = shift + XRM - (xa + 1 + e.minX1 - e.minD) + dfirst;
    const int par = (shift + XRM - (x0 - SW2 + 1 + e.minX1 - e.minD) + dfirst) & 1;
    const unsigned roff = (unsigned)max(e1 - par, 0) * 4u;                    // (columns beyond the image: clamped, replaced later)
    const int opos = b2s_cell_pos(dfirst, e.R, e.G, e.swz);
    const int xo = x0 + 2 * lane;                                             // output column of slot 0
    const bool outa = 2 * lane < TW && xo < e.width1, outb = 2 * lane + 1 < TW && xo + 1 < e.width1;
    const size_t ostride = (size_t)e.width1 * Dp;
    uint16_t* optr = cost + (((size_t)b * e.H + yb) * e.width1 + min(xo, e.width1 - 1)) * Dp + opos;
    const unsigned bsrc = b / (unsigned)e.nst;
    const int s0 = b2s_stripe_start(e, (int)(b % (unsigned)e.nst));
    const unsigned plane = (unsigned)(e.Hsrc * e.W);
    const unsigned bimg = bsrc * CN * plane + (unsigned)(s0 * e.W);
    const unsigned rowbase = bsrc * CN * (unsigned)e.Hsrc + (unsigned)s0;
    const int hrel = e.Hsrc - 1 - s0;
    const int nitems = nrows * CN;
    const int xla = min(max(xa + e.minX1, 0), e.W - 1), xlb = min(max(xa + 1 + e.minX1, 0), e.W - 1);   // left pixels (addresses kept inside the row)
    auto src_row = [&](int r) { return min(max(yb - SW2 + r, 0), hrel); };
    auto left_words = [&](int q, uint2& wa, uint2& wb) {
        const int r = q / CN, c = q - r * CN;
        const uint2* row = descL + (bimg + (unsigned)c * plane + (unsigned)(src_row(r) * e.W));
        wa = __ldg(row + xla); wb = __ldg(row + xlb);
    };
    const int gal8 = gal >> 2;                                                // in 16-byte units (arithmetic shift: may be negative, TMA zero-fills)
    auto issue = [&](int q) {
        const int r = q / CN, c = q - r * CN;
        const int yr = src_row(r);
        const unsigned st = (unsigned)(q & (NSTG - 1));
        const unsigned full = sm0 + P::bars + st * 8, dst = sm0 + st * P::STAGE;
        if (q >= NSTG) cmbar_wait(sm0 + P::bars + (NSTG + st) * 8, (unsigned)((q / NSTG) - 1) & 1u);
        cmbar_expect_tx(full, (unsigned)P::STAGE);
        const int row = (int)((rowbase + (unsigned)(c * e.Hsrc + yr)) * 3u);
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(dst), "l"(&tmap), "r"(0), "r"(gal8), "r"(row), "r"(full) : "memory");
    };
    if (threadIdx.x == 0) {
        for (int i = 0; i < NSTG; i++) {
            cmbar_init(sm0 + P::bars + i * 8, 1);
            cmbar_init(sm0 + P::bars + (NSTG + i) * 8, nwarps);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const bool producer = threadIdx.x == blockDim.x - 32;
    if (producer)
        for (int q = 0; q < NSTG - 1 && q < nitems; q++) issue(q);
    uint32_t acc[2][4], ring[BS][2][4], vmax = 0;
#pragma unroll
    for (int w = 0; w < 4; w++) {
        acc[0][w] = acc[1][w] = 0;
#pragma unroll
        for (int s = 0; s < BS; s++) ring[s][0][w] = ring[s][1][w] = 0;
    }
    uint2 wla_next, wlb_next;
    left_words(0, wla_next, wlb_next);
    // border strips: window column c takes the pixel cost of column min(max(c, clo), chi)
    const int ca = min(max(2 * lane, clo), chi), cb = min(max(2 * lane + 1, clo), chi);

    // BT cost of the two columns, both channels, against the three packed planes: PAR selects where the needed five
    // words sit among the six fetched ones (slot 1: PAR .. PAR + 3, slot 0: PAR + 1 .. PAR + 4); every word is
    // unpacked into two u16x2 operands (bytes 0, 1 and bytes 2, 3) with one PRMT each
    auto costs = [&](auto parc, unsigned base, const uint2 wordA, const uint2 wordB, uint32_t (&pa4)[4], uint32_t (&pb4)[4]) {
        constexpr int PAR = decltype(parc)::value;
        uint32_t W0[6], W1[6], W2[6];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            lds64(base + 8 * k, W0[2 * k], W0[2 * k + 1]);
            lds64(base + NELP * 4 + 8 * k, W1[2 * k], W1[2 * k + 1]);
            lds64(base + 2 * NELP * 4 + 8 * k, W2[2 * k], W2[2 * k + 1]);
        }
        const uint32_t uga = __byte_perm(wordA.x, 0, 0x4040), loga = __byte_perm(wordA.x, 0, 0x4141), higa = __byte_perm(wordA.x, 0, 0x4242);
        const uint32_t ura = __byte_perm(wordA.y, 0, 0x4040), lora = __byte_perm(wordA.y, 0, 0x4141), hira = __byte_perm(wordA.y, 0, 0x4242);
        const uint32_t ugb = __byte_perm(wordB.x, 0, 0x4040), logb = __byte_perm(wordB.x, 0, 0x4141), higb = __byte_perm(wordB.x, 0, 0x4242);
        const uint32_t urb = __byte_perm(wordB.y, 0, 0x4040), lorb = __byte_perm(wordB.y, 0, 0x4141), hirb = __byte_perm(wordB.y, 0, 0x4242);
        uint32_t cgp = 0, crp = 0;                         // slot 0 of word j uses the same plane words as slot 1 of word j + 1
#pragma unroll
        for (int j = 0; j < 5; j++) {
            const uint32_t w0 = W0[j + PAR], w1 = W1[j + PAR], w2 = W2[j + PAR];
            const uint32_t gv = __byte_perm(w0, 0, 0x4140), glo = __byte_perm(w0, 0, 0x4342), ghi = __byte_perm(w1, 0, 0x4140);
            const uint32_t rv = __byte_perm(w1, 0, 0x4342), rlo = __byte_perm(w2, 0, 0x4140), rhi = __byte_perm(w2, 0, 0x4342);
            // (cr - bias) >> 2 == ((cr >> 2) & 0x7f) - 0x40 per half
            if (j < 4) pb4[j] += bt2(ugb, logb, higb, gv, glo, ghi) + ((bt2(urb, lorb, hirb, rv, rlo, rhi) >> 2) & 0x007f007fu) - (B2S_BIAS + 0x00400040u);
            if (j > 0) pa4[j - 1] += bt2(uga, loga, higa, gv, glo, ghi) + ((bt2(ura, lora, hira, rv, rlo, rhi) >> 2) & 0x007f007fu) - (B2S_BIAS + 0x00400040u);
        }
        (void)cgp; (void)crp;
    };

    for (int r0 = 0; r0 < nrows; r0 += BS) {
#pragma unroll
        for (int s = 0; s < BS; s++) {
            const int r = r0 + s;
            if (r >= nrows) break;                                            // uniform
            // ---- phase A: pixel costs of (2 columns, 8 cells), summed over the colour channels ----
            uint32_t pa[4] = { 0, 0, 0, 0 }, pb[4] = { 0, 0, 0, 0 };
#pragma unroll
            for (int c = 0; c < CN; c++) {
                const int q = r * CN + c;
                if (producer && q + NSTG - 1 < nitems) issue(q + NSTG - 1);      // keep NSTG-1 items ahead
                const uint2 wa = wla_next, wb = wlb_next;
                if (q + 1 < nitems) left_words(q + 1, wla_next, wlb_next);
                const unsigned st = (unsigned)(q & (NSTG - 1));
                cmbar_wait(sm0 + P::bars + st * 8, (unsigned)(q / NSTG) & 1u);
                const unsigned rb = sm0 + st * P::STAGE + roff;
                if (par) costs(std::integral_constant<int, 1>{}, rb, wa, wb, pa, pb);
                else costs(std::integral_constant<int, 0>{}, rb, wa, wb, pa, pb);
                __syncwarp();
                if (lane == 0) cmbar_arrive(sm0 + P::bars + (NSTG + st) * 8);  // this warp is done with the stage
            }
            if (border) {
#pragma unroll
                for (int w = 0; w < 4; w++) {
                    const uint32_t a0 = __shfl_sync(0xffffffffu, pa[w], ca >> 1), a1 = __shfl_sync(0xffffffffu, pb[w], ca >> 1);
                    const uint32_t b0 = __shfl_sync(0xffffffffu, pa[w], cb >> 1), b1 = __shfl_sync(0xffffffffu, pb[w], cb >> 1);
                    pa[w] = (ca & 1) ? a1 : a0;
                    pb[w] = (cb & 1) ? b1 : b0;
                }
            }
            // ---- phase B: window sums over column pairs, vertical running sums, stores ----
#pragma unroll
            for (int w = 0; w < 4; w++) {
                uint32_t sa, sb;
                if (SW2 == 0) { sa = pa[w]; sb = pb[w]; }
                else {
                    const uint32_t q2 = pa[w] + pb[w];
                    uint32_t Q = q2;
#pragma unroll
                    for (int j = 1; j < SW2; j++) Q += __shfl_down_sync(0xffffffffu, q2, j);
                    sa = Q + __shfl_down_sync(0xffffffffu, pa[w], SW2);
                    sb = sa - pa[w] + __shfl_down_sync(0xffffffffu, pb[w], SW2);
                }
                acc[0][w] += sa - ring[s][0][w];           // halves never borrow: acc >= old contribution
                ring[s][0][w] = sa;
                acc[1][w] += sb - ring[s][1][w];
                ring[s][1][w] = sb;
            }
            // a half can only wrap past 0xffff after first being seen >= 0x8000
            vmax = __vimax3_u16x2(vmax, __vmaxu2(acc[0][0], acc[0][1]), __vmaxu2(acc[0][2], acc[0][3]));
            vmax = __vimax3_u16x2(vmax, __vmaxu2(acc[1][0], acc[1][1]), __vmaxu2(acc[1][2], acc[1][3]));
            if (r >= 2 * SW2) {                            // output row yb + r - 2*SW2
                if (outa) *reinterpret_cast<uint4*>(optr) = make_uint4(acc[0][0], acc[0][1], acc[0][2], acc[0][3]);
                if (outb) *reinterpret_cast<uint4*>(optr + Dp) = make_uint4(acc[1][0], acc[1][1], acc[1][2], acc[1][3]);
                optr += ostride;
            }
        }
    }
    uint32_t m = max(vmax & 0xffffu, vmax >> 16);
    if (2 * lane >= TW) m = 0;                             // lanes whose window leaves the warp hold garbage
    if (2 * (int)m + 3 * e.P2 > 32767) atomicOr(overflow_flag, 1);
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link against libcuda)
typedef CUresult (*tmap_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static tmap_encode_fn tmap_encoder()
{
    static tmap_encode_fn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess && qr == cudaDriverEntryPointSuccess)
            fn = (tmap_encode_fn)p;
    }
    return fn;
}

template <int BS, int MAXW, bool P8, int CN>
static int launch_cost_fast(const uint2* descL, const uint16_t* planesR, const SgbmEff& e, uint16_t* cost, int* flag, cudaStream_t st)
{
    using P = CostPlan<MAXW>;
    constexpr int TW = 32 - (BS - 1);
    // rows per band: tall bands re-read fewer halo rows (BS-1 per band); shorter ones when the batch is too small
    // to fill the machine otherwise
    int TR = 125;
    for (const int cand : { 125, 75, 47, 25 }) {
        TR = cand;
        if ((long long)b2s_cdiv(e.width1, TW) * b2s_cdiv(e.H, TR) * e.batch >= 4 * 148) break;
    }
    dim3 grid(b2s_cdiv(e.width1, TW), b2s_cdiv(e.H, TR), e.batch);
    const int nthreads = 32 * (e.Dp / 8);
    const int PW = plane_pitch(e.W);
    tmap_encode_fn enc = tmap_encoder();
    if (!enc) { b2s_set_error("cuTensorMapEncodeTiled is not available from this driver"); return B2S_E_CUDA; }
    // plane rows as a 3-d tensor of 16-byte groups: {8 u16, groups per plane row, plane rows}
    CUtensorMap tmap;
    const cuuint64_t rowu16 = P8 ? 2 * (cuuint64_t)PW : (cuuint64_t)PW, npl = P8 ? 6 : 12;
    const cuuint64_t gdim[3] = { 8, rowu16 / 8, (cuuint64_t)(e.batch / e.nst) * CN * e.Hsrc * npl };   // the source images
    const cuuint64_t gstr[2] = { 16, rowu16 * 2 };
    const cuuint32_t box[3] = { 8, (cuuint32_t)(P8 ? 2 * P::NELP : P::NELP) / 8, (cuuint32_t)npl };
    const cuuint32_t estr[3] = { 1, 1, 1 };
    CUresult cr = enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT16, 3, (void*)planesR, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) { b2s_set_error("cuTensorMapEncodeTiled failed (%d)", (int)cr); return B2S_E_CUDA; }
    if (e.D != e.Dp) {
        B2S_CUDA_CHECK(cudaFuncSetAttribute(k_cost_fast<BS, MAXW, true, P8, CN>, cudaFuncAttributeMaxDynamicSharedMemorySize, P::total));
        k_cost_fast<BS, MAXW, true, P8, CN><<<grid, nthreads, P::total, st>>>(descL, tmap, e, TR, cost, flag);
    } else {
        B2S_CUDA_CHECK(cudaFuncSetAttribute(k_cost_fast<BS, MAXW, false, P8, CN>, cudaFuncAttributeMaxDynamicSharedMemorySize, P::total));
        k_cost_fast<BS, MAXW, false, P8, CN><<<grid, nthreads, P::total, st>>>(descL, tmap, e, TR, cost, flag);
    }
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

// two columns per thread (k_cost_fast2): interleaved records with every cell real; B2S_COST2=0 falls back (A/B runs)
template <int BS, int MAXW, int CN, int NH>
static int launch_cost_fast2(const uint2* descL, const uint16_t* planesR, const SgbmEff& e, uint16_t* cost, int* flag, cudaStream_t st)
{
    using P = CostPlan2<MAXW * NH>;
    constexpr int TW = 64 - (BS - 1);
    int TR = 125;
    for (const int cand : { 125, 75, 47, 25 }) {
        TR = cand;
        if ((long long)b2s_cdiv(e.width1, TW) * b2s_cdiv(e.H, TR) * e.batch * NH >= 4 * 148) break;
    }
    dim3 grid(b2s_cdiv(e.width1, TW), b2s_cdiv(e.H, TR) * NH, e.batch);
    const int nthreads = 32 * MAXW;
    const int PW = plane_pitch(e.W);
    tmap_encode_fn enc = tmap_encoder();
    if (!enc) { b2s_set_error("cuTensorMapEncodeTiled is not available from this driver"); return B2S_E_CUDA; }
    CUtensorMap tmap;
    const cuuint64_t rowu16 = 2 * (cuuint64_t)PW;
    const cuuint64_t gdim[3] = { 8, rowu16 / 8, (cuuint64_t)(e.batch / e.nst) * CN * e.Hsrc * 3 };      // three packed planes per row
    const cuuint64_t gstr[2] = { 16, rowu16 * 2 };
    const cuuint32_t box[3] = { 8, (cuuint32_t)(2 * P::NELP) / 8, 3 };
    const cuuint32_t estr[3] = { 1, 1, 1 };
    CUresult cr = enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT16, 3, (void*)planesR, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) { b2s_set_error("cuTensorMapEncodeTiled failed (%d)", (int)cr); return B2S_E_CUDA; }
    B2S_CUDA_CHECK(cudaFuncSetAttribute(k_cost_fast2<BS, MAXW, CN, NH>, cudaFuncAttributeMaxDynamicSharedMemorySize, P::total));
    k_cost_fast2<BS, MAXW, CN, NH><<<grid, nthreads, P::total, st>>>(descL, tmap, e, TR, cost, flag);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

static bool cost2_enabled()
{
    static const bool on = [] { const char* s = getenv("B2S_COST2"); return !(s && s[0] == '0'); }();
    return on;
}
// k_cost_fast2 (and with it the packed format of the right descriptor planes): interleaved records, every cell real,
// Dp = 128 or 256 (CTAs of 16 warps; 256 disparities are split over two CTAs per strip, 1024-thread CTAs would be capped
// at 64 registers), blockSize <= 5 (larger blocks need BS * 8 registers for the vertical ring and spill).
static bool cost_uses_fast2(const SgbmEff& e)
{
    return e.swz && e.D == e.Dp && (e.Dp == 128 || e.Dp == 256) && 2 * e.SW2 + 1 <= 5 && cost2_enabled();
}

template <int BS, int MAXW>
static int launch_cost_fast(const uint2* descL, const uint16_t* planesR, const SgbmEff& e, uint16_t* cost, int* flag, cudaStream_t st)
{
    if constexpr (BS <= 5) {
        constexpr int NH = MAXW / 16;                      // Dp = 256: two CTAs of 16 warps per strip
        if (cost_uses_fast2(e))
            return e.cn == 3 ? launch_cost_fast2<BS, 16, 3, NH>(descL, planesR, e, cost, flag, st)
                             : launch_cost_fast2<BS, 16, 1, NH>(descL, planesR, e, cost, flag, st);
    }
    if (e.cn == 3)
        return e.swz ? launch_cost_fast<BS, MAXW, true, 3>(descL, planesR, e, cost, flag, st)
                     : launch_cost_fast<BS, MAXW, false, 3>(descL, planesR, e, cost, flag, st);
    return e.swz ? launch_cost_fast<BS, MAXW, true, 1>(descL, planesR, e, cost, flag, st)
                 : launch_cost_fast<BS, MAXW, false, 1>(descL, planesR, e, cost, flag, st);
}

// which of the two cost kernels serves a configuration (decides the format of the right descriptors)
static bool cost_is_fast(const SgbmEff& e)
{
    const int bs = 2 * e.SW2 + 1;
    return (e.Dp <= 128 && bs <= 11) || (e.Dp <= 256 && bs <= 7);
}

size_t b2s_desc_right_bytes(const SgbmEff& e)
{
    if (cost_is_fast(e)) return (size_t)e.batch * e.H * 12 * plane_pitch(e.W) * sizeof(uint16_t);
    return (size_t)e.batch * e.H * e.W * sizeof(uint2);
}

int b2s_launch_descriptors_right(const uint8_t* img, const SgbmEff& e, void* desc, cudaStream_t st)
{
    if (!cost_is_fast(e)) return b2s_launch_descriptors(img, e, (uint2*)desc, st);
    const int PW = plane_pitch(e.W);
    dim3 grid(b2s_cdiv(PW, B2S_PLANE_EPB), e.H, e.batch);
    if (cost_uses_fast2(e)) k_descriptor_planes8b<<<grid, 256, 0, st>>>(img, e.H, e.W, e.ftzero, PW, (uint32_t*)desc);
    else if (e.swz) k_descriptor_planes8<<<grid, 256, 0, st>>>(img, e.H, e.W, e.ftzero, PW, (uint32_t*)desc);
    else k_descriptor_planes<<<grid, 256, 0, st>>>(img, e.H, e.W, e.ftzero, PW, (uint16_t*)desc);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

static int b2s_launch_cost_generic(const uint2* descL, const uint2* descR, const SgbmEff& e, uint16_t* cost,
                                   int* overflow_flag, cudaStream_t st);

int b2s_launch_cost(const uint2* descL, const void* descRv, const SgbmEff& e, uint16_t* cost,
                    int* overflow_flag, cudaStream_t st)
{
    const int bs = 2 * e.SW2 + 1;
    const uint16_t* descR = (const uint16_t*)descRv;      // planes when cost_is_fast(e)
    if (!cost_is_fast(e)) return b2s_launch_cost_generic(descL, (const uint2*)descRv, e, cost, overflow_flag, st);
    if (e.Dp <= 128) {
        switch (bs) {
        case 1: return launch_cost_fast<1, 16>(descL, descR, e, cost, overflow_flag, st);
        case 3: return launch_cost_fast<3, 16>(descL, descR, e, cost, overflow_flag, st);
        case 5: return launch_cost_fast<5, 16>(descL, descR, e, cost, overflow_flag, st);
        case 7: return launch_cost_fast<7, 16>(descL, descR, e, cost, overflow_flag, st);
        case 9: return launch_cost_fast<9, 16>(descL, descR, e, cost, overflow_flag, st);
        case 11: return launch_cost_fast<11, 16>(descL, descR, e, cost, overflow_flag, st);
        }
    } else if (e.Dp <= 256) {
        switch (bs) {
        case 1: return launch_cost_fast<1, 32>(descL, descR, e, cost, overflow_flag, st);
        case 3: return launch_cost_fast<3, 32>(descL, descR, e, cost, overflow_flag, st);
        case 5: return launch_cost_fast<5, 32>(descL, descR, e, cost, overflow_flag, st);
        case 7: return launch_cost_fast<7, 32>(descL, descR, e, cost, overflow_flag, st);
        }
    }
    return B2S_E_UNSUPPORTED;                              // unreachable: cost_is_fast covers exactly the cases above
}

static int b2s_launch_cost_generic(const uint2* descL, const uint2* descR, const SgbmEff& e, uint16_t* cost,
                    int* overflow_flag, cudaStream_t st)
{
    const int DP2 = e.Dp / 2, bs = 2 * e.SW2 + 1;
    int TW = (COST_NT * COST_IPT) / DP2;
    if (TW < 1) TW = 1;
    if (TW > 32) TW = 32;
    if (TW > e.width1) TW = e.width1;
    // TW*DP2 <= COST_NT*COST_IPT must hold: Dp <= 2048
    if (TW * DP2 > COST_NT * COST_IPT) {
        b2s_set_error("numDisparities too large for the cost kernel (Dp=%d)", e.Dp);
        return B2S_E_UNSUPPORTED;
    }
    size_t smem = ((size_t)(TW + 2 * e.SW2) * DP2 + (size_t)bs * TW * DP2) * sizeof(uint32_t);
    while (smem > 200 * 1024 && TW > 1) {
        TW = (TW + 1) / 2;
        smem = ((size_t)(TW + 2 * e.SW2) * DP2 + (size_t)bs * TW * DP2) * sizeof(uint32_t);
    }
    if (smem > 220 * 1024) {
        b2s_set_error("blockSize*numDisparities too large for the cost kernel (needs %zu B smem)", smem);
        return B2S_E_UNSUPPORTED;
    }
    int TR = 64;
    // per device / context attribute: set before every launch, not once per process
    if (smem > 48 * 1024) B2S_CUDA_CHECK(cudaFuncSetAttribute(k_cost, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid(b2s_cdiv(e.width1, TW), b2s_cdiv(e.H, TR), e.batch);
    k_cost<<<grid, COST_NT, smem, st>>>(descL, descR, e, TW, TR, cost, overflow_flag);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}
