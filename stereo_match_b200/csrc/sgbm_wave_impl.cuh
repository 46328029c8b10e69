// sgbm_wave_impl.cuh -- the wavefront kernel template (see sgbm_wave.cuh for the design).
#pragma once
#include <type_traits>
#include "sgbm_wave.cuh"

namespace wave {


// ---- R packed registers <-> memory (generic address space; 16/8/4-byte vectors) ---------------
template <int R, int CSB = 16>
__device__ __forceinline__ void ld_regs(const uint16_t* p, uint32_t (&r)[R])
{
    if constexpr (R % 4 == 0) {
#pragma unroll
        for (int i = 0; i < R / 4; i++) {
            uint4 v = *reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(p) + i * CSB);
            r[4 * i] = v.x; r[4 * i + 1] = v.y; r[4 * i + 2] = v.z; r[4 * i + 3] = v.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < R / 2; i++) {
            uint2 v = reinterpret_cast<const uint2*>(p)[i];
            r[2 * i] = v.x; r[2 * i + 1] = v.y;
        }
    }
}

template <int R, int CSB = 16>
__device__ __forceinline__ void ldg_regs_stream(const uint16_t* p, uint32_t (&r)[R])
{
    if constexpr (R % 4 == 0) {
#pragma unroll
        for (int i = 0; i < R / 4; i++) {
            uint4 v = __ldcs(reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(p) + i * CSB));
            r[4 * i] = v.x; r[4 * i + 1] = v.y; r[4 * i + 2] = v.z; r[4 * i + 3] = v.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < R / 2; i++) {
            uint2 v = __ldcs(reinterpret_cast<const uint2*>(p) + i);
            r[2 * i] = v.x; r[2 * i + 1] = v.y;
        }
    }
}

template <int R, int CSB = 16>
__device__ __forceinline__ void st_regs(uint16_t* p, const uint32_t (&r)[R])
{
    if constexpr (R % 4 == 0) {
#pragma unroll
        for (int i = 0; i < R / 4; i++)
            *reinterpret_cast<uint4*>(reinterpret_cast<char*>(p) + i * CSB) = make_uint4(r[4 * i], r[4 * i + 1], r[4 * i + 2], r[4 * i + 3]);
    } else {
#pragma unroll
        for (int i = 0; i < R / 2; i++) reinterpret_cast<uint2*>(p)[i] = make_uint2(r[2 * i], r[2 * i + 1]);
    }
}

// same, through a 32-bit shared-window address (no generic-pointer arithmetic in the hot loop)
template <int R, int CSB = 16>
__device__ __forceinline__ void lds_regs(unsigned addr, uint32_t (&r)[R])
{
    if constexpr (R % 4 == 0) {
#pragma unroll
        for (int i = 0; i < R / 4; i++)
            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                         : "=r"(r[4 * i]), "=r"(r[4 * i + 1]), "=r"(r[4 * i + 2]), "=r"(r[4 * i + 3]) : "r"(addr + CSB * i));
    } else {
#pragma unroll
        for (int i = 0; i < R / 2; i++)
            asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(r[2 * i]), "=r"(r[2 * i + 1]) : "r"(addr + 8 * i));
    }
}
template <int R, int CSB = 16>
__device__ __forceinline__ void sts_regs(unsigned addr, const uint32_t (&r)[R])
{
    if constexpr (R % 4 == 0) {
#pragma unroll
        for (int i = 0; i < R / 4; i++)
            asm volatile("st.shared.v4.u32 [%4], {%0, %1, %2, %3};"
                         ::"r"(r[4 * i]), "r"(r[4 * i + 1]), "r"(r[4 * i + 2]), "r"(r[4 * i + 3]), "r"(addr + CSB * i) : "memory");
    } else {
#pragma unroll
        for (int i = 0; i < R / 2; i++)
            asm volatile("st.shared.v2.u32 [%2], {%0, %1};" ::"r"(r[2 * i]), "r"(r[2 * i + 1]), "r"(addr + 8 * i) : "memory");
    }
}
__device__ __forceinline__ uint32_t lds_u32(unsigned addr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts_u32(unsigned addr, uint32_t v)
{
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

#ifndef B2S_WAVE_DEP
#define B2S_WAVE_DEP 1
#endif
// pixels of the edge row between two release stores of the progress counter (each costs the publishing warp a
// gpu-scope fence, about 1000 cycles)
#ifndef B2S_WAVE_PUBW
#define B2S_WAVE_PUBW 32
#endif
#ifndef B2S_WAVE_PFD
#define B2S_WAVE_PFD 6
#endif

template <int G>
__device__ __forceinline__ uint32_t group_min_x2(uint32_t v)     // two independent u16 minima at once
{
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v = __vminu2(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__device__ __forceinline__ uint32_t half_min(uint32_t v) { return min(v & 0xffffu, v >> 16); }

// One recurrence step of one direction for this lane's 2R disparities:
//   L(d) <- C(d) + min( L(d), min(L(d-1), L(d+1), minL + P2 - P1) + P1 ) - minL
// L in/out; returns the packed lane-local minimum of the new L.
// IL (interleaved lanes, R == 8 records): register i packs the cells (2R*lane + i, 2R*lane + R + i), so the
// neighbours d-1 and d+1 of both halves are registers i-1 and i+1 and only the two end registers need a PRMT;
// otherwise register i packs (2R*lane + 2i, 2R*lane + 2i + 1) and every register needs its shifted copy.
template <int R, int G, bool PAD, bool IL>
__device__ __forceinline__ uint32_t step(uint32_t (&L)[R], const uint32_t (&C)[R], int lane, uint32_t p1x2,
                                         uint32_t p2m1x2, uint32_t minx2, const uint32_t (&padm)[R])
{
    uint32_t up = __shfl_up_sync(0xffffffffu, L[R - 1], 1, G);
    uint32_t dn = __shfl_down_sync(0xffffffffu, L[0], 1, G);
    if (lane == 0) up = 0x7fff0000u;          // L(q,-1) = 32767
    if (lane == G - 1) dn = 0x00007fffu;      // L(q, D) = 32767
    const uint32_t dmp = minx2 + p2m1x2;      // minL + P2 - P1  (P2 > P1)
    uint32_t m[R];
    if (IL) {
        // L[-1] = (cell 2R*lane - 1, cell 2R*lane + R - 1) ; L[R] = (cell 2R*lane + R, cell 2R*lane + 2R)
        const uint32_t lm1 = __byte_perm(up, L[R - 1], 0x5432), lr = __byte_perm(L[0], dn, 0x5432);
#pragma unroll
        for (int i = 0; i < R; i++) m[i] = __vimin3_u16x2(i == 0 ? lm1 : L[i - 1], i == R - 1 ? lr : L[i + 1], dmp);
    } else {
        uint32_t sh[R + 1];                   // sh[i] = (L(d-1), L(d)) for register i's low cell
        sh[0] = __byte_perm(up, L[0], 0x5432);
#pragma unroll
        for (int i = 1; i < R; i++) sh[i] = __byte_perm(L[i - 1], L[i], 0x5432);
        sh[R] = __byte_perm(L[R - 1], dn, 0x5432);
#pragma unroll
        for (int i = 0; i < R; i++) m[i] = __vimin3_u16x2(sh[i], sh[i + 1], dmp);
    }
    uint32_t mn = 0xffffffffu;
#pragma unroll
    for (int i = 0; i < R; i++) {
        uint32_t u = __viaddmin_u16x2(m[i], p1x2, L[i]);
        uint32_t l = C[i] + u - minx2;        // halves never borrow: u >= minL
        if (PAD) l |= padm[i];                // padded cells (d >= D) stay pinned at 0x7fff
        L[i] = l;
        if (i & 1) mn = __vimin3_u16x2(mn, L[i - 1], l);
    }
    return mn;
}

__device__ __forceinline__ int ld_acquire(const int* p)
{
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(int* p, int v)
{
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// out of line on purpose: the fence sequence of a release store must not be if-converted into the step
static __device__ __noinline__ void publish_progress(int* p, int v) { st_release(p, v); }
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem)
{
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar)
{
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, 0x989680;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// shared-memory plan (bytes), shared by host and device
template <int R, int G, int NT, bool VERT, bool FINAL>
struct Plan {
    static constexpr int RB = NT / G;
    static constexpr int DP = 2 * R * G;
    static constexpr int CH = (G == 8 && NT >= 256) ? 8 : 4;
    static constexpr int ring1 = 0;                                        // [RB][4][DP] u16
    static constexpr int ring2 = ring1 + (VERT ? RB * 4 * DP * 2 : 0);     // [RB][4][DP]
    static constexpr int ring3 = ring2 + (VERT ? RB * 4 * DP * 2 : 0);     // [RB][2][DP]
    static constexpr int stage = ring3 + (VERT ? RB * 2 * DP * 2 : 0);     // [2][CH][3][DP]
    static constexpr int zero = stage + (VERT ? 2 * CH * 3 * DP * 2 : 0);  // [DP]
    static constexpr int scratch = zero + (VERT ? DP * 2 : 0);             // [RB][DP] (FINAL)
    static constexpr int ringmin = scratch + (FINAL ? RB * DP * 2 : 0);    // [RB][4][4] u32
    static constexpr int stagemin = ringmin + (VERT ? RB * 16 * 4 : 0);    // [2][CH][4] u32
    static constexpr int zeromin = stagemin + (VERT ? 2 * CH * 4 * 4 : 0); // [4] u32
    static constexpr int item = zeromin + 16;
    static constexpr int bar = item + 8;                                   // mbarrier (8 bytes, 8-aligned)
    static constexpr int total = item + 16;
};

// ---- fused winner-take-all on the registers of one group (first loop) --------
template <int R, int G, bool PAD, int CSB>
__device__ __forceinline__ void wta(const uint32_t (&Sv)[R], int lane, const SgbmEff& e, unsigned long long magic,
                                    uint16_t* scr, bool act, size_t rowpix, int xreal, uint2* rec,
                                    uint32_t* disp2key)
{
    constexpr bool IL = CSB != 16;                                 // register i = cells (i, R + i) instead of (2i, 2i + 1)
    const int dbase = lane * 2 * R;
    uint32_t key = 0xffffffffu;
#pragma unroll
    for (int i = 0; i < R; i++) {
        uint32_t k0 = __byte_perm(Sv[i], IL ? i : 2 * i, 0x1054);          // S(lo) << 16 | cell
        uint32_t k1 = __byte_perm(Sv[i], IL ? R + i : 2 * i + 1, 0x3254);  // S(hi) << 16 | cell
        key = __vimin3_u32(key, k0, k1);
    }
    key += dbase;                                                  // cell -> disparity index
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) key = min(key, __shfl_xor_sync(0xffffffffu, key, o));
    int best = (int)(key & 0xffffu);                               // first minimum (lowest d on ties)
    const int minS = (int)(key >> 16);
    if (e.way3) {
        // MODE_SGBM_3WAY: a tied minimum goes to cv2's 8-lane winner (b2s_tie8_*, common.cuh).  Rare: the whole warp
        // takes the slow path when any of its pixels has a tie (it gives the first minimum again where there is none).
        bool tie = false;
#pragma unroll
        for (int i = 0; i < R; i++) {
            const int d0 = dbase + (IL ? i : 2 * i), d1 = dbase + (IL ? R + i : 2 * i + 1);
            tie = tie || ((int)(Sv[i] & 0xffffu) == minS && d0 != best && d0 < e.D) || ((int)(Sv[i] >> 16) == minS && d1 != best && d1 < e.D);
        }
        if (__any_sync(0xffffffffu, tie)) {
            unsigned long long w8 = 0ull;
#pragma unroll
            for (int i = 0; i < R; i++) {
                const int d0 = dbase + (IL ? i : 2 * i), d1 = dbase + (IL ? R + i : 2 * i + 1);
                if ((int)(Sv[i] & 0xffffu) == minS && d0 < e.D) w8 = b2s_tie8_merge(w8, b2s_tie8_one(d0));
                if ((int)(Sv[i] >> 16) == minS && d1 < e.D) w8 = b2s_tie8_merge(w8, b2s_tie8_one(d1));
            }
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) w8 = b2s_tie8_merge(w8, __shfl_xor_sync(0xffffffffu, w8, o));
            best = b2s_tie8_best(w8);
        }
    }

    // uniqueness: exists d, |d - best| > 1, S(d) * (100 - uniq) < minS * 100
    const int mul = 100 - e.uniq;
    bool bad_lane = false;
    if (mul > 0) {
        // S * mul < minS * 100  <=>  S < T,  T = ceil(minS * 100 / mul)   (exact magic division)
        // (MODE_SGBM_3WAY, e.uqle: S * mul <= minS * 100  <=>  S < floor(minS * 100 / mul) + 1)
        const unsigned a = (unsigned)(minS * 100 + mul - 1 + e.uqle);
        unsigned T = (unsigned)(((unsigned long long)a * magic) >> 40);
        if (e.uqle) T = (T & 0x8000u) ? 0u : (T & 0xffffu);        // cv2's 3WAY compares against (int16)T: wraps, <= 0 rejects nothing
        T = min(T, 0x8000u);
        const uint32_t K = (0x8000u - T) * 0x10001u;               // S + K has bit 15 set  <=>  S >= T
        uint32_t acc = 0;
#pragma unroll
        for (int i = 0; i < R; i++) acc = (acc >> 1) | ((Sv[i] + K) & 0x80008000u);
        // register i, low cell -> bit 16-R+i ; high cell -> bit 32-R+i
        uint32_t cells = (((1u << R) - 1u) << (16 - R)) * 0x10001u;
        if (PAD) {
            cells = 0;
#pragma unroll
            for (int i = 0; i < R; i++) {
                if (dbase + (IL ? i : 2 * i) < e.D) cells |= 1u << (16 - R + i);
                if (dbase + (IL ? R + i : 2 * i + 1) < e.D) cells |= 1u << (32 - R + i);
            }
        }
        uint32_t excl = 0;
        const int bl = best - dbase;
        if (IL && R == 8) {
            // cells bl-1 .. bl+1 of this lane as a 16-bit linear mask (bit c = cell c), then cell c -> bit 8 + c (c < 8)
            // or 16 + c (c >= 8): bytes 0 / 1 of the linear mask become bytes 1 / 3
            const unsigned sh = (unsigned)min(max(bl + 2, 0), 31);
            excl = __byte_perm(((7u << sh) >> 3) & 0xffffu, 0u, 0x1404);
        } else {
#pragma unroll
            for (int k = -1; k <= 1; k++) {
                int c = bl + k;
                if (c >= 0 && c < 2 * R) excl |= IL ? 1u << ((c / R) * 16 + 16 - R + (c % R)) : 1u << ((c & 1) * 16 + 16 - R + (c >> 1));
            }
        }
        bad_lane = (~acc & cells & ~excl) != 0;
    } else {
        const int thr = minS * 100 + e.uqle;
#pragma unroll
        for (int i = 0; i < R; i++) {
#pragma unroll
            for (int h = 0; h < 2; h++) {
                int d = dbase + (IL ? h * R + i : 2 * i + h);
                int val = h ? (int)(Sv[i] >> 16) : (int)(Sv[i] & 0xffffu);
                int ad = d - best;
                if (d < e.D && (ad > 1 || ad < -1) && val * mul < thr) bad_lane = true;
            }
        }
    }
    const unsigned gshift = (threadIdx.x & 31) & ~(G - 1);
    const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << (G & 31)) - 1u) << gshift);
    // every S saturated: cv2's search (strict compare from bestDisp = -1) has no winner and writes INVALID; not MODE_SGBM_3WAY
    const bool bad = (__ballot_sync(0xffffffffu, bad_lane) & gmask) != 0 || (minS >= 32767 && !e.way3);

    // neighbours of the winner through the group's scratch row
    st_regs<R, CSB>(scr + (CSB == 16 ? dbase : lane * 8), Sv);
    __syncwarp();
    if (lane == 0 && act) {
        const uint32_t sm1 = best > 0 ? scr[b2s_cell_pos(best - 1, R, G, CSB != 16)] : 0u;
        const uint32_t sp1 = best < e.D - 1 ? scr[b2s_cell_pos(best + 1, R, G, CSB != 16)] : 0u;
        const int xi = xreal + e.minX1;
        rec[rowpix + xi] = make_uint2((uint32_t)minS | (bad ? 0xffff0000u : (uint32_t)best << 16), sm1 | (sp1 << 16));
        if (!bad && minS < 32767) {       // cv2: disp2cost starts at SHRT_MAX and is only replaced by a smaller cost
            const int x2 = xi - best - e.minD;
            atomicMin(disp2key + rowpix + x2, ((uint32_t)minS << 16) | (uint32_t)(0xffff - xreal));
        }
    }
    __syncwarp();
}

// FA ("fast accumulate"): the host proved that S1 + the four L_r of a sweep stay below 65536, so S is summed with
// plain adds and clamped to 32767 once (all terms are >= 0: same result as cv2's saturating add per direction)
// DIAG = false (MODE_HH4): only the straight vertical direction comes from the row above; the two diagonal
// slots of the rings and of the edge stream stay unused.
template <int R, int G, int NT, bool PAD, bool VERT, bool FINAL, bool FA, bool DIAG = true>
__global__ void __launch_bounds__(NT, (G == 32 ? 1 : 576 / NT))
k_wave(SgbmEff e, WaveArgs a)
{
    using P = Plan<R, G, NT, VERT, FINAL>;
    constexpr int RB = P::RB, DP = P::DP, CH = P::CH;
    // Layout of one [Dp] record (cost, S, rings, edge stream): with R == 8 a lane's two 16-byte chunks sit
    // G*16 bytes apart (chunk c of lane j at (c*G + j)*16, see b2s_cell_pos), so that every vector access of
    // a group covers 128 contiguous bytes: conflict-free in shared memory, whole sectors in global memory.
    constexpr bool SWZ = (R == 8);
    constexpr int CSB = SWZ ? G * 16 : 16;
    extern __shared__ __align__(16) unsigned char smem[];
    uint16_t* const ring1 = reinterpret_cast<uint16_t*>(smem + P::ring1);
    uint16_t* const ring2 = reinterpret_cast<uint16_t*>(smem + P::ring2);
    uint16_t* const ring3 = reinterpret_cast<uint16_t*>(smem + P::ring3);
    uint16_t* const stage = reinterpret_cast<uint16_t*>(smem + P::stage);
    uint16_t* const zero = reinterpret_cast<uint16_t*>(smem + P::zero);
    uint16_t* const scratch = reinterpret_cast<uint16_t*>(smem + P::scratch);
    uint32_t* const ringmin = reinterpret_cast<uint32_t*>(smem + P::ringmin);
    uint32_t* const stagemin = reinterpret_cast<uint32_t*>(smem + P::stagemin);
    uint32_t* const zeromin = reinterpret_cast<uint32_t*>(smem + P::zeromin);
    int* const s_item = reinterpret_cast<int*>(smem + P::item);
    unsigned long long* const s_bar = reinterpret_cast<unsigned long long*>(smem + P::bar);

    if (threadIdx.x == 0) {
        *s_item = VERT ? (int)atomicAdd(a.ticket, 1u) : (int)blockIdx.x;
        mbar_init((unsigned)__cvta_generic_to_shared(s_bar), NT / 32);        // one arrival per warp and step
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (VERT) {
        for (int i = threadIdx.x; i < DP / 2; i += NT) reinterpret_cast<uint32_t*>(zero)[i] = 0u;
        if (threadIdx.x < 4) zeromin[threadIdx.x] = 0u;
    }
    __syncthreads();
    const int item = *s_item;
    // band-major order: every band-k item precedes all band-(k+1) items, so a consumer's producer
    // holds an earlier ticket and, with more than a few pairs in flight, is already far ahead
    const int band = item / e.batch, b = item - band * e.batch;
    const int eitem = b * a.nbands + band;              // slot in the edge / progress arrays
    const int g = threadIdx.x / G, lane = threadIdx.x % G;
    const int w1 = e.width1;
    const int ys = band * RB + g;
    const bool rowact = ys < e.H;
    const int nrows = min(RB, e.H - band * RB);
    const int gl = nrows - 1;
    const bool has_above = VERT && band > 0;
    const bool has_below = VERT && band < a.nbands - 1;
    constexpr int SKEW = VERT ? 2 : 0;
    const int T = w1 + SKEW * (nrows - 1);
    const int yreal = a.flip ? e.H - 1 - min(ys, e.H - 1) : min(ys, e.H - 1);
    const int xsign = a.flip ? -1 : 1, xorg = a.flip ? w1 - 1 : 0;      // xreal = xorg + xsign * xs

    const int lcell = SWZ ? lane * 8 : lane * 2 * R;                   // first cell slot of this lane in a record
    const size_t rowoff = ((size_t)b * e.H + yreal) * w1 * DP + (size_t)lcell;
    const uint16_t* const cb = a.cost + rowoff;
    uint16_t* const sb = a.S + rowoff;
    const size_t rowpix = ((size_t)b * e.H + yreal) * e.W;

    uint32_t padm[R];
#pragma unroll
    for (int i = 0; i < R; i++) {
        const int dlo = lane * 2 * R + (SWZ ? i : 2 * i), dhi = SWZ ? dlo + R : dlo + 1;   // cells of register i
        padm[i] = PAD ? ((dlo >= e.D ? 0x7fffu : 0u) | (dhi >= e.D ? 0x7fff0000u : 0u)) : 0u;
    }
    const uint32_t p1x2 = (uint32_t)e.P1 * 0x10001u;
    const uint32_t p2m1x2 = (uint32_t)(e.P2 - e.P1) * 0x10001u;

    uint32_t Lh[R], C[R], Cn[R], S1[R], S1n[R];
    uint32_t minh = 0;
#pragma unroll
    for (int i = 0; i < R; i++) { Lh[i] = 0; C[i] = 0; Cn[i] = 0; S1[i] = 0; S1n[i] = 0; }

    // edge stream of the band above: producer item = item - 1
    const uint16_t* const eabove = a.edge + (size_t)(eitem - 1) * w1 * 3 * DP;
    const uint32_t* const eabove_min = a.edge_min + (size_t)(eitem - 1) * w1 * 4;
    uint16_t* const ebelow = a.edge + (size_t)eitem * w1 * 3 * DP;
    uint32_t* const ebelow_min = a.edge_min + (size_t)eitem * w1 * 4;
    auto stage_chunk = [&](int c) {          // cp.async chunk c of the edge stream into stage[c & 1]
        const int x0 = c * CH, npx = min(CH, w1 - x0);
        const int n16 = npx * 3 * DP * 2 / 16;
        const unsigned char* src = reinterpret_cast<const unsigned char*>(eabove + (size_t)x0 * 3 * DP);
        unsigned char* dst = reinterpret_cast<unsigned char*>(stage + (size_t)(c & 1) * CH * 3 * DP);
        for (int i = threadIdx.x; i < n16; i += NT) cp_async16(dst + 16 * i, src + 16 * i);
        if ((int)threadIdx.x < npx)
            cp_async16(stagemin + ((c & 1) * CH + threadIdx.x) * 4, eabove_min + (size_t)(x0 + threadIdx.x) * 4);
    };
    // thread 0 remembers the last progress value it saw: once the producer band is far ahead (the steady state of the
    // band-major ticket order) whole stretches of chunks need no acquire load at all
    int known = 0;
    auto wait_progress = [&](int need) {
        if (known >= need) return;
        const int* p = a.progress + (eitem - 1);
        int v;
        while ((v = ld_acquire(p)) < need) __nanosleep(64);
        known = v;
    };
    if (has_above) {
        if (threadIdx.x == 0) wait_progress(min(CH, w1));
        __syncthreads();
        stage_chunk(0);
        cp_async_wait_all();
        __syncthreads();
    }

    // ---- loop-invariant addressing (32-bit shared-window addresses, running global pointers) -------
    const unsigned sm0 = (unsigned)__cvta_generic_to_shared(smem);
    const unsigned laneb = (unsigned)lcell * 2;                        // byte offset of this lane's cells
    // predecessor sources of the three vertical directions: the row above in the rings, or, for the
    // band's first row, the staged edge stream; `rowok` false (very first row of the sweep): zeros
    const bool rowok = g > 0 || has_above;
    unsigned rsrc[3], rmask[3], msrc[3];
    const unsigned rstr = (g > 0) ? DP * 2 : 3 * DP * 2;               // bytes between consecutive slots
    const unsigned mmask = (g > 0) ? 3u : (unsigned)(2 * CH - 1);
    if (g > 0) {
        rsrc[0] = sm0 + P::ring1 + (g - 1) * 4 * DP * 2 + laneb; rmask[0] = 3;
        rsrc[1] = sm0 + P::ring2 + (g - 1) * 4 * DP * 2 + laneb; rmask[1] = 3;
        rsrc[2] = sm0 + P::ring3 + (g - 1) * 2 * DP * 2 + laneb; rmask[2] = 1;
#pragma unroll
        for (int k = 0; k < 3; k++) msrc[k] = sm0 + P::ringmin + (g - 1) * 64 + (k + 1) * 4;
    } else {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            rsrc[k] = sm0 + P::stage + k * DP * 2 + laneb; rmask[k] = 2 * CH - 1;
            msrc[k] = sm0 + P::stagemin + k * 4;
        }
    }
    const unsigned zsrc = sm0 + P::zero + laneb, zmin = sm0 + P::zeromin;
    const unsigned wdst1 = sm0 + P::ring1 + g * 4 * DP * 2 + laneb, wdst2 = sm0 + P::ring2 + g * 4 * DP * 2 + laneb,
                   wdst3 = sm0 + P::ring3 + g * 2 * DP * 2 + laneb, wmin = sm0 + P::ringmin + g * 64;
    const bool edgerow = has_below && g == gl;
    const unsigned pubthread = (unsigned)gl * G;                        // lane 0 of the band's last row

    // running pointers: element xs of this thread's row (linear in xs, also while xs is out of range)
    const long long xstep = (long long)xsign * DP;
    const uint16_t* cptr = cb + ((long long)xorg + (long long)xsign * (-SKEW * g + 1)) * DP;   // next pixel's C
    const uint16_t* s1ptr = sb + ((long long)xorg + (long long)xsign * (-SKEW * g + 1)) * DP;  // next pixel's S1
    uint16_t* optr = (FINAL ? a.S_out + rowoff : sb) + ((long long)xorg + (long long)xsign * (-SKEW * g)) * DP;
    uint16_t* eptr = ebelow + (long long)(-SKEW * g) * 3 * DP + lcell;
    // first operands
    {
        const int xs = -SKEW * g;
        if (rowact && xs >= 0 && xs < w1) {
            const size_t off = (size_t)(xorg + xsign * xs) * DP;
            ldg_regs_stream<R, CSB>(cb + off, C);
            if (FINAL) ldg_regs_stream<R, CSB>(sb + off, S1);
        }
    }
    unsigned phase = 0;
    const unsigned bar_addr = sm0 + P::bar;
    constexpr int PFD = B2S_WAVE_PFD;                                            // L2 prefetch distance in steps
    const int pfoff = (SWZ && (lane & 1)) ? CSB / 2 : 0;                // in u16 elements

    // One wavefront step.  Cc/S1c: operands of this step, Cx/S1x: prefetch targets for the next one
    // (the caller alternates the two register sets, so nothing is copied).
    // `fast` (compile-time tag): every row of this warp is strictly inside the image at this step, has a row
    // above and stays in range for the prefetches -- all per-pixel range tests and the zero-record selects vanish.
    // `odd` (compile-time tag): parity of t.  The rare events of the band pipeline happen at fixed residues of t mod CH
    // (CH is 4 or 8) and of (t - 2*gl) mod B2S_WAVE_PUBW, so each body only carries the tests of its own parity.
    auto body = [&](auto fast, auto odd, const int t, uint32_t (&Cc)[R], uint32_t (&Cx)[R], uint32_t (&S1c)[R], uint32_t (&S1x)[R]) {
        constexpr bool FAST = decltype(fast)::value;
        constexpr bool ODD = decltype(odd)::value;
        if (VERT && t > 0) { mbar_wait(bar_addr, phase); phase ^= 1u; }   // every warp finished step t-1's ring stores
        if (!ODD && has_below) {
            // uniform: the band's last row finished `xdone` pixels in the steps before this one.  The release
            // store of one thread is cumulative over the edge stores the other lanes made before they arrived.
            // (xdone == w1 is only reached after the loop, see the end of the kernel.)
            const int xdone = t - SKEW * gl;
            if (xdone > 0 && (xdone % B2S_WAVE_PUBW) == 0 && threadIdx.x == pubthread)
                publish_progress(a.progress + eitem, xdone);
        }
        const int xs = t - SKEW * g;
        const bool act = FAST || (rowact && (unsigned)xs < (unsigned)w1);
        if (!ODD && has_above && (t % CH) == 2 && t >= 2) {      // chunk staging (polled during step t-1)
            const int c = (t - 2) / CH + 1;
            if (c * CH < w1) stage_chunk(c);
        }
        // prefetch the next pixel's operands
        {
            const bool actn = FAST || (rowact && (unsigned)(xs + 1) < (unsigned)w1);
            if (!FAST) {
#pragma unroll
                for (int i = 0; i < R; i++) { Cx[i] = 0; }
            }
            if (actn) {
                // ptxas tracks every LDG of this loop on ONE scoreboard, and a scoreboard wait means "all of them
                // returned": if the loads of the next step were issued before the first use of this step's operands,
                // that use would wait for loads issued a few instructions ago instead of a whole step ago
                // (10 % of pass 1's stall samples).  The always-zero `dep` makes the new addresses depend on the old
                // data, so the wait is taken while only the old loads are in flight.
#if B2S_WAVE_DEP
                const unsigned dep = (FINAL ? (Cc[0] | S1c[0]) : Cc[0]) & a.zero;
#else
                const unsigned dep = 0;
#endif
                ldg_regs_stream<R, CSB>(cptr + dep, Cx);
                if (FINAL) ldg_regs_stream<R, CSB>(s1ptr + dep, S1x);
            }
            // pull the lines of a few steps ahead into L2: the register prefetch above then waits for an
            // L2 hit instead of a DRAM round trip (one step is about as long as the loaded DRAM latency)
            if (FAST || (rowact && (unsigned)(xs + 1 + PFD) < (unsigned)w1)) {
                // (permuted records: a lane's two 16-byte chunks sit in different 128-byte lines, and a
                // prefetch touches one 32-byte sector: even lanes cover chunk 0's line, odd lanes chunk 1's)
                prefetch_l2(cptr + PFD * xstep + pfoff);
                if (FINAL) prefetch_l2(s1ptr + PFD * xstep + pfoff);
            }
            cptr += xstep;
            if (FINAL) s1ptr += xstep;
        }

        uint32_t Sv[R];
        if (VERT) {
            // ---- the three directions that come from the row above: ring -> registers -> ring -----
            uint32_t Lv[3][R], minv[3], mnv[3];
#pragma unroll
            for (int k = 0; k < 3; k++) {
                if (!DIAG && k != 1) {
#pragma unroll
                    for (int i = 0; i < R; i++) Lv[k][i] = 0u;
                    minv[k] = 0u; mnv[k] = 0u;
                    continue;
                }
                const int xp = xs + k - 1;                       // (x-1, x, x+1) of the row above
                const bool has_pred = FAST || (rowok && (unsigned)xp < (unsigned)w1);
                const unsigned src = has_pred ? rsrc[k] + ((unsigned)xp & rmask[k]) * rstr : zsrc;
                const unsigned ms = has_pred ? msrc[k] + ((unsigned)xp & mmask) * 16 : zmin;
                lds_regs<R, CSB>(src, Lv[k]);
                minv[k] = lds_u32(ms);
            }
#pragma unroll
            for (int k = 0; k < 3; k++)
                if (DIAG || k == 1) mnv[k] = step<R, G, PAD, SWZ>(Lv[k], Cc, lane, p1x2, p2m1x2, minv[k], padm);
            if (act) {
                if (DIAG) sts_regs<R, CSB>(wdst1 + (xs & 3) * (DP * 2), Lv[0]);
                sts_regs<R, CSB>(wdst2 + (xs & 3) * (DP * 2), Lv[1]);
                if (DIAG) sts_regs<R, CSB>(wdst3 + (xs & 1) * (DP * 2), Lv[2]);
                if (edgerow) {
                    if (DIAG) st_regs<R, CSB>(eptr, Lv[0]);
                    st_regs<R, CSB>(eptr + DP, Lv[1]);
                    if (DIAG) st_regs<R, CSB>(eptr + 2 * DP, Lv[2]);
                }
            }
            eptr += 3 * DP;
            const uint32_t r12 = group_min_x2<G>(half_min(mnv[0]) | (half_min(mnv[1]) << 16));
            const uint32_t r3 = DIAG ? group_min_x2<G>(half_min(mnv[2])) : 0u;
            if (act && lane == 0) {
                const uint32_t m1 = (r12 & 0xffffu) * 0x10001u, m2 = (r12 >> 16) * 0x10001u, m3 = (r3 & 0xffffu) * 0x10001u;
                const unsigned md = wmin + (xs & 3) * 16;
                sts_u32(md + 4, m1); sts_u32(md + 8, m2); sts_u32(md + 12, m3);
                if (edgerow) {
                    uint32_t* ge = ebelow_min + (size_t)xs * 4;
                    ge[0] = m1; ge[1] = m2; ge[2] = m3;
                }
            }
            if (has_above) {
                if (!ODD && (t % CH) == CH - 2) cp_async_wait_all();
                if (ODD && (t % CH) == 1 && threadIdx.x == 0) {
                    const int c = t / CH + 1;
                    if (c * CH < w1) wait_progress(min((c + 1) * CH, w1));
                }
            }
            // this warp's part of step t is visible: arrive, and keep working on what no other row needs
            __syncwarp();
            if ((threadIdx.x & 31) == 0) mbar_arrive(bar_addr);
#pragma unroll
            for (int i = 0; i < R; i++) {
                if (FA) {
                    Sv[i] = FINAL ? S1c[i] + Lv[0][i] + Lv[1][i] + Lv[2][i] : Lv[0][i] + Lv[1][i] + Lv[2][i];
                } else {
                    uint32_t v = __viaddmin_u16x2(Lv[0][i], Lv[1][i], 0x7fff7fffu);
                    v = __viaddmin_u16x2(v, Lv[2][i], 0x7fff7fffu);
                    Sv[i] = FINAL ? __viaddmin_u16x2(v, S1c[i], 0x7fff7fffu) : v;
                }
            }
        } else {
#pragma unroll
            for (int i = 0; i < R; i++) Sv[i] = FINAL ? S1c[i] : 0u;
        }
        // ---- direction 0: along the row, state in registers -----------------------------------
        const uint32_t mn0 = step<R, G, PAD, SWZ>(Lh, Cc, lane, p1x2, p2m1x2, minh, padm);
#pragma unroll
        for (int i = 0; i < R; i++) Sv[i] = __viaddmin_u16x2(Sv[i], Lh[i], 0x7fff7fffu);
        minh = (group_min_x2<G>(half_min(mn0)) & 0xffffu) * 0x10001u;

        // ---- S ----------------------------------------------------------------------------------
        if (FINAL) {
            wta<R, G, PAD, CSB>(Sv, lane, e, a.magic, scratch + g * DP, act, rowpix, xorg + xsign * xs, a.rec, a.disp2key);
            if (a.S_out != nullptr && act) st_regs<R, CSB>(optr, Sv);
        } else {
            if (act) st_regs<R, CSB>(optr, Sv);
        }
        optr += xstep;
    };

    uint32_t C2[R], S2[R];
#pragma unroll
    for (int i = 0; i < R; i++) { C2[i] = 0; S2[i] = 0; }
    (void)Cn; (void)S1n; (void)ring1; (void)ring2; (void)ring3; (void)ringmin; (void)zero; (void)zeromin;
    // steps at which all rows of this warp are interior (see `fast` above)
    const int gw0 = (int)(threadIdx.x & ~31u) / G, gw3 = gw0 + 32 / G - 1;
    const bool warp_full = VERT && (gw3 < nrows) && (gw0 > 0 || has_above);
    const int tfast_lo = warp_full ? SKEW * gw3 + 1 : 0x7fffffff;
    const int tfast_hi = w1 - 2 - PFD + SKEW * gw0;                   // inclusive
#pragma unroll 1
    for (int t = 0; t < T; t += 2) {
        if (t >= tfast_lo && t + 1 <= tfast_hi) {
            body(std::true_type{}, std::false_type{}, t, C, C2, S1, S2);
            body(std::true_type{}, std::true_type{}, t + 1, C2, C, S2, S1);
        } else {
            body(std::false_type{}, std::false_type{}, t, C, C2, S1, S2);
            if (t + 1 < T) body(std::false_type{}, std::true_type{}, t + 1, C2, C, S2, S1);
        }
    }
    if (has_below) {                       // the last chunk(s) of the edge stream
        mbar_wait(bar_addr, phase);
        if (threadIdx.x == pubthread) publish_progress(a.progress + eitem, w1);
    }
}

template <int R, int G, int NT, bool PAD, bool VERT, bool FINAL, bool FA, bool DIAG = true>
int launch_one(const SgbmEff& e, const WaveArgs& a, cudaStream_t st)
{
    using P = Plan<R, G, NT, VERT, FINAL>;
    // the attribute belongs to the current device's context (one process may drive several GPUs): set it every launch
    B2S_CUDA_CHECK(cudaFuncSetAttribute(k_wave<R, G, NT, PAD, VERT, FINAL, FA, DIAG>, cudaFuncAttributeMaxDynamicSharedMemorySize, P::total));
    const int grid = e.batch * a.nbands;
    k_wave<R, G, NT, PAD, VERT, FINAL, FA, DIAG><<<grid, NT, P::total, st>>>(e, a);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

template <int R, int G, int NT, bool FA>
int launch_rgf(const SgbmEff& e, const WaveArgs& a, int vert, int final, cudaStream_t st)
{
    const bool pad = e.D != e.Dp;
    if (vert && !a.diag) {      // MODE_HH4, first sweep of MODE_SGBM_3WAY
        if (!final) return pad ? launch_one<R, G, NT, true, true, false, FA, false>(e, a, st) : launch_one<R, G, NT, false, true, false, FA, false>(e, a, st);
        return pad ? launch_one<R, G, NT, true, true, true, FA, false>(e, a, st) : launch_one<R, G, NT, false, true, true, FA, false>(e, a, st);
    }
    if (vert && !final) return pad ? launch_one<R, G, NT, true, true, false, FA>(e, a, st) : launch_one<R, G, NT, false, true, false, FA>(e, a, st);
    if (vert && final) return pad ? launch_one<R, G, NT, true, true, true, FA>(e, a, st) : launch_one<R, G, NT, false, true, true, FA>(e, a, st);
    if (!vert && final) return pad ? launch_one<R, G, NT, true, false, true, false>(e, a, st) : launch_one<R, G, NT, false, false, true, false>(e, a, st);
    b2s_set_error("wave: unsupported pass combination");
    return B2S_E_INVALID;
}

template <int R, int G, int NT>
int launch_rg(const SgbmEff& e, const WaveArgs& a, int vert, int final, cudaStream_t st)
{
    // the fast-accumulate variant is only built for the widest lanes (R = 8: D = 128 / 256 / 512)
    if (R == 8 && a.fastacc) return launch_rgf<R, G, NT, (R == 8)>(e, a, vert, final, st);
    return launch_rgf<R, G, NT, false>(e, a, vert, final, st);
}

// NT threads per CTA = NT / G rows per band (b2s_wave_threads picks it on the host)
template <int G, int NT>
int launch_g(const SgbmEff& e, const WaveArgs& a, int vert, int final, cudaStream_t st)
{
    switch (e.R) {
    case 2: return launch_rg<2, G, NT>(e, a, vert, final, st);
    case 4: return launch_rg<4, G, NT>(e, a, vert, final, st);
    case 6: return launch_rg<6, G, NT>(e, a, vert, final, st);
    case 8: return launch_rg<8, G, NT>(e, a, vert, final, st);
    }
    b2s_set_error("wave: no kernel for R=%d G=%d", e.R, G);
    return B2S_E_UNSUPPORTED;
}

}  // namespace wave
