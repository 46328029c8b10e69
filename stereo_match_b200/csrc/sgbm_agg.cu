// sgbm_agg.cu -- SGBM path aggregation, one direction per launch: the generic path for
// 512 < numDisparities <= 1024 and for MODE_SGBM_3WAY with numDisparities % 8 != 0 (cv2's scalar-tail rules, TM below);
// everything else runs the fused wavefront in sgbm_wave*.cu.
// Replaces the L_r recurrence inside cv2.StereoSGBM.compute (MODE_SGBM: 5 paths,
// MODE_HH: 8 paths; reference call site stereo_vision/stereo_vision.py:178).
//
//   L_r(p,d) = Craw(p,d) + min( L_r(q,d), L_r(q,d-1)+P1, L_r(q,d+1)+P1, minL_r(q)+P2 ) - minL_r(q)
//   (identical to cv2's  C + min(...) - (minL+P2)  because cv2's C carries +P2)
//   S(p,d)   = min(32767, sum_r L_r(p,d))        -- order independent since every L_r >= 0
//
// Work decomposition: every scanline of the direction is an independent serial scan.  A group
// of G lanes owns one scanline; lane j keeps disparities [2R*j, 2R*j+2R) of L_r(q,.) in R
// registers as packed u16x2.  Per step: two shuffles fetch the d-1 / d+1 neighbours across
// lanes, VIMNMX/VIADDMNMX.U16x2 evaluate the recurrence two disparities per instruction, a
// G-lane butterfly gives minL.  Diagonal and vertical scanlines are indexed by their column
// in the current row and wrap around the image side (a path that enters from the side starts
// from L = Craw), so all W scanlines have the same length H and a warp touches contiguous
// memory.
#include "common.cuh"

template <int R>
__device__ __forceinline__ void load_regs(const uint16_t* p, uint32_t (&r)[R])
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

template <int R>
__device__ __forceinline__ void load_regs_rw(const uint16_t* p, uint32_t (&r)[R])
{
    if constexpr (R % 4 == 0) {
#pragma unroll
        for (int i = 0; i < R / 4; i++) {
            uint4 v = reinterpret_cast<const uint4*>(p)[i];
            r[4 * i] = v.x; r[4 * i + 1] = v.y; r[4 * i + 2] = v.z; r[4 * i + 3] = v.w;
        }
    } else if constexpr (R % 2 == 0) {
#pragma unroll
        for (int i = 0; i < R / 2; i++) {
            uint2 v = reinterpret_cast<const uint2*>(p)[i];
            r[2 * i] = v.x; r[2 * i + 1] = v.y;
        }
    } else {
#pragma unroll
        for (int i = 0; i < R; i++) r[i] = reinterpret_cast<const uint32_t*>(p)[i];
    }
}

template <int R>
__device__ __forceinline__ void store_regs(uint16_t* p, const uint32_t (&r)[R])
{
    if constexpr (R % 4 == 0) {
#pragma unroll
        for (int i = 0; i < R / 4; i++)
            reinterpret_cast<uint4*>(p)[i] = make_uint4(r[4 * i], r[4 * i + 1], r[4 * i + 2], r[4 * i + 3]);
    } else if constexpr (R % 2 == 0) {
#pragma unroll
        for (int i = 0; i < R / 2; i++) reinterpret_cast<uint2*>(p)[i] = make_uint2(r[2 * i], r[2 * i + 1]);
    } else {
#pragma unroll
        for (int i = 0; i < R; i++) reinterpret_cast<uint32_t*>(p)[i] = r[i];
    }
}

template <int G>
__device__ __forceinline__ uint32_t group_min(uint32_t v)
{
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// One aggregation step for a lane: Lp (in/out), C (cost), returns the lane-local packed min.
//
// TM (MODE_SGBM_3WAY with numDisparities % 8 != 0; cv2 4.13.0, probed -- oracle/stereo_oracle.c aggregate_dir_ex): cv2 runs
// d < t0 = 8 * (D / 8) on SIMD lanes and the rest in a scalar tail whose FIRST element t0 takes its d - 1 neighbour from
// the wrong place in the two paths cv2 updates in place.  qm[i] = 0xffff where register i's low cell is t0 (t0 is even).
//   TM = 1 (down): the NEW L_r(p, t0 - 1) of this pixel (the SIMD loop has already stored it) instead of L_r(q, t0 - 1)
//   TM = 2 (<-)  : L_r(q, t0) itself: + P1 it can never win, the neighbour is lost
template <int R, int G, bool PAD, int TM>
__device__ __forceinline__ uint32_t agg_step(uint32_t (&L)[R], const uint32_t (&C)[R], int lane,
                                             uint32_t p1x2, uint32_t deltax2, uint32_t minx2,
                                             const uint32_t (&padm)[R], const uint32_t (&qm)[R])
{
    // neighbours across lanes: d-1 of my first cell, d+1 of my last cell
    uint32_t up = __shfl_up_sync(0xffffffffu, L[R - 1], 1, G);
    uint32_t dn = __shfl_down_sync(0xffffffffu, L[0], 1, G);
    if (lane == 0) up = 0x7fff0000u;          // L(q,-1) = 32767
    if (lane == G - 1) dn = 0x00007fffu;      // L(q, D) = 32767
    uint32_t mn = 0xffffffffu;
    uint32_t prev = up;
#pragma unroll
    for (int i = 0; i < R; i++) {
        uint32_t cur = L[i];
        uint32_t next = (i + 1 < R) ? L[i + 1] : dn;
        uint32_t dm1 = __byte_perm(prev, cur, 0x5432);    // (d-1) for both halves
        uint32_t dp1 = __byte_perm(cur, next, 0x5432);    // (d+1) for both halves
        if (TM == 2) dm1 = (dm1 & ~qm[i]) | (cur & qm[i]);
        uint32_t m = __vminu2(dm1, dp1);
        uint32_t t = __viaddmin_u16x2(m, p1x2, deltax2);  // min(m + P1, minL + P2)
        uint32_t u = __vminu2(cur, t);
        uint32_t l = C[i] + u - minx2;                    // halves never borrow: u >= minL
        if (PAD) l |= padm[i];                            // padded cells (d >= D) pinned at 0x7fff
        prev = cur;
        L[i] = l;
        mn = __vminu2(mn, l);
    }
    return mn;
}

// TM = 1: the down path of the scalar-tail case.  With its d - 1 neighbour taken from the CURRENT pixel, L_r(p, t0) can
// fall below zero (cv2 and the restatement keep it as a signed int16; every other cell stays >= its cost), so this path
// runs cell by cell in signed 32-bit arithmetic -- it is a corner of a corner, not a hot path.  Returns the lane's
// signed minimum; minprev is the signed minimum of the predecessor.
template <int R, int G, bool PAD>
__device__ __forceinline__ int agg_step_tail_down(uint32_t (&L)[R], const uint32_t (&C)[R], int lane, int P1, int P2,
                                                  int minprev, int D, int t0)
{
    const uint32_t up = __shfl_up_sync(0xffffffffu, L[R - 1], 1, G);
    const uint32_t dn = __shfl_down_sync(0xffffffffu, L[0], 1, G);
    const int vm1 = lane == 0 ? 32767 : (int)(int16_t)(up >> 16);
    const int vp = lane == G - 1 ? 32767 : (int)(int16_t)(dn & 0xffffu);
    const int delta = minprev + P2, dbase = lane * 2 * R;
    int v[2 * R], nv[2 * R];
#pragma unroll
    for (int i = 0; i < R; i++) { v[2 * i] = (int)(int16_t)(L[i] & 0xffffu); v[2 * i + 1] = (int)(int16_t)(L[i] >> 16); }
#pragma unroll
    for (int c = 0; c < 2 * R; c++) {
        const int cost = (int)((C[c >> 1] >> (16 * (c & 1))) & 0xffffu);
        int b = (c > 0 ? v[c > 0 ? c - 1 : 0] : vm1) + P1;
        if (c > 0 && dbase + c == t0) b = nv[c > 0 ? c - 1 : 0] + P1;              // the new value of this pixel
        const int cc = (c < 2 * R - 1 ? v[c < 2 * R - 1 ? c + 1 : c] : vp) + P1;
        int l = cost + min(min(v[c], b), min(cc, delta)) - minprev;
        if (PAD && dbase + c >= D) l = 32767;
        nv[c] = (int)(int16_t)l;
    }
    // t0 is the first cell of this lane: its neighbour is the previous lane's new last cell
    const int nup = __shfl_up_sync(0xffffffffu, nv[2 * R - 1], 1, G);
    if (dbase == t0 && lane > 0) {
        const int cost = (int)(C[0] & 0xffffu);
        const int cc = (2 * R > 1 ? v[1] : vp) + P1;
        nv[0] = (int)(int16_t)(cost + min(min(v[0], nup + P1), min(cc, delta)) - minprev);
    }
    int mn = 32767;
#pragma unroll
    for (int i = 0; i < R; i++) {
        L[i] = ((uint32_t)nv[2 * i] & 0xffffu) | ((uint32_t)nv[2 * i + 1] << 16);
        mn = min(mn, min(nv[2 * i], nv[2 * i + 1]));
    }
    return mn;
}

// signed saturating S + L per half (the scalar-tail paths: S may hold the negative L of the down path)
__device__ __forceinline__ uint32_t sat_add_s16x2(uint32_t a, uint32_t b)
{
    const int lo = max(-32768, min(32767, (int)(int16_t)(a & 0xffffu) + (int)(int16_t)(b & 0xffffu)));
    const int hi = max(-32768, min(32767, (int)(int16_t)(a >> 16) + (int)(int16_t)(b >> 16)));
    return ((uint32_t)lo & 0xffffu) | ((uint32_t)hi << 16);
}

#define AGG_NT 256
template <int R, int G, bool PAD, bool FIRST, int TM>
__global__ void __launch_bounds__(AGG_NT)
k_aggregate(const uint16_t* __restrict__ cost, uint16_t* __restrict__ S, SgbmEff e, int dx, int dy)
{
    const int lane = threadIdx.x % G;
    const int nlines = dy == 0 ? e.H : e.width1;
    const int nsteps = dy == 0 ? e.width1 : e.H;
    const long long gid = ((long long)blockIdx.x * AGG_NT + threadIdx.x) / G;
    const long long total = (long long)nlines * e.batch;
    // whole groups retire together (G divides 32); keep idle groups alive for the shuffles
    const bool active = gid < total;
    const int b = active ? (int)(gid / nlines) : 0;
    const int line = active ? (int)(gid - (long long)b * nlines) : 0;

    int x, y;
    if (dy == 0) { y = line; x = dx > 0 ? 0 : e.width1 - 1; }
    else         { x = line; y = dy > 0 ? 0 : e.H - 1; }

    const size_t volsz = (size_t)e.H * e.width1 * e.Dp;
    const uint16_t* cb = cost + (size_t)b * volsz + (size_t)lane * 2 * R;
    uint16_t* sb = S + (size_t)b * volsz + (size_t)lane * 2 * R;

    uint32_t padm[R];
#pragma unroll
    for (int i = 0; i < R; i++) {
        int d0 = lane * 2 * R + 2 * i;
        padm[i] = PAD ? ((d0 >= e.D ? 0x7fffu : 0u) | (d0 + 1 >= e.D ? 0x7fff0000u : 0u)) : 0u;
    }
    uint32_t qm[R];
#pragma unroll
    for (int i = 0; i < R; i++) qm[i] = (TM != 0 && lane * 2 * R + 2 * i == e.simdD) ? 0xffffu : 0u;
    const uint32_t p1x2 = (uint32_t)e.P1 * 0x10001u;
    const uint32_t p2x2 = (uint32_t)e.P2 * 0x10001u;

    uint32_t L[R], C[R], Cn[R], Sv[R], Sn[R];
#pragma unroll
    for (int i = 0; i < R; i++) { L[i] = 0; Sv[i] = 0; Sn[i] = 0; }
    uint32_t minx2 = 0;
    int minprev = 0;                               // TM = 1: signed

    size_t off = ((size_t)y * e.width1 + x) * e.Dp;
    load_regs<R>(cb + off, C);
    if (!FIRST) load_regs_rw<R>(sb + off, Sv);

    for (int s = 0; s < nsteps; s++) {
        // next position (wrap around the image side for vertical/diagonal scanlines)
        int xn = x + dx, yn = y + dy;
        bool wrap = false;
        if (dy != 0) {
            if (xn < 0) { xn = e.width1 - 1; wrap = true; }
            else if (xn >= e.width1) { xn = 0; wrap = true; }
        }
        size_t offn = ((size_t)yn * e.width1 + xn) * e.Dp;
        if (s + 1 < nsteps) {                      // prefetch the next step's operands
            load_regs<R>(cb + offn, Cn);
            if (!FIRST) load_regs_rw<R>(sb + offn, Sn);
        }
        if (TM == 1) {
            int mn = agg_step_tail_down<R, G, PAD>(L, C, lane, e.P1, e.P2, minprev, e.D, e.simdD);
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
            minprev = mn;
#pragma unroll
            for (int i = 0; i < R; i++) Sv[i] = sat_add_s16x2(Sv[i], L[i]);
            if (active) store_regs<R>(sb + off, Sv);
        } else {
            uint32_t mn = agg_step<R, G, PAD, TM>(L, C, lane, p1x2, minx2 + p2x2, minx2, padm, qm);
            // S accumulate
#pragma unroll
            for (int i = 0; i < R; i++)
                Sv[i] = FIRST ? L[i] : (TM == 2 ? sat_add_s16x2(Sv[i], L[i]) : __viaddmin_u16x2(Sv[i], L[i], 0x7fff7fffu));
            if (active) store_regs<R>(sb + off, Sv);
            uint32_t m16 = min(mn & 0xffffu, mn >> 16);
            m16 = group_min<G>(m16);
            minx2 = m16 * 0x10001u;
        }
        if (wrap) {                                // the path enters from the image side
#pragma unroll
            for (int i = 0; i < R; i++) L[i] = 0;
            minx2 = 0;
            minprev = 0;
        }
#pragma unroll
        for (int i = 0; i < R; i++) { C[i] = Cn[i]; Sv[i] = Sn[i]; }
        x = xn; y = yn; off = offn;
    }
}

template <int R, int G>
static int launch_agg_rg(const uint16_t* cost, uint16_t* S, const SgbmEff& e, int dir, int first,
                         cudaStream_t st)
{
    int dx, dy;
    b2s_dir(dir, dx, dy);
    const long long nlines = (dy == 0 ? e.H : e.width1) * (long long)e.batch;
    const int grid = b2s_cdiv(nlines * G, AGG_NT);
    const bool pad = e.D != e.Dp;
    // cv2's scalar tail of MODE_SGBM_3WAY (numDisparities % 8 != 0, at least one SIMD group): down = dir 2, <- = dir 4
    const int tm = (e.simdD != e.D && e.simdD > 0) ? (dir == 2 ? 1 : (dir == 4 ? 2 : 0)) : 0;
    if (tm == 1) {
        if (pad) k_aggregate<R, G, true, false, 1><<<grid, AGG_NT, 0, st>>>(cost, S, e, dx, dy);
        else     k_aggregate<R, G, false, false, 1><<<grid, AGG_NT, 0, st>>>(cost, S, e, dx, dy);
    } else if (tm == 2) {
        if (pad) k_aggregate<R, G, true, false, 2><<<grid, AGG_NT, 0, st>>>(cost, S, e, dx, dy);
        else     k_aggregate<R, G, false, false, 2><<<grid, AGG_NT, 0, st>>>(cost, S, e, dx, dy);
    } else if (pad) {
        if (first) k_aggregate<R, G, true, true, 0><<<grid, AGG_NT, 0, st>>>(cost, S, e, dx, dy);
        else       k_aggregate<R, G, true, false, 0><<<grid, AGG_NT, 0, st>>>(cost, S, e, dx, dy);
    } else {
        if (first) k_aggregate<R, G, false, true, 0><<<grid, AGG_NT, 0, st>>>(cost, S, e, dx, dy);
        else       k_aggregate<R, G, false, false, 0><<<grid, AGG_NT, 0, st>>>(cost, S, e, dx, dy);
    }
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

int b2s_launch_aggregate(const uint16_t* cost, uint16_t* S, const SgbmEff& e, int dir, int first,
                         cudaStream_t st)
{
#define AGG_CASE(RR, GG) if (e.R == RR && e.G == GG) return launch_agg_rg<RR, GG>(cost, S, e, dir, first, st);
    AGG_CASE(12, 32) AGG_CASE(16, 32)          // D > 512; smaller D runs sgbm_wave, except
    AGG_CASE(2, 32) AGG_CASE(4, 32) AGG_CASE(8, 32)   // MODE_SGBM_3WAY with numDisparities % 8 != 0 (api.cu sgbm_effective)
#undef AGG_CASE
    b2s_set_error("no aggregation kernel for R=%d G=%d (numDisparities=%d)", e.R, e.G, e.D);
    return B2S_E_UNSUPPORTED;
}
