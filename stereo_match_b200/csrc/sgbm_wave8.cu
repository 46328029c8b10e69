// sgbm_wave8.cu -- the lean wavefront kernel for the widest lanes (R = 8 packed registers per lane, D = 16*G:
// D = 128 with G = 8, D = 256 with G = 16), both sweeps of MODE_HH and the first sweep of MODE_SGBM
// (reference call site stereo_vision/stereo_vision.py:178).
//
// Same decomposition, band pipeline and results as k_wave (sgbm_wave.cuh / sgbm_wave_impl.cuh); what differs is the
// instruction budget of the steady state.  The generic kernel spends 372 (sweep 1) / 573 (sweep 2) warp
// instructions per step where the recurrence itself needs about 150; this one is organised so that almost nothing
// else is left inside the loop:
//   * steps come in blocks of four with every ring slot a compile-time constant: the producer of step t writes slot
//     (t+3)&3 / (t+2)&3 / (t+1)&1 of its three rings, so the consumer reads slot t&3 / t&3 / t&1 at step t;
//   * the first row of a band reads the band above through a stage with the SAME slot rule (the edge stream in HBM
//     is indexed by the consumer's step, one array per direction), its base registers toggle by XOR once per block,
//     so row 0 and the other rows run identical code;
//   * the three minima of the row above travel as one 16-byte record per consumer step (one LDS.128);
//   * lane minima of two directions share one butterfly (the horizontal direction joins the next step's);
//   * global addresses are one IMAD.WIDE off a 32-bit running offset that C, S1 and S share;
//   * the sweep direction is a template parameter (the final sweep is the flipped one);
//   * the winner-take-all builds half of its keys and its whole uniqueness mask on the fma pipe
//     (IMAD with a power of two) and tests S >= T with one VIADDMNMX.S16x2.RELU per register.
// Steps at which a warp touches the image border run a guarded copy of the same body.
// Bands are 32 rows with 8 lanes per pixel (two 256-thread CTAs per SM) and 16 rows with 16 lanes; a band publishes its
// progress every WaveArgs::pubw pixels (32, or 8 when every work item is resident: small batches).
#include "sgbm_wave_impl.cuh"
#include <cstdlib>

namespace wave8 {

using wave::ld_acquire;
using wave::publish_progress;
using wave::mbar_init;
using wave::mbar_arrive;
using wave::mbar_wait;
using wave::group_min_x2;
using wave::step;

// step index modulo 4: a compile-time constant in the fast blocks, a run-time value in the guarded ones
template <int V> struct Jc { __device__ constexpr operator int() const { return V; } };
struct Jr { int v; __device__ operator int() const { return v; } };

constexpr int PFD = 6;       // L2 prefetch distance in steps
// (pixels between two publications of a band's progress: WaveArgs::pubw, a power of two chosen on the host)

template <int G, int NT, bool FINAL>
struct Plan {
    static constexpr int RB = NT / G;
    static constexpr int REC = 32 * G;                                  // bytes of one [Dp] record
    static constexpr int ALIGN = 8 * REC;                               // the stage toggles by XOR with 2/4 records
    static constexpr int stage = 0;                                     // [3][2][4][REC]
    static constexpr int ringA = stage + 24 * REC;                      // [RB][4][REC]   predecessor x-1
    static constexpr int ringB = ringA + RB * 4 * REC;                  // [RB][4][REC]   predecessor x
    static constexpr int ringC = ringB + RB * 4 * REC;                  // [RB][2][REC]   predecessor x+1
    static constexpr int scratch = ringC + RB * 2 * REC;                // [RB][REC]      (FINAL)
    static constexpr int ringmin = scratch + (FINAL ? RB * REC : 0);    // [RB][4] uint4  (-, m0, m1, m2) per consumer step
    static constexpr int MINROW = 80;                                   // bytes per row of ringmin: 64 + 16, so that the four rows of a warp hit different banks
    static constexpr int stagemin = ringmin + RB * MINROW;              // [2][4] uint4, 128-byte aligned (toggles by XOR 64)
    static_assert((RB * MINROW) % 128 == 0, "stagemin must stay 128-byte aligned");
    static constexpr int zero = stagemin + 128;                         // [REC] zeros, then 16 zero bytes
    static constexpr int misc = zero + REC + 16;                        // work item (int), step mbarrier at +8, stage mbarrier at +16
    static constexpr int total = misc + 32 + ALIGN;                     // + slack to align the stage
};

__device__ __forceinline__ void lds128(unsigned addr, uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d)
{
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr));
}
__device__ __forceinline__ void lds_rec(unsigned addr, uint32_t (&r)[8], int csb)
{
    lds128(addr, r[0], r[1], r[2], r[3]);
    lds128(addr + csb, r[4], r[5], r[6], r[7]);
}
__device__ __forceinline__ void sts_rec(unsigned addr, const uint32_t (&r)[8], int csb)
{
    asm volatile("st.shared.v4.u32 [%4], {%0, %1, %2, %3};" ::"r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(addr) : "memory");
    asm volatile("st.shared.v4.u32 [%4], {%0, %1, %2, %3};" ::"r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(addr + csb) : "memory");
}
__device__ __forceinline__ void sts32(unsigned addr, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t lds16(unsigned addr)
{
    uint32_t v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void cp16(unsigned dst, const void* src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
// this thread's earlier cp.async copies arrive on the mbarrier when they land (no scoreboard wait in the issuing warp)
__device__ __forceinline__ void cp_async_arrive(unsigned bar)
{
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void pf_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void ldg_rec(const uint4* p, uint32_t (&r)[8], int g)
{
    const uint4 v0 = __ldcs(p), v1 = __ldcs(p + g);
    r[0] = v0.x; r[1] = v0.y; r[2] = v0.z; r[3] = v0.w; r[4] = v1.x; r[5] = v1.y; r[6] = v1.z; r[7] = v1.w;
}
__device__ __forceinline__ void stg_rec(uint4* p, const uint32_t (&r)[8], int g)
{
    p[0] = make_uint4(r[0], r[1], r[2], r[3]);
    p[g] = make_uint4(r[4], r[5], r[6], r[7]);
}
// streaming store (evict-first in L2): the S1 volume is read again only by the next launch, long after it left L2
__device__ __forceinline__ void stg_rec_cs(uint4* p, const uint32_t (&r)[8], int g)
{
    __stcs(p, make_uint4(r[0], r[1], r[2], r[3]));
    __stcs(p + g, make_uint4(r[4], r[5], r[6], r[7]));
}
// per 16-bit half: 1 if (short)s + (short)k >= 1 else 0   (VIADDMNMX.S16x2.RELU)
__device__ __forceinline__ uint32_t ge_flag_x2(uint32_t s, uint32_t k) { return __viaddmin_s16x2_relu(s, k, 0x00010001u); }

// ---- fused winner-take-all on the registers of one group (first loop) ----------------------------
// Sv: register i packs S of the cells (16*lane + i, 16*lane + 8 + i).  scr: this row's scratch record (shared window).
template <int G>
__device__ __forceinline__ void wta8(const uint32_t (&Sv)[8], int lane, const SgbmEff& e, unsigned magic32, unsigned scr,
                                     bool act, unsigned pix, int xcode, uint2* __restrict__ rec, uint32_t* __restrict__ disp2key,
                                     uint32_t k65536)
{
    const unsigned dbase = (unsigned)lane * 16u;
    uint32_t key = 0xffffffffu;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t k0 = Sv[i] * k65536 + (uint32_t)i;                    // S(lo) << 16 | cell   (IMAD: fma pipe)
        const uint32_t k1 = __byte_perm(Sv[i], 8 + i, 0x3254);               // S(hi) << 16 | cell
        key = __vimin3_u32(key, k0, k1);
    }
    key += dbase;
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) key = min(key, __shfl_xor_sync(0xffffffffu, key, o));
    const unsigned best = key & 0xffffu, minS = key >> 16;                   // first minimum (lowest d on ties)

    // the scratch copy of S serves the winner's two neighbours
    sts_rec(scr + lane * 16, Sv, G * 16);                                    // chunk c of lane j at (c*G + j)*16: no bank conflicts

    // uniqueness: exists d, |d - best| > 1, S(d) * (100 - uniq) < minS * 100   <=>   S(d) < T = ceil(minS*100 / mul)
    const unsigned mul = 100u - (unsigned)e.uniq;
    const unsigned T = min(__umulhi(minS * 100u + mul - 1u, magic32), 0x8000u);   // exact for minS*100 + 99 < 2^25
    const uint32_t kneg = ((1u - max(T, 1u)) & 0xffffu) * 0x10001u;          // S + (1 - T) >= 1  <=>  S >= T
    uint32_t ge = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) ge = ge_flag_x2(Sv[i], kneg) * (1u << i) + ge;     // bit i: cell i, bit 16+i: cell 8+i
    // cells best-1 .. best+1 of this lane: 16-bit linear mask, byte 0 -> bits 0-7, byte 1 -> bits 16-23
    const int bl2 = (int)best - (int)dbase + 2;
    const unsigned lin = ((7u << min(max(bl2, 0), 31)) >> 3) & 0xffffu;
    const unsigned excl = __byte_perm(lin, 0u, 0x4140);
    const bool bad_lane = ((~ge) & 0x00ff00ffu & ~excl) != 0u && T != 0u;
    const unsigned gshift = (threadIdx.x & 31u) & ~(unsigned)(G - 1);
    const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << (G & 31)) - 1u) << gshift);
    const bool bad = (__ballot_sync(0xffffffffu, bad_lane) & gmask) != 0u;

    __syncwarp();
    if (lane == 0 && act) {
        // cell d of the scratch record: lane j = d>>4, register i = d&7 (chunk i>>2, word i&3), half (d>>3)&1
        const unsigned dm = best - 1u, dp = best + 1u;
        const unsigned am = (dm & 0xf0u) + ((dm & 4u) ? G * 16u : 0u) + ((dm & 3u) << 2) + ((dm >> 2) & 2u);
        const unsigned ap = (dp & 0xf0u) + ((dp & 4u) ? G * 16u : 0u) + ((dp & 3u) << 2) + ((dp >> 2) & 2u);
        const uint32_t sm1 = best > 0u ? lds16(scr + am) : 0u;
        const uint32_t sp1 = best < (unsigned)e.D - 1u ? lds16(scr + ap) : 0u;
        // every S saturated (minS = 32767): cv2's search starts from bestDisp = -1 with a strict compare, finds no winner,
        // writes INVALID and leaves disp2 alone (its disp2cost starts at SHRT_MAX and only a smaller cost replaces it)
        const bool keep = !bad && minS < 32767u;
        rec[pix] = make_uint2(minS | (keep ? best << 16 : 0xffff0000u), sm1 | (sp1 << 16));
        if (keep)
            atomicMin(disp2key + (pix - best - (unsigned)e.minD), (minS << 16) | (uint32_t)xcode);
    }
    __syncwarp();
}

template <int G, int NT, bool FINAL>
__global__ void __launch_bounds__(NT, 512 / NT)
k_wave8(const SgbmEff e, const WaveArgs a)
{
    using P = Plan<G, NT, FINAL>;
    constexpr int RB = P::RB, REC = P::REC, RECV = REC / 16, CSB = G * 16;
    constexpr int XS = FINAL ? -1 : 1;                     // memory direction of the sweep (the final sweep is flipped)
    static_assert((RB % 2) == 0, "the publication step assumes an even number of rows per band");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const unsigned sm_raw = (unsigned)__cvta_generic_to_shared(smem_raw);
    const unsigned sm0 = (sm_raw + P::ALIGN - 1) & ~(unsigned)(P::ALIGN - 1);
    unsigned char* const smem = smem_raw + (sm0 - sm_raw);
    int* const s_item = reinterpret_cast<int*>(smem + P::misc);
    const unsigned bar_addr = sm0 + P::misc + 8, sbar = sm0 + P::misc + 16;     // step barrier, stage barrier
    const int tid = (int)threadIdx.x;

    if (tid == 0) {
        *s_item = (int)atomicAdd(a.ticket, 1u);
        mbar_init(bar_addr, NT / 32);
        mbar_init(sbar, NT);                          // every thread arrives once per staged chunk
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = tid; i < 24 * REC / 16; i += NT) reinterpret_cast<uint4*>(smem + P::stage)[i] = make_uint4(0u, 0u, 0u, 0u);
    for (int i = tid; i < (128 + REC + 16) / 16; i += NT) reinterpret_cast<uint4*>(smem + P::stagemin)[i] = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();
    const int item = *s_item;
    // band-major ticket order: a consumer's producer holds an earlier ticket (see sgbm_wave.cuh)
    const int band = item / e.batch, b = item - band * e.batch;
    const int eitem = b * a.nbands + band;
    const int g = tid / G, lane = tid % G;
    const int w1 = e.width1;
    const int ys = band * RB + g;
    const bool rowact = ys < e.H;
    const int nrows = min(RB, e.H - band * RB);
    const int gl = nrows - 1;
    const bool has_above = band > 0, has_below = band < a.nbands - 1;
    const int T = w1 + 2 * (nrows - 1);
    const int yreal = FINAL ? e.H - 1 - min(ys, e.H - 1) : min(ys, e.H - 1);
    const int y0 = FINAL ? e.H - 1 - min(band * RB + RB - 1, e.H - 1) : band * RB;      // lowest image row of the band

    // ---- global streams: one 32-bit running offset (uint4 units) from three per-CTA bases -------------------------
    const size_t cta_vec = ((size_t)b * e.H + y0) * (size_t)w1 * RECV;
    const uint4* const cbase = reinterpret_cast<const uint4*>(a.cost) + cta_vec;
    uint4* const sbase = reinterpret_cast<uint4*>(a.S) + cta_vec;
    uint4* const dbgbase = (FINAL && a.S_out != nullptr) ? reinterpret_cast<uint4*>(a.S_out) + cta_vec : nullptr;
    const unsigned rowdelta = (unsigned)(yreal - y0) * (unsigned)w1 * RECV;
    const int xorg = FINAL ? w1 - 1 : 0;
    unsigned vcur = rowdelta + (unsigned)lane + (unsigned)((xorg - XS * 2 * g) * RECV);   // record of pixel xs = t - 2g
    unsigned pix = (unsigned)(((size_t)b * e.H + yreal) * e.W) + (unsigned)(e.minX1 + xorg - XS * 2 * g);
    // edge stream of this band's last row (producer side) and of the band above (consumer side); record r of
    // direction k holds pixel r + k - 1 of the edge row: indexed by the consumer's step
    uint4* const ebelow = reinterpret_cast<uint4*>(a.edge) + (size_t)eitem * 3 * w1 * RECV;
    uint32_t* const ebelow_min = a.edge_min + (size_t)eitem * w1 * 4;
    const uint4* const eabove = reinterpret_cast<const uint4*>(a.edge) + (size_t)(eitem - 1) * 3 * w1 * RECV;
    const uint4* const eabove_min = reinterpret_cast<const uint4*>(a.edge_min) + (size_t)(eitem - 1) * w1;
    const unsigned wrecv = (unsigned)w1 * RECV;

    // ---- shared-memory addresses (32-bit shared window) ---------------------------------------------------------
    const unsigned laneb = (unsigned)lane * 16u;
    unsigned rA, rB, rC, rM;
    const unsigned togA = g == 0 ? 4u * REC : 0u, togC2 = g == 0 ? 2u * REC : 0u, togC0 = g == 0 ? 6u * REC : 0u,
                   togM = g == 0 ? 64u : 0u;
    if (g > 0) {
        rA = sm0 + P::ringA + (g - 1) * 4 * REC + laneb;
        rB = sm0 + P::ringB + (g - 1) * 4 * REC + laneb;
        rC = sm0 + P::ringC + (g - 1) * 2 * REC + laneb;
        rM = sm0 + P::ringmin + (g - 1) * P::MINROW;
    } else {
        rA = sm0 + P::stage + laneb;
        rB = sm0 + P::stage + 8 * REC + laneb;
        rC = sm0 + P::stage + 16 * REC + laneb;
        rM = sm0 + P::stagemin;
    }
    const unsigned wA = sm0 + P::ringA + g * 4 * REC + laneb, wB = sm0 + P::ringB + g * 4 * REC + laneb,
                   wC = sm0 + P::ringC + g * 2 * REC + laneb, wM = sm0 + P::ringmin + g * P::MINROW;
    const unsigned zrec = sm0 + P::zero + laneb;
    const unsigned scr = sm0 + P::scratch + g * REC;
    const bool edgerow = has_below && g == gl;
    const bool wedge = has_below && (tid / 32) == (gl * G) / 32;        // this warp holds the band's last row
    const int pubthread = gl * G;
    // chunk staging (consumer side): the 12 * RECV 16-byte granules of [3 directions][4 records] are spread over the
    // CTA (NT == 16 G: the lower half of the threads moves a second granule, direction 2), four threads add the minima;
    // every thread's copies arrive on the stage mbarrier when they land (cp.async.wait_all would wait on the scoreboard
    // the global loads of the step share: 6 % of the first sweep's stall samples)
    static_assert(NT == 16 * G || NT == 32 * G, "chunk staging: 24 G granules over NT threads");
    const int ck = tid / (4 * RECV), crem = tid % (4 * RECV), cr = crem / RECV;
    const bool cvalid = tid < 12 * RECV, csecond = NT == 16 * G && tid < 4 * RECV;
    const uint4* csrc = eabove + (size_t)ck * wrecv + crem;             // chunk 0
    unsigned cdst = sm0 + P::stage + ck * 8 * REC + crem * 16;
    unsigned cdstm = sm0 + P::stagemin + (unsigned)(tid - (NT - 4)) * 16u;
    auto stage_chunk = [&](int c4) {                                    // records c4 .. c4+3 of the band above
        if (cvalid && c4 + cr < w1) {
            cp16(cdst, csrc);
            if (csecond) cp16(cdst + 16 * REC, csrc + 2 * (size_t)wrecv);
        }
        if (tid >= NT - 4 && c4 + (tid - (NT - 4)) < w1) cp16(cdstm, eabove_min + c4 + (tid - (NT - 4)));
        cp_async_arrive(sbar);
        csrc += 4 * RECV;
        cdst ^= 4u * REC;
        cdstm ^= 64u;
    };
    int known = 0;                                                     // thread 0: last progress value seen
    auto wait_progress = [&](int need) {
        if (known >= need) return;
        const int* p = a.progress + (eitem - 1);
        int v;
        while ((v = ld_acquire(p)) < need) __nanosleep(64);
        known = v;
    };
    if (has_above) {
        if (tid == 0) wait_progress(min(5, w1));
        __syncthreads();
        stage_chunk(0);
    }
    unsigned sphase = 0;                                               // parity of the stage barrier (first row's warp)

    const uint32_t p1x2 = (uint32_t)e.P1 * 0x10001u;
    const uint32_t p2m1x2 = (uint32_t)(e.P2 - e.P1) * 0x10001u;
    const unsigned magic32 = FINAL ? (unsigned)(((1ull << 32) + (unsigned)(100 - e.uniq) - 1ull) / (unsigned)(100 - e.uniq)) : 0u;
    const uint32_t k65536 = 65536u + a.zero;            // opaque to the compiler: keeps the key build an IMAD (fma pipe)
    uint32_t nopad[8];
#pragma unroll
    for (int i = 0; i < 8; i++) nopad[i] = 0u;

    uint32_t Lh[8], C[8], C2[8], S1[8], S2[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { Lh[i] = 0; C[i] = 0; C2[i] = 0; S1[i] = 0; S2[i] = 0; }
    uint32_t minh = 0, mnh = 0;                                         // horizontal direction: splat minimum, packed lane minimum
    if (rowact && g == 0) {                                             // first operands (row 0 starts at t = 0)
        ldg_rec(cbase + vcur, C, G);
        if (FINAL) ldg_rec(sbase + vcur, S1, G);
    }
    unsigned phase = 0;

    // One wavefront step.  fast: every row of this warp is strictly inside the image, has all its predecessors and
    // stays in range for the prefetches.  J = t & 3 (a constant in the fast blocks).
    auto body = [&](auto fastc, const auto jc, const int t, uint32_t (&Cc)[8], uint32_t (&Cx)[8], uint32_t (&S1c)[8], uint32_t (&S1x)[8]) {
        constexpr bool FAST = decltype(fastc)::value;
        const int J = jc;
        if (FAST || t > 0) { mbar_wait(bar_addr, phase); phase ^= 1u; }      // every warp finished step t-1's ring stores
        const int xs = t - 2 * g;
        const bool act = FAST || (rowact && (unsigned)xs < (unsigned)w1);
        if (J == 2) {
            if (has_below) {
                // the band's last row finished xdone pixels before this step; one release store is cumulative over the
                // edge stores its warp made before it arrived (xdone == w1 is published after the loop)
                const int xdone = t - 2 * gl;
                if (tid == pubthread && xdone > 0 && (xdone & (a.pubw - 1)) == 0 && xdone < w1) publish_progress(a.progress + eitem, xdone);
            }
            if (has_above && t + 2 < w1) stage_chunk(t + 2);                  // records t+2 .. t+5, first read at step t+2
        }
        if (J == 1 && has_above && tid == 0 && t + 3 < w1) wait_progress(min(t + 8, w1));     // for the chunk staged at t+1
        // the warp of the band's first row reads records t .. t+3 from now on: their chunk (staged two steps ago) landed?
        if (J == 0 && has_above && tid < 32 && t < w1) { mbar_wait(sbar, sphase); sphase ^= 1u; }

        // ---- operands of the row above (issued first: their latency hides behind the global loads) ---------
        uint32_t Lv[3][8], m0, m1, m2;
        {
            uint32_t mx;
            if (FAST) {
                lds_rec(rA + J * REC, Lv[0], CSB);
                lds_rec(rB + J * REC, Lv[1], CSB);
                lds_rec(rC + (J & 1) * REC, Lv[2], CSB);
                lds128(rM + J * 16, mx, m0, m1, m2);
            } else {
                const bool p0 = (unsigned)(xs - 1) < (unsigned)w1, p1 = (unsigned)xs < (unsigned)w1, p2 = (unsigned)(xs + 1) < (unsigned)w1;
                lds_rec(p0 ? rA + J * REC : zrec, Lv[0], CSB);
                lds_rec(p1 ? rB + J * REC : zrec, Lv[1], CSB);
                lds_rec(p2 ? rC + (J & 1) * REC : zrec, Lv[2], CSB);
                lds128(rM + J * 16, mx, m0, m1, m2);
                if (!p0) m0 = 0u;
                if (!p1) m1 = 0u;
                if (!p2) m2 = 0u;
            }
            (void)mx;
        }
        // ---- next pixel's C (and S1): registers one step ahead, L2 a few steps ahead -------------------------
        if (FAST) {
            // (the always-zero dep makes the new addresses depend on the old data: the scoreboard wait for this
            //  step's operands is then taken while only the old loads are in flight, see sgbm_wave_impl.cuh)
            const unsigned dep = (FINAL ? (Cc[0] | S1c[0]) : Cc[0]) & a.zero;
            const uint4* pc = cbase + (vcur + dep);
            ldg_rec(pc + XS * RECV, Cx, G);
            pf_l2(pc + XS * (1 + PFD) * RECV);
            pf_l2(pc + XS * (1 + PFD) * RECV + G);
            if (FINAL) {
                const uint4* ps = sbase + (vcur + dep);
                ldg_rec(ps + XS * RECV, S1x, G);
                pf_l2(ps + XS * (1 + PFD) * RECV);
                pf_l2(ps + XS * (1 + PFD) * RECV + G);
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++) { Cx[i] = 0u; S1x[i] = 0u; }
            if (rowact && (unsigned)(xs + 1) < (unsigned)w1) {
                const unsigned vn = vcur + (unsigned)(XS * RECV);
                ldg_rec(cbase + vn, Cx, G);
                if (FINAL) ldg_rec(sbase + vn, S1x, G);
            }
            if (rowact && (unsigned)(xs + 1 + PFD) < (unsigned)w1) {
                const unsigned vp = vcur + (unsigned)(XS * (1 + PFD) * RECV);
                pf_l2(cbase + vp); pf_l2(cbase + vp + G);
                if (FINAL) { pf_l2(sbase + vp); pf_l2(sbase + vp + G); }
            }
        }

        // ---- the three directions that come from the row above --------------------------------------------
        const uint32_t mn0 = step<8, G, false, true>(Lv[0], Cc, lane, p1x2, p2m1x2, m0, nopad);
        if (act) sts_rec(wA + ((J + 3) & 3) * REC, Lv[0], CSB);
        const uint32_t mn1 = step<8, G, false, true>(Lv[1], Cc, lane, p1x2, p2m1x2, m1, nopad);
        if (act) sts_rec(wB + ((J + 2) & 3) * REC, Lv[1], CSB);
        const uint32_t mn2 = step<8, G, false, true>(Lv[2], Cc, lane, p1x2, p2m1x2, m2, nopad);
        if (act) sts_rec(wC + ((J + 1) & 1) * REC, Lv[2], CSB);
        // lane minima of two directions per butterfly; the horizontal direction of the previous step rides along
        const uint32_t r12 = group_min_x2<G>(__vminu2(__byte_perm(mn0, mn1, 0x5410), __byte_perm(mn0, mn1, 0x7632)));
        const uint32_t r3h = group_min_x2<G>(__vminu2(__byte_perm(mn2, mnh, 0x5410), __byte_perm(mn2, mnh, 0x7632)));
        const uint32_t o0 = __byte_perm(r12, 0u, 0x1010), o1 = __byte_perm(r12, 0u, 0x3232), o2 = __byte_perm(r3h, 0u, 0x1010);
        minh = __byte_perm(r3h, 0u, 0x3232);
        if (act && lane == 0) {
            sts32(wM + ((J + 3) & 3) * 16 + 4, o0);
            sts32(wM + ((J + 2) & 3) * 16 + 8, o1);
            sts32(wM + ((J + 1) & 3) * 16 + 12, o2);
        }
        if (wedge) {
            if (edgerow && act) {
                // record r of direction k holds pixel r + k - 1:  pixel xs goes to records xs+1 / xs / xs-1
                uint4* const pe = ebelow + ((unsigned)xs * RECV + (unsigned)lane);
                if (FAST || xs + 1 < w1) stg_rec(pe + RECV, Lv[0], G);
                stg_rec(pe + wrecv, Lv[1], G);
                if (FAST || xs >= 1) stg_rec(pe + 2 * (size_t)wrecv - RECV, Lv[2], G);
                if (lane == 0) {
                    uint32_t* const pm = ebelow_min + (size_t)xs * 4;
                    if (FAST || xs + 1 < w1) pm[4 + 1] = o0;
                    pm[2] = o1;
                    if (FAST || xs >= 1) pm[-4 + 3] = o2;
                }
                if (!FAST) {
                    const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
                    if (xs == 0) {                         // pixel -1 of direction 0
                        pe[0] = z4; pe[G] = z4;
                        if (lane == 0) ebelow_min[1] = 0u;
                    }
                    if (xs == w1 - 1) {                    // pixel w1 of direction 2
                        pe[2 * (size_t)wrecv] = z4; pe[2 * (size_t)wrecv + G] = z4;
                        if (lane == 0) ebelow_min[(size_t)xs * 4 + 3] = 0u;
                    }
                }
            }
        }
        // this warp's part of step t is visible: arrive, and keep working on what no other row needs
        __syncwarp();
        if ((tid & 31) == 0) mbar_arrive(bar_addr);

        // ---- direction 0: along the row, state in registers ---------------------------------------------
        mnh = step<8, G, false, true>(Lh, Cc, lane, p1x2, p2m1x2, minh, nopad);
        uint32_t Sv[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const uint32_t v = Lv[0][i] + Lv[1][i] + Lv[2][i];
            Sv[i] = __viaddmin_u16x2(FINAL ? v + S1c[i] : v, Lh[i], 0x7fff7fffu);
        }
        if (FINAL) {
            wta8<G>(Sv, lane, e, magic32, scr, act, pix, 0xffff - (xorg + XS * xs), a.rec, a.disp2key, k65536);
            if (dbgbase != nullptr && act) stg_rec(dbgbase + vcur, Sv, G);
        } else {
            if (act) stg_rec_cs(sbase + vcur, Sv, G);
        }
        vcur += (unsigned)(XS * RECV);
        pix += (unsigned)XS;
        if (J == 1) rC ^= togC2;
        if (J == 3) { rA ^= togA; rB ^= togA; rC ^= togC0; rM ^= togM; }
    };

    // steps at which all rows of this warp are interior (see `fast` above)
    const int gw0 = (tid & ~31) / G, gw3 = gw0 + 32 / G - 1;
    const int tfast_lo = (gw3 < nrows) ? 2 * gw3 + 1 : 0x7fffffff;
    const int tfast_hi = w1 - 2 - PFD + 2 * gw0;                       // inclusive
    using I0 = Jc<0>; using I1 = Jc<1>; using I2 = Jc<2>; using I3 = Jc<3>;
#pragma unroll 1
    for (int t = 0; t < T; t += 4) {
        if (t >= tfast_lo && t + 3 <= tfast_hi) {
            body(std::true_type{}, I0{}, t, C, C2, S1, S2);
            body(std::true_type{}, I1{}, t + 1, C2, C, S2, S1);
            body(std::true_type{}, I2{}, t + 2, C, C2, S1, S2);
            body(std::true_type{}, I3{}, t + 3, C2, C, S2, S1);
        } else {
            // (the steps t >= T of the last block only keep the toggles and the barrier phase in step)
#pragma unroll 1
            for (int j = 0; j < 4; j += 2) {
                body(std::false_type{}, Jr{j}, t + j, C, C2, S1, S2);
                body(std::false_type{}, Jr{j + 1}, t + j + 1, C2, C, S2, S1);
            }
        }
    }
    if (has_below) {                       // the rest of the edge stream
        mbar_wait(bar_addr, phase);
        if (tid == pubthread) publish_progress(a.progress + eitem, w1);
    }
}

template <int G, int NT, bool FINAL>
static int launch(const SgbmEff& e, const WaveArgs& a, cudaStream_t st)
{
    using P = Plan<G, NT, FINAL>;
    B2S_CUDA_CHECK(cudaFuncSetAttribute(k_wave8<G, NT, FINAL>, cudaFuncAttributeMaxDynamicSharedMemorySize, P::total));
    k_wave8<G, NT, FINAL><<<e.batch * a.nbands, NT, P::total, st>>>(e, a);
    B2S_LAUNCH_CHECK();
    return B2S_OK;
}

}  // namespace wave8

// the lean kernel covers: R = 8 with every disparity cell real (D = 16 G), G = 8 or 16, the sweeps with all three
// vertical directions, sums that cannot overflow (fastacc), a uniqueness ratio below 99 (the 32-bit reciprocal of
// 100 - ratio must fit), rows long enough for chunks
bool b2s_wave8_ok(const SgbmEff& e, const WaveArgs& a, int vert)
{
    static const bool off = [] { const char* s = getenv("B2S_WAVE8"); return s && s[0] == '0'; }();
    if (off) return false;
    return vert && a.diag && a.fastacc && e.R == 8 && e.D == e.Dp && (e.G == 8 || e.G == 16) && e.uniq < 99 &&
           e.width1 >= 16 && e.uqle == 0;
}

// rows per band (= threads / G).  G = 8: 32 rows (two CTAs of eight warps per SM, half the edge traffic of 16-row bands
// and half as many bands in a pair's pipeline); measured per 128 KITTI pairs 16.54 ms against 16.9 ms for 16 rows, and
// 4.44 against 4.64 ms for a single pair.  B2S_WAVE8_ROWS=16 selects the 16-row variant for A/B measurements.
static int wave8_rows8()
{
    static const int rows = [] { const char* s = getenv("B2S_WAVE8_ROWS"); return (s && atoi(s) == 16) ? 16 : 32; }();
    return rows;
}
int b2s_wave8_rows_per_band(const SgbmEff& e) { return e.G == 8 ? wave8_rows8() : 16; }

int b2s_launch_wave8(const SgbmEff& e, const WaveArgs& a, int final, cudaStream_t st)
{
    if (e.G == 8 && wave8_rows8() == 32) return final ? wave8::launch<8, 256, true>(e, a, st) : wave8::launch<8, 256, false>(e, a, st);
    if (e.G == 8) return final ? wave8::launch<8, 128, true>(e, a, st) : wave8::launch<8, 128, false>(e, a, st);
    return final ? wave8::launch<16, 256, true>(e, a, st) : wave8::launch<16, 256, false>(e, a, st);
}
