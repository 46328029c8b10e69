// sgbm_wave.cuh -- fused multi-direction SGBM path aggregation as a pixel wavefront
// (SURVEY.md A.2/A.3).  Replaces the two passes of cv2.StereoSGBM.compute (reference call site
// stereo_vision/stereo_vision.py:178):
//
//   pass 1 (forward sweep) : directions  -> , down-right, down, down-left   summed into S1
//   pass 2 (flipped sweep) : directions  <- , up-left,   up,   up-right     S = S1 + ..., then
//                            winner-take-all + uniqueness + disp2 scatter fused on the registers
//   MODE_SGBM              : pass 1, then a horizontal-only flipped sweep (<-) with the same WTA
//   MODE_HH4               : both sweeps without their diagonal directions (-> , down | <- , up)
//
// so the cost volume is read once per pass and S is written once and read once: 10 B per cost
// cell of HBM traffic including the cost kernel's write (the reference algorithm's own 8 B plus
// one re-read of C), instead of one read-modify-write of S per direction.
//
// Decomposition.  A CTA owns a band of RB consecutive rows (in sweep order) of one pair; a
// group of G lanes owns one row and walks it pixel by pixel, lane j holding disparities
// [2R*j, 2R*j + 2R) as packed u16x2.  Rows are skewed by two pixels per row: at step t the
// group of row g is at x = t - 2g, so the three predecessors in the row above, (x-1), (x),
// (x+1), were finished 3, 2 and 1 steps earlier.  They travel through small shared-memory rings;
// the horizontal direction never leaves the registers.  One split-phase mbarrier per step (a warp arrives as soon as its ring stores are done, waits at the start of the next step).
// Bands of the same pair run concurrently as a pipeline: the last row of a band streams its
// three vertical L_r vectors to an HBM edge buffer, the next band stages them into shared
// memory chunk-wise with cp.async, gated by a release/acquire progress counter.  Bands are
// handed out through an atomic ticket in (pair, band) order, so a consumer can only ever wait
// for a CTA that has already been scheduled (no co-residency assumption, no deadlock).
#pragma once
#include "common.cuh"

struct WaveArgs {
    const uint16_t* cost;   // [batch][H][width1][Dp]  raw block cost
    uint16_t* S;            // pass 1: S1 out ; final pass: S1 in
    uint16_t* S_out;        // final pass: full S (debug tap) or nullptr
    uint16_t* edge;         // [batch*nbands][width1][3][Dp]
    uint32_t* edge_min;     // [batch*nbands][width1][4]   (min L_r, replicated in both halves)
    int* progress;          // [batch*nbands]  pixels of the band's last row published so far
    unsigned int* ticket;   // work-item dispenser
    uint2* rec;             // final pass: [batch][H][W] WTA records
    uint32_t* disp2key;     // final pass: [batch][H][W] (cost << 16 | 0xffff - x) scatter-min keys
    unsigned long long magic;  // ceil(2^40 / (100 - uniquenessRatio)), 0 if the ratio is >= 100
    int flip;               // 0: forward sweep, 1: both axes reversed
    int fastacc;            // S can be summed without per-direction saturation (see k_wave, FA)
    int nbands;
    int diag;               // 1: the sweeps carry the two diagonal directions (MODE_SGBM / MODE_HH), 0: MODE_HH4
    unsigned int zero;      // always 0, opaque to the compiler (see `dep` in k_wave)
    int pubw;               // k_wave8: pixels between two publications of a band's progress (a power of two)
};

// Threads per CTA of the wavefront kernel (a CTA owns threads / G rows): small CTAs keep the per-step barrier cheap
// and let several independent bands share an SM.  With G = 8 lanes per pixel both 256 (32 rows) and 128 (16 rows) are
// built; measured on the B200 per 128 KITTI-shape pairs: MODE_HH 19.36 / 19.05 ms, MODE_HH4 14.68 / 15.11,
// MODE_SGBM_3WAY (stripe images of 107 rows: 128 / 112 rows computed) 13.16 / 12.38, MODE_SGBM 640x480 D = 64
// 10.12 / 10.36 ms per step; 64 threads: 29.8 ms (four times the edge traffic).
int b2s_wave_threads(const SgbmEff& e);
int b2s_wave_rows_per_band(const SgbmEff& e);
size_t b2s_wave_edge_bytes(const SgbmEff& e);       // edge + edge_min + progress + ticket
int b2s_launch_wave(const SgbmEff& e, const uint16_t* cost, uint16_t* S, uint16_t* S_dbg, void* edge_ws,
                    uint2* rec, uint32_t* disp2key, cudaStream_t st, cudaEvent_t between_sweeps = nullptr);
int b2s_launch_wave_g8(const SgbmEff& e, const WaveArgs& a, int vert, int final, cudaStream_t st);        // 256 threads
int b2s_launch_wave_g8s(const SgbmEff& e, const WaveArgs& a, int vert, int final, cudaStream_t st);       // 128 threads
int b2s_launch_wave_g16(const SgbmEff& e, const WaveArgs& a, int vert, int final, cudaStream_t st);
int b2s_launch_wave_g32(const SgbmEff& e, const WaveArgs& a, int vert, int final, cudaStream_t st);
// the lean kernel for R = 8, D = 16 G (sgbm_wave8.cu): sweeps with the three vertical directions
bool b2s_wave8_ok(const SgbmEff& e, const WaveArgs& a, int vert);
int b2s_wave8_rows_per_band(const SgbmEff& e);
int b2s_launch_wave8(const SgbmEff& e, const WaveArgs& a, int final, cudaStream_t st);
