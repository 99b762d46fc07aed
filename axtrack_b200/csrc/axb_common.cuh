// Shared declarations of the axtrack_b200 CUDA translation units (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include "../../include/axtrack_b200.h"

namespace axb {

// ---- K2 geometry ---------------------------------------------------------------------------
// Search window of one source: the L1 ball of radius RW = max_px around it (cells at hop
// max_px are detected, cells at hop <= max_px-1 are expanded).  The diamond is stored FOLDED:
// rows dy <= 0 live in folded row dy+RW centred at column 512, rows dy > 0 share folded row
// dy-1 centred at column 0 (mod 1024).  Both halves of a folded row together hold 2*RW cells,
// so the whole window is (RW+1) rows x 1024 bits: 64 KB per bitmap at max_px = 500 instead of
// 125 KB for the bounding box.
constexpr int kRowBits = 1024;
constexpr int kRowWords = 32;
constexpr int kK2Threads = 256;
constexpr int kQCap = 4096;          // frontier entries kept in shared memory per buffer
constexpr int kTCap = 2048;          // in-range targets per pass kept in shared memory
constexpr int kGCap = 1 << 19;       // per-CTA global overflow entries (>= cells of the window)
constexpr int kSeedEnd = 512;        // entries of the per-level seed calendar (j <= max_px)
constexpr uint32_t kUnset = 0xFFFFFFFFu;

struct K2Params {
    const uint32_t* mask_bits;   // [n_masks][H][Wp]
    int H, W, Wp;
    const int32_t* dets_yx;      // [n][2] (y, x)
    const axb_pair_desc* pairs;  // device
    int n_pairs;
    const int* work_prefix;      // [n_pairs+1] exclusive prefix of n_src (device)
    int max_px;
    int32_t* path_len;
    uint8_t* unreachable;
    float* wcost;
    uint32_t* scratch;           // per CTA: 4 * kGCap words
    int* work_counter;           // atomic ticket
    unsigned long long* err_key; // atomicMin: (work item << 26 | target << 2 | code)
};

inline size_t k2_smem_bytes(int max_px) {
    const size_t nrows = (size_t)max_px + 1;
    return (2 * nrows * kRowWords + 2 * kQCap + 2 * kSeedEnd + 3 * kTCap) * sizeof(uint32_t);
}

void launch_k2(const K2Params& p, int grid, cudaStream_t stream);
void launch_work_prefix(const axb_pair_desc* pairs, int n_pairs, int* prefix, cudaStream_t stream);
void launch_pack_mask_u8(const uint8_t* mask, int n_masks, int H, int W, uint32_t* bits, cudaStream_t s);
void launch_pack_mask_f32(const float* w, int n_masks, int H, int W, uint32_t* bits,
                          unsigned long long* err_flags, cudaStream_t s);

// ---- K1 ------------------------------------------------------------------------------------
struct K1Params {
    const float* Y;              // [F][ntiles][Sx][Sy][3]
    const int32_t* tile_yx;      // [ntiles][2]
    int F, ntiles, Sx, Sy, tilesize;
    float conf_thr;
    int nms_d2;                  // nms_dist^2
    int cap;
    float* pre_conf; int32_t* pre_yx; int32_t* pre_tile; int32_t* n_pre;
    float* conf; int32_t* yx; int32_t* keep_idx; int32_t* n_out;
    uint8_t* scratch;            // per frame workspace
    size_t scratch_per_frame;
    int npad;                    // power of two >= ncell (sort / hash size)
    unsigned long long* err_flags;  // atomicOr: bit 2 = capacity
};
size_t k1_scratch_per_frame(int ncell, int npad);
void launch_k1(const K1Params& p, cudaStream_t stream);

}  // namespace axb
