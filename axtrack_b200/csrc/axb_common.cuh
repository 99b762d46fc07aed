// Shared declarations of the axtrack_b200 CUDA translation units (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include "../../include/axtrack_b200.h"

namespace axb {

// ---- K2 geometry ---------------------------------------------------------------------------
// Search window of one source: the L1 ball of radius RW = max_px around it (cells at hop
// max_px are detected, cells at hop <= max_px-1 are expanded).  The diamond is stored FOLDED
// on a circle of 1024 columns: rows dy <= 0 live in folded row dy+RW with their centre at
// column 512, rows dy > 0 share folded row dy-1 with their centre at column 0; columns are
// measured from the 32-pixel-aligned word that holds the source (cx = x - (sx & ~31)), so one
// shared-memory word is one aligned word of the packed mask.  Both halves of a folded row
// together hold 2*RW <= 1022 cells, so the window is (RW+1) rows x 1024 bits: 64 KB per
// bitmap at max_px = 500 instead of 125 KB for the bounding box.
constexpr int kRowBits = 1024;
constexpr int kRowWords = 32;
constexpr int kMaxChain = 4;          // frame pairs that share one source search (forward members of a chain)
constexpr int kMaxRev = 4;            // pairs answered from their TARGET side by one search (reverse members, k2_prep_kernel "cover")
constexpr int kMaxMem = kMaxChain + kMaxRev;
constexpr int kHashSize = 1024;       // target-cell hash (u32 entries) in shared memory
constexpr int kHashMaxLoad = 640;     // candidates per search pass
constexpr int kCandCap = 2048;        // per-CTA global candidate table (>= the hash loads)
constexpr int kK2Threads = 256;       // wide all-levels kernel: one CTA (8 warps) per source search
constexpr int kGenThreads = 512;      // all-levels kernel (one CTA per SM): 16 warps per source search
constexpr int kSeedCal = 512;         // general kernel: per-j seed calendar entries (j <= max_px)
constexpr int kSeedCalWide = 1024;
constexpr int kFastMaxPx = 511;       // largest radius of the 1024-column folded window (tile-word kernels, shared-memory bitmaps)
static_assert(AXB_MAX_PX == 1023, "wide layout: 2048-bit folded rows");
// fast kernel (level 0 only): 3 CTAs per SM
#ifndef AXB_FAST_THREADS
#define AXB_FAST_THREADS 96
#endif
constexpr int kFastThreads = AXB_FAST_THREADS;      // fast kernels: one CTA (3 warps) per source search, one thread per frontier entry
constexpr int kFastGQCap = 2048;      // per-CTA global overflow frontier entries per buffer
// dense window (every tile word of the folded diamond, 64.5 KB): 3 CTAs per SM
constexpr int kFastQCap = 384;        // frontier entries (tile word, bits: 8 B) per buffer in shared memory
constexpr int kFastCtasPerSm = 3;
// sparse window (only the non-zero tile words + a presence/base table): 10 CTAs per SM
constexpr int kSparseWords = 3896;    // non-zero tile words a sparse window can hold (15.25 KB; with the tables, queues and control block 10 CTAs fill an SM at max_px = 500)
constexpr int kSparseQCap = 144;
constexpr int kSparseHash = 512;
constexpr int kSparseHashMaxLoad = 320;
constexpr int kSparseCtasPerSm = 10;
// sparse window on the 256-column circle (max_px 512 .. 1023): 2 CTAs per SM
constexpr int kWideWords = 19456;     // non-zero tile words a wide sparse window can hold (76 KB; the whole diamond: 65 792)
constexpr int kWideQCap = 384;
constexpr int kWideSparseCtasPerSm = 2;
constexpr int kWideHash = 2048;       // target hash of the wide sparse kernel (shared memory): a radius of 1000 px holds four times the targets,
constexpr int kWideHashMaxLoad = 1280; // and the chains of the cover twice as many again; a second pass is a second search
// general kernel (all levels): both bitmaps, 1 CTA per SM
constexpr int kGenQCap = 2048;
constexpr int kGCap = 1 << 19;        // >= cells of the window: worst-case frontier / seed lists
constexpr int kGCapWide = 1 << 21;
constexpr int kWideCtasPerSm = 4;     // wide all-levels kernel: bound by the latency of L2 atomics, 28 KB of shared memory
constexpr uint32_t kEmpty = 0xFFFFFFFFu;

struct K2Params {
    const uint32_t* mask_bits;   // [n_masks][H][Wp]
    int H, W, Wp;
    const uint32_t* mask_tiles;  // [n_masks][TH][TWp] 8x4-pixel tile words (bit = (y&3)*8 + (x&7)), built per call
    int TH, TWp;                 // TH = ceil(H/4), TWp = 4*Wp
    const int32_t* dets_yx;      // [n][2] (y, x)
    const axb_pair_desc* pairs;  // device
    int n_pairs;
    const int* work_prefix;      // [n_pairs+1] exclusive prefix of n_src over chain LEADERS
    const int* all_prefix;       // [n_pairs+1] exclusive prefix of n_src over all pairs (error order)
    const int* next_same;        // [n_pairs] next pair with the same (mask, sources), or -1
    int max_px;
    int hash_smem;               // k2_sparse_kernel: target hash in shared memory (set by launch_k2_sparse)
    int sparse_words;            // k2_sparse_kernel: non-zero tile words a window can hold (set by launch_k2_sparse)
    int32_t* path_len;
    uint8_t* unreachable;
    float* wcost;
    uint32_t* fast_scratch;      // per fast CTA: kCandCap + 2*kFastGQCap words
    uint32_t* gen_scratch;       // per general CTA: kCandCap + 4*kGCap words
    int* counters;               // [0] sparse ticket, [1] general list length, [2] general ticket, [3] dense list length, [4] dense ticket
    int* gen_list;               // work items deferred to the general kernel
    int* dense_list;             // work items whose window does not fit the sparse layout
    // path tracing (k2_general_kernel only; NULL otherwise): one (source, target) request per pair
    uint8_t* trace_par;          // per general CTA kTraceParBytes: direction a window cell was entered from
    int32_t* trace_yx;           // [n_pairs][max_px][2] (row, col), start -> goal
    int32_t* trace_len;          // [n_pairs] pixels on the path, 0 = None
    unsigned long long* err_key; // atomicMin: (item in reference order << 26 | target << 2 | code)
    long long orig_base;         // sources of the pairs that precede this call's pairs in the caller's batch (chunked host entry)
    // cover (k2_prep_kernel): which sets are searched, and the pairs answered from their target side
    const int* p1_prefix;        // [n_pairs+1] exclusive prefix of n_src over the chain leaders of SEARCHED sets: the tickets of the first pass
    const int* rev_head;         // [n_pairs] head pair of a searched set -> its first reverse member, or -1
    const int* rev_next;         // [n_pairs] next reverse member
    const int* lead_of;          // [n_pairs] leader of the chain a pair belongs to (its directed work items)
    uint32_t* item_bits;         // bitmap over the work items: already on the directed list
    int* dir_list;               // work items that need their own (directed) search after all
    const int32_t* comp_label;   // [n_masks][H][W] root pixel of the mask component, -1 off the mask (axb_components.cu); NULL: no cover
    const int4* comp_ext;        // [n_masks][H * W] L1 extents of the component, at its root pixel
    int reverse;                 // this launch attaches the reverse members (first pass and its dense list)
    // where a fast kernel takes its work items from and where it puts the windows that do not fit
    const int* src_list;         // NULL: tickets over p1_prefix
    int src_len_ctr, src_ticket_ctr;
    int* dense_out;
    int dense_out_ctr;
    // the all-levels kernel's work list
    const int* gen_src;
    int gen_len_ctr, gen_ticket_ctr;
    int early;                   // in-grid off-mask detections were handed to the all-levels kernel before the level-0 kernels ran
};

// outputs of k2_prep_kernel (each [n_pairs + 1] ints)
struct K2Prep {
    int *work_prefix, *all_prefix, *next_same, *p1_prefix, *rev_head, *rev_next, *lead_of;
    int *dst_head, *state;       // scratch of the cover decision
    uint32_t* item_bits;         // [(total_src + 31) / 32]
};
constexpr int kPrepArrays = 9;

// counters[]: [0] first-pass ticket, [1] general list length, [2] general ticket, [3] dense list A length, [4] its ticket,
// [5] directed list length, [6] its ticket, [7] dense list B length, [8] its ticket, [9] early list length, [10] its ticket
enum { kCtrP1Ticket = 0, kCtrGenLen = 1, kCtrGenTicket = 2, kCtrDenseALen = 3, kCtrDenseATicket = 4, kCtrDirLen = 5,
       kCtrDirTicket = 6, kCtrDenseBLen = 7, kCtrDenseBTicket = 8, kCtrEarlyLen = 9, kCtrEarlyTicket = 10 };

inline int k2_tile_rows(int max_px) { return (max_px + 3) / 4 + 1; }
inline size_t k2_fast_smem_bytes(int max_px) {      // dense window + queues + hash
    return ((size_t)k2_tile_rows(max_px) * 128 + 4 * kFastQCap + kHashSize) * sizeof(uint32_t);
}
inline size_t k2_sparse_smem_bytes(int max_px) {    // packed words + (presence, base) per 32 tile columns + queues + hash
    return ((size_t)kSparseWords + (size_t)k2_tile_rows(max_px) * 8 + 4 * kSparseQCap) * sizeof(uint32_t);   // the target hash lives in the CTA's global scratch
}
inline size_t k2_sparse_wide_smem_bytes(int max_px) {   // packed words + (presence, base) per 32 of 256 tile columns + queues
    return ((size_t)kWideWords + (size_t)k2_tile_rows(max_px) * 16 + 4 * kWideQCap) * sizeof(uint32_t);
}
inline size_t k2_gen_smem_bytes(int max_px) {
    return (2 * (size_t)(max_px + 1) * kRowWords + 2 * kGenQCap + kHashSize + 2 * kSeedCal) * sizeof(uint32_t);
}
inline size_t k2_gen_wide_smem_bytes() {            // bitmaps in global scratch
    return (2 * (size_t)kGenQCap + kHashSize + 2 * kSeedCalWide) * sizeof(uint32_t);
}
constexpr size_t kFastScratchWords = (size_t)kCandCap + 4 * (size_t)kFastGQCap + kSparseHash + kCandCap / 4;   // candidates | overflow queues | sparse target hash | cell bit of a candidate (bytes)
constexpr size_t kTraceParBytes = (size_t)(kFastMaxPx + 1) * kRowBits;
constexpr size_t kTraceParBytesWide = (size_t)(AXB_MAX_PX + 1) * 2 * kRowBits;
constexpr size_t kGenScratchWords = (size_t)kCandCap + 4 * (size_t)kGCap;
constexpr size_t kGenScratchWordsWide = (size_t)kCandCap + 4 * (size_t)kGCapWide + 2 * (size_t)(AXB_MAX_PX + 1) * 2 * kRowWords;

// Function attributes are per device: true the first time the calling thread's current device is
// seen for `flags` (one 64-bit word per kernel), so that the opt-in shared-memory size is set once
// on every device a process uses.
inline bool first_use_on_device(unsigned long long& flags) {
    int dev = 0;
    cudaGetDevice(&dev);
    const unsigned long long bit = 1ull << (dev & 63);
    if (flags & bit) return false;
    flags |= bit;
    return true;
}

void launch_k2_prep(const axb_pair_desc* pairs, int n_pairs, const K2Prep& o, int* counters, int n_masks, int n_dets,
                    long long out_capacity, int total_src, int allow_reverse, unsigned long long* err_flags, cudaStream_t stream);
void launch_k2_components(const uint32_t* mask_bits, int n_masks, int H, int W, int Wp, int* labels, int4* ext, cudaStream_t stream);
void launch_k2_tiles(const uint32_t* mask_bits, int n_masks, int H, int Wp, uint32_t* tiles, cudaStream_t stream);
void launch_k2_sparse(const K2Params& p, int grid, cudaStream_t stream);
void launch_k2_fast(const K2Params& p, int grid, cudaStream_t stream);
void launch_k2_sparse_wide(const K2Params& p, int grid, bool second, cudaStream_t stream);   // max_px 512 .. 1023
void launch_k2_classify(const K2Params& p, int* early, int total_src, cudaStream_t stream);
void launch_k2_fill_reverse(const K2Params& p, const int* pair_searched, const int* state, cudaStream_t stream);
void launch_k2_general(const K2Params& p, int grid, cudaStream_t stream);     // max_px > kFastMaxPx: the wide kernel
void launch_k2_all_to_general(const int* work_prefix, int n_pairs, int* gen_list, int* counters, cudaStream_t stream);
void launch_pack_mask_u8(const uint8_t* mask, int n_masks, int H, int W, uint32_t* bits, cudaStream_t s);
void launch_pack_mask_f32(const float* w, int n_masks, int H, int W, uint32_t* bits,
                          unsigned long long* err_flags, cudaStream_t s);

void launch_k3(const int32_t* path_len, const axb_pair_desc* pairs, int n_pairs, const double* cost_tab,
               const long long* arc_tab, int n_gaps, int max_px, double* cost, long long* arc,
               unsigned long long* err_flags, int max_elems_per_pair, cudaStream_t stream);

void launch_k5(const float* images, int n_frames, int H, int W, const int32_t* boxes, const int32_t* box_frame, int n_boxes,
               int bins, float lo, float hi, float* out, cudaStream_t stream);
void launch_k7(const float* f_src, int n_src, const float* f_dst, int n_dst, int bins, const int32_t* path_len, int max_px,
               double* vis, cudaStream_t stream);

constexpr int kK6FlagStride = 8;      // uint32 words between two K6 vote flags: one flag per 32-byte sector
void launch_k6_nonempty(const float* X, int T, int C, int H, int W, int ts, uint32_t* flags, uint8_t* nonempty, cudaStream_t stream);
cudaError_t launch_k6_nonempty_static(const float* X, int T, int C, int H, int W, int ts, uint8_t* nonempty, cudaStream_t stream);   // flags in a static device array
void launch_k6_gather(const float* X, int T, int C, int H, int W, int ts, const int32_t* tile_yx, int ntiles,
                      const int32_t* frames, int B, int tc, int chan_lo, int nch, float* out,
                      unsigned long long* err_flags, int n_sms, cudaStream_t stream);

void launch_k6_frame(const float* X, int C, int H, int W, int ts, const uint8_t* keep, int frame, int chan_lo, int nch,
                     float* out, int n_sms, cudaStream_t stream);

// ---- K1 ------------------------------------------------------------------------------------
struct K1Params {
    const float* Y;              // [F][ntiles][Sx][Sy][3]
    const int32_t* tile_yx;      // [ntiles][2]
    int F, ntiles, Sx, Sy, tilesize;
    float conf_thr;
    int nms_d2;                  // nms_dist^2
    int cap;
    float* pre_conf; int32_t* pre_yx; int32_t* pre_cell; int32_t* n_pre;
    float* conf; int32_t* yx; int32_t* keep_idx; int32_t* n_out;
    uint8_t* scratch;            // per frame workspace
    size_t scratch_per_frame;
    int npad;                    // power of two >= ncell (sort / hash size)
    unsigned long long* err_flags;  // atomicOr: bit 2 = capacity
    int* wide_ctl;                  // wide kernel: 8 ints of grid-wide control (zeroed by the launcher)
    int num_sms;
};
size_t k1_scratch_per_frame(int ncell, int npad, int ntiles);
void launch_k1(const K1Params& p, cudaStream_t stream);
bool launch_k1_wide(const K1Params& p, cudaStream_t stream);      // false: not applicable, use the one-CTA-per-frame kernel

}  // namespace axb
