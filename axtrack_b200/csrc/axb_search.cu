// K2 -- batched, radius-bounded, multi-target grid shortest-path search (sm_100a).
//
// Replaces the per-pair A* of the reference (pyastar2d/src/cpp/astar.cpp:46-172 driven by
// axtrack/AxonDetections.py:553-579,600-629): ONE bounded search per SOURCE detection answers
// every target of that source -- and, when several frame pairs share the same source frame and
// mask (t_bef is the predecessor of both t_bef+1 and t_bef+2, AxonDetections.py:560), every
// target of all of them ("chain", up to kMaxChain pairs).
//
// Semantics (DESIGN.md "K2 semantics"; the closed form oracle/astar_oracle.cpp implements and
// tests/test_oracle_golden.py pins to the compiled reference).  Weights are 1 on the mask and
// 65536 off it (AxonDetections.py:598), so the exact cost of a path is n_fg + 65536*n_bg over the
// pixels it enters and its hop count is n_fg + n_bg.  Cells are settled in lexicographic
// (n_bg, n_fg) order: "level" k = n_bg, and inside a level a breadth-first wave over mask pixels
// indexed by j = n_fg.  Off-mask neighbours of level-k cells seed level k+1 with the j of the
// cell that touched them first.  A cell at hop == max_px is never expanded (the reference pushes
// it with path_length max_px+1, astar.cpp:136, and aborts when it reaches the top of the heap,
// astar.cpp:88-92).  The search stops after the first level k* that contains such a cell: every
// target of a higher level is "None" because that heap entry has a smaller priority than any
// goal of a higher level (65536 >> 2*max_px).  Targets settled at (k, j), k <= k*, get
// path_length = k + j + 1 if that is <= max_px.
//
// Mapping to the SM.  A search is ~max_px dependent wavefronts of a few hundred cells: latency
// bound.  The level-0 searches (nearly all sources on real data) run on the tile-word kernels of
// axb_search_fast.cu; this file holds
//   k2_general_kernel       all levels, one CTA of kGenThreads (16 warps) per source, one lane per
//                           (frontier cell, direction): per-cell bitmaps "open mask pixel" and
//                           "open off-mask pixel" in shared memory (folded-diamond layout of
//                           axb_common.cuh, rows rotated by their index over the 32 banks),
//                           per-level seed lists in global memory with a per-j calendar in shared
//                           memory; for the sources the fast kernels defer (level 0 exhausted
//                           below hop max_px with targets still open: source off the mask or in
//                           a small component) and for path tracing (K4).
//   k2_general_wide_kernel  the same search for 512 <= max_px <= 1023: 2048-bit folded rows,
//                           bitmaps in the CTA's global scratch, every source.
//   k2_prep_kernel          chains, work prefix sums, counter reset (one CTA, tiny).
#include "axb_common.cuh"
#include <algorithm>

namespace axb {

namespace {

constexpr unsigned kFull = 0xFFFFFFFFu;

__device__ __forceinline__ uint32_t span_mask(int a, int b) {
    // bits i in [a, b] clipped to [0, 31]
    a = a < 0 ? 0 : a;
    b = b > 31 ? 31 : b;
    if (a > b) return 0u;
    return (0xFFFFFFFFu >> (31 - b)) & (0xFFFFFFFFu << a);
}

__device__ __forceinline__ uint32_t lanemask_lt() {
    uint32_t m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// Window geometry of the all-levels kernel.  Up to max_px = 511 the folded rows are 1024 bits and
// both bitmaps live in shared memory; the WIDE variant (512 <= max_px <= 1023: the reference honours
// any max_px_assoc_dist, AxonDetections.py:623-625,736) folds on 2048-bit rows and keeps the bitmaps
// in the CTA's global scratch (512 KB, L2 resident).
template <bool WIDE> struct Geo {
    static constexpr int LOG = WIDE ? 11 : 10;       // log2(bits of a folded row)
    static constexpr int RB = 1 << LOG, RWD = RB / 32, HALF = RB / 2;
    static constexpr int CSH = LOG + 1;              // queue entry: (dy + HALF) << CSH | (cx + HALF)
    static constexpr int ESH = WIDE ? 11 : 13;       // hash entry: cell index << ESH | candidate slot
    static constexpr int CAL = WIDE ? kSeedCalWide : kSeedCal;
    static constexpr int GQ = WIDE ? kGCapWide : kGCap;
    static constexpr size_t SCRATCH = WIDE ? kGenScratchWordsWide : kGenScratchWords;
    static constexpr size_t PAR = WIDE ? kTraceParBytesWide : kTraceParBytes;
};
static_assert(kCandCap <= (1 << Geo<true>::ESH) && kCandCap <= (1 << Geo<false>::ESH), "hash entries keep ESH bits of candidate slot");

// queue entry; cx = x - (sx & ~31) in [-(HALF-1), HALF+30]
template <bool WIDE> __device__ __forceinline__ uint32_t pack_cell(int dy, int cx) {
    return ((uint32_t)(dy + Geo<WIDE>::HALF) << Geo<WIDE>::CSH) | (uint32_t)(cx + Geo<WIDE>::HALF);
}

// logical cell index of the folded window: row * RB + column (the hash key of a target)
template <bool WIDE> __device__ __forceinline__ int fold_index(int dy, int cx, int RW) {
    const int row = dy <= 0 ? dy + RW : dy - 1;
    const int col = (cx + (dy <= 0 ? Geo<WIDE>::HALF : 0)) & (Geo<WIDE>::RB - 1);
    return row * Geo<WIDE>::RB + col;
}

// bitmap word of a logical cell index: the words of row r are rotated by r (shared-memory banks)
template <bool WIDE> __device__ __forceinline__ int phys_word(int idx) {
    const int row = idx >> Geo<WIDE>::LOG;
    return row * Geo<WIDE>::RWD + (((idx >> 5) + row) & (Geo<WIDE>::RWD - 1));
}

__device__ __forceinline__ uint32_t hash_slot(uint32_t key) { return (key * 0x9E3779B1u) >> 22; }
static_assert(kHashSize == 1024, "hash_slot keeps 10 bits");

struct Src {
    const uint32_t* mbits;
    int H, W, Wp, sy, sx, iw0, o, RW;
};

// OR the bits of packed-mask word iw of image row y, restricted to x in [xlo, xhi], into fg/bg
__device__ __forceinline__ void gather(const Src& s, int y, int iw, int xlo, int xhi, uint32_t& fg, uint32_t& bg) {
    if (iw < 0 || iw >= s.Wp) return;
    const uint32_t valid = span_mask(xlo - 32 * iw, xhi - 32 * iw);
    if (!valid) return;
    const uint32_t m = __ldg(s.mbits + (size_t)y * s.Wp + iw);
    fg |= m & valid;
    bg |= ~m & valid;
}

// Stage the folded window: one warp per folded row, lane = logical word of the row (two passes
// for the 64-word rows of the WIDE layout).  One bitmap word is one aligned word of the packed
// mask: the upper half (dy <= 0) of folded row r sits at columns cx+HALF, the lower half
// (dy = r+1) at columns cx (mod RB); only word 0 / word RWD/2 can see a second, wrapped word of
// the same half.
template <bool GEN, int NT, bool WIDE>
__device__ __forceinline__ void stage_window(const Src& s, uint32_t* OF, uint32_t* OB) {
    constexpr int RWD = Geo<WIDE>::RWD;
    const int RW = s.RW;
    const int lane = threadIdx.x & 31;
    #pragma unroll 2
    for (int r = threadIdx.x >> 5; r <= RW; r += NT / 32) {
        #pragma unroll
        for (int wl = lane; wl < RWD; wl += 32) {
            uint32_t fg = 0, bg = 0;
            {
                const int y = s.sy + r - RW;                       // upper half, half width r
                if (y >= 0 && y < s.H) {
                    const int xlo = max(0, s.sx - r), xhi = min(s.W - 1, s.sx + r);
                    gather(s, y, s.iw0 - RWD / 2 + wl, xlo, xhi, fg, bg);
                    if (wl == 0) gather(s, y, s.iw0 + RWD / 2, xlo, xhi, fg, bg);
                }
            }
            if (r + 1 <= RW) {
                const int y = s.sy + r + 1;                        // lower half, half width RW-r-1
                if (y >= 0 && y < s.H) {
                    const int hw = RW - r - 1;
                    const int xlo = max(0, s.sx - hw), xhi = min(s.W - 1, s.sx + hw);
                    gather(s, y, s.iw0 + (wl < RWD / 2 ? wl : wl - RWD), xlo, xhi, fg, bg);
                    if (wl == RWD / 2) gather(s, y, s.iw0 + RWD / 2, xlo, xhi, fg, bg);
                }
            }
            const int pw = r * RWD + ((wl + r) & (RWD - 1));
            OF[pw] = fg;
            if (GEN) OB[pw] = bg;
        }
    }
}

struct Chain {
    long long orow[kMaxChain];      // first output element of this source's row, per member pair
    long long item_orig[kMaxChain]; // work index in the reference's iteration order (errors)
    int dst_off[kMaxChain];
    int cum[kMaxChain + 1];         // flattened target index space
    int nm;
};

// per-CTA search state in static shared memory
struct Ctl {
    Chain ch;
    int cnt[3];        // frontier sizes, rotating: wave s reads cnt[s%3], appends to cnt[(s+1)%3]
    int unsettled;     // in-range targets not yet reached
    int ncand;
    int seedadd[3];    // general: seeds appended for the next level per wave, rotating like cnt[]
    int settled[3];    // targets reached per wave, rotating like cnt[]
    int item;
};

// a target cell was reached at level k with j mask pixels on the path: hop = k + j
template <bool WIDE>
__device__ __forceinline__ int probe_settle(const K2Params& p, const uint32_t* hash, const uint32_t* cand,
                                            const Chain& ch, uint32_t key, int k, int j) {
    constexpr int ESH = Geo<WIDE>::ESH;
    int n = 0;
    uint32_t h = hash_slot(key);
    for (;;) {
        const uint32_t e = hash[h];
        if (e == kEmpty) break;
        if ((e >> ESH) == key) {
            const uint32_t c = cand[e & ((1u << ESH) - 1u)];
            const long long o = ch.orow[c >> 30] + (long long)(c & 0x3FFFFFFFu);
            p.path_len[o] = k + j + 1;
            p.unreachable[o] = 0;
            if (p.wcost) p.wcost[o] = (float)(k * 65536 + j);
            ++n;
        }
        h = (h + 1) & (kHashSize - 1);
    }
    return n;
}

// One work item = one source of a chain leader, processed by the whole CTA.  Returns true when
// the item must be redone by the general kernel (fast kernel only).  All control flow is
// CTA-uniform (decisions are read from shared memory after a barrier).
template <bool GEN, int NT, bool WIDE>
__device__ bool run_item(const K2Params& p, const int item, uint32_t* smem, uint32_t* gscr, Ctl& ctl) {
    using G = Geo<WIDE>;
    constexpr int kCal = G::CAL;
    static_assert(NT <= kHashMaxLoad, "a group of NT targets must fit one pass of the target hash");
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const uint32_t lt = lanemask_lt();
    const int RW = p.max_px;
    const int nwords = (RW + 1) * G::RWD;
    constexpr int QCAP = kGenQCap;
    constexpr int GQCAP = G::GQ;

    // global scratch: candidates | frontier overflow [2][GQCAP] | seed lists [2][GQCAP] | (WIDE) both bitmaps
    uint32_t* const cand = gscr;
    uint32_t* const gq0 = gscr + kCandCap;
    uint32_t* const gs0 = gq0 + 2 * (size_t)GQCAP;       // general only: [2][GQCAP]
    // shared memory: (not WIDE) both bitmaps | frontier [2][QCAP] | target hash | seed calendar [2][kCal]
    uint32_t* const OF = WIDE ? gs0 + 2 * (size_t)GQCAP : smem;
    uint32_t* const OB = OF + (GEN ? nwords : 0);
    uint32_t* const q0 = WIDE ? smem : OB + nwords;
    uint32_t* const hash = q0 + 2 * QCAP;
    int* const cal0 = (int*)(hash + kHashSize);
    Chain& ch = ctl.ch;

    // ---- which chain / source -------------------------------------------------------------
    int lo = 0, hi = p.n_pairs - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (p.work_prefix[mid] <= item) lo = mid; else hi = mid - 1;
    }
    const int si = item - p.work_prefix[lo];
    const axb_pair_desc lead = p.pairs[lo];
    __syncthreads();                   // the previous item is finished everywhere
    if (tid == 0) {
        int q = lo, nm = 0;
        ch.cum[0] = 0;
        while (q >= 0 && nm < kMaxChain) {
            const axb_pair_desc pd = p.pairs[q];
            ch.orow[nm] = pd.out_off + (long long)si * pd.n_dst;
            ch.item_orig[nm] = p.orig_base + (long long)p.all_prefix[q] + si;
            ch.dst_off[nm] = pd.dst_off;
            ch.cum[nm + 1] = ch.cum[nm] + max(pd.n_dst, 0);
            ++nm;
            q = p.next_same[q];
        }
        ch.nm = nm;
    }
    __syncthreads();
    const int nm = ch.nm;
    const int total_t = ch.cum[nm];
    if (total_t <= 0) return false;

    Src s;
    s.sy = p.dets_yx[2 * ((size_t)lead.src_off + si)];
    s.sx = p.dets_yx[2 * ((size_t)lead.src_off + si) + 1];
    s.H = p.H; s.W = p.W; s.Wp = p.Wp; s.RW = RW;
    s.mbits = p.mask_bits + (size_t)lead.mask_idx * p.H * p.Wp;
    s.iw0 = s.sx >> 5;                 // only used when the source is in the grid
    s.o = s.sx & 31;
    const bool src_in = s.sy >= 0 && s.sy < p.H && s.sx >= 0 && s.sx < p.W;
    const long long r2 = (long long)RW * RW;
    const int o = s.o;
    uint8_t* const par = p.trace_par ? p.trace_par + (size_t)blockIdx.x * G::PAR : nullptr;

    int tt = 0;                        // next target of the flattened chain target list
    for (;;) {                         // passes of <= kHashMaxLoad candidates (one on real data)
        for (int i = tid; i < kHashSize; i += NT) hash[i] = kEmpty;
        if (tid == 0) ctl.ncand = 0;
        for (;;) {                     // groups of NT targets
            __syncthreads();
            if (tt >= total_t || ctl.ncand > kHashMaxLoad - NT) break;
            const int t = tt + tid;
            bool is_cand = false, self = false;
            const bool have = t < total_t;
            int m = 0, j = 0, dy = 0, dx = 0;
            if (have) {
                while (m + 1 < nm && t >= ch.cum[m + 1]) ++m;
                j = t - ch.cum[m];
                const int ty = p.dets_yx[2 * ((size_t)ch.dst_off[m] + j)];
                const int tx = p.dets_yx[2 * ((size_t)ch.dst_off[m] + j) + 1];
                dy = ty - s.sy; dx = tx - s.sx;
                const long long d2 = (long long)dy * dy + (long long)dx * dx;
                if (d2 < r2) {                                        // AxonDetections.py:620-623
                    const bool t_in = ty >= 0 && ty < p.H && tx >= 0 && tx < p.W;
                    const unsigned long long ek = ((unsigned long long)ch.item_orig[m] << 26) | ((unsigned long long)j << 2);
                    if (!src_in) atomicMin(p.err_key, ek | 1ull);     // astar_wrapper.py:39-42
                    else if (!t_in) atomicMin(p.err_key, ek | 2ull);  // astar_wrapper.py:43-45
                    else if (dy == 0 && dx == 0) self = true;         // start == goal: L = 1 (astar.cpp:67,99)
                    else is_cand = abs(dy) + abs(dx) < RW;            // farther cells need > max_px pixels
                }
                const long long oo = ch.orow[m] + j;
                p.path_len[oo] = self ? 1 : RW;                       // None -> max_px (AxonDetections.py:736)
                p.unreachable[oo] = self ? 0 : 1;
                if (p.wcost) p.wcost[oo] = self ? 0.f : __int_as_float(0x7f800000);
            }
            const uint32_t bal = __ballot_sync(kFull, is_cand);
            if (bal) {
                int base = 0;
                if (lane == 0) base = atomicAdd(&ctl.ncand, __popc(bal));
                base = __shfl_sync(kFull, base, 0);
                if (is_cand) {
                    const uint32_t slot = (uint32_t)base + __popc(bal & lt);
                    cand[slot] = ((uint32_t)m << 30) | (uint32_t)j;
                    const uint32_t entry = ((uint32_t)fold_index<WIDE>(dy, dx + o, RW) << G::ESH) | slot;
                    uint32_t h = hash_slot(entry >> G::ESH);
                    while (atomicCAS(&hash[h], kEmpty, entry) != kEmpty) h = (h + 1) & (kHashSize - 1);
                }
            }
            tt += NT;
        }
        const int ncand = ctl.ncand;
        if (ncand > 0) {
            stage_window<GEN, NT, WIDE>(s, OF, OB);
            __syncthreads();
            if (tid == 0) {            // close the source cell, seed the search with it
                const int c = fold_index<WIDE>(0, o, RW);
                OF[phys_word<WIDE>(c)] &= ~(1u << (c & 31));
                if (GEN) OB[phys_word<WIDE>(c)] &= ~(1u << (c & 31));
                q0[0] = pack_cell<WIDE>(0, o);
                if (GEN) gs0[0] = pack_cell<WIDE>(0, o);
                ctl.cnt[0] = GEN ? 0 : 1; ctl.cnt[1] = 0; ctl.cnt[2] = 0;
                ctl.settled[0] = 0; ctl.settled[1] = 0; ctl.settled[2] = 0;
                ctl.unsettled = ncand;
                ctl.seedadd[0] = 0; ctl.seedadd[1] = 0; ctl.seedadd[2] = 0;
            }
            if (GEN) for (int jj = tid; jj < kCal; jj += NT) cal0[jj] = 1;    // the source: one seed tagged j = 0
            __threadfence_block();
            __syncthreads();

            int unsettled = ncand;     // CTA-uniform: updated from ctl.settled[] after each barrier
            // One lane per (frontier cell, direction): the direction of a thread never changes.
            const int dir = tid & 3;
            const int ddy = dir == 0 ? -1 : dir == 3 ? 1 : 0;
            const int dcx = dir == 1 ? -1 : dir == 2 ? 1 : 0;
            // Expanded cells sit at hop <= max_px-1, so every neighbour lies inside the diamond: no
            // bound check; pixels outside the image were staged as closed.
            {
                // ---- all levels -------------------------------------------------------------
                int n_seeds = 1;
                for (int k = 0; k <= RW && n_seeds > 0 && unsettled > 0; ++k) {
                    const uint32_t* const gs_cur = gs0 + (size_t)(k & 1) * GQCAP;
                    uint32_t* const gs_nxt = gs0 + (size_t)((k + 1) & 1) * GQCAP;
                    const int* const cal_cur = cal0 + (k & 1) * kCal;          // #seeds tagged <= j
                    int* const cal_nxt = cal0 + ((k + 1) & 1) * kCal;
                    int j = 0;
                    while (cal_cur[j] == 0) ++j;                      // first seeded j of this level
                    const int jstart = j;
                    int sp = 0, nlist = 0, step = 0, seed_total = 0;
                    bool hit = false;
                    for (;;) {
                        const int send = cal_cur[j];
                        const int tot_n = nlist + (send - sp);
                        if (tot_n == 0) {
                            if (sp >= n_seeds) break;                 // level exhausted
                            if (tid == 0) cal_nxt[j] = seed_total;
                            ++j;
                            continue;
                        }
                        if (k + j >= RW) { hit = true; break; }       // cells at hop == max_px exist
                        if (unsettled <= 0) break;
                        const uint32_t* const qcur = q0 + (step & 1) * QCAP;
                        uint32_t* const qnxt = q0 + ((step + 1) & 1) * QCAP;
                        const uint32_t* const gcur = gq0 + (size_t)(step & 1) * GQCAP;
                        uint32_t* const gnxt = gq0 + (size_t)((step + 1) & 1) * GQCAP;
                        int* const cnt_nxt = &ctl.cnt[(step + 1) % 3];
                        int* const set_nxt = &ctl.settled[(step + 1) % 3];
                        int* const seed_nxt = &ctl.seedadd[(step + 1) % 3];
                        const int ntask = tot_n * 4;
                        for (int base = (tid & ~31); base < ntask; base += NT) {
                            const int i = (base + lane) >> 2;
                            bool got_fg = false, got_bg = false;
                            int idx = 0, my_set = 0;
                            uint32_t ne = 0;
                            if (i < tot_n) {
                                uint32_t e;
                                bool seed = false;
                                if (i < nlist) e = i < QCAP ? qcur[i] : gcur[i - QCAP];
                                else { e = gs_cur[sp + (i - nlist)]; seed = true; }
                                const int dy = (int)(e >> G::CSH) - G::HALF, cx = (int)(e & ((1u << G::CSH) - 1u)) - G::HALF;
                                // off-mask cells settle when their level starts consuming them: by then no
                                // lower level has reached hop max_px
                                if (seed && k > 0 && dir == 0)
                                    my_set += probe_settle<WIDE>(p, hash, cand, ch, (uint32_t)fold_index<WIDE>(dy, cx, RW), k, j);
                                idx = fold_index<WIDE>(dy + ddy, cx + dcx, RW);
                                ne = pack_cell<WIDE>(dy + ddy, cx + dcx);
                                const int pw = phys_word<WIDE>(idx);
                                const uint32_t bit = 1u << (idx & 31);
                                // (WIDE: the bitmaps are global memory changed by atomics: read around L1)
                                const uint32_t wf = WIDE ? __ldcg(&OF[pw]) : OF[pw];
                                if (wf & bit) got_fg = (atomicAnd(&OF[pw], ~bit) & bit) != 0;
                                else if ((WIDE ? __ldcg(&OB[pw]) : OB[pw]) & bit) got_bg = (atomicAnd(&OB[pw], ~bit) & bit) != 0;
                                if (par && (got_fg || got_bg)) par[idx] = (uint8_t)dir;   // entered from (dy, cx)
                            }
                            const uint32_t bf = __ballot_sync(kFull, got_fg);
                            const uint32_t bb = __ballot_sync(kFull, got_bg);
                            if (bf) {
                                int pos = 0;
                                if (lane == 0) pos = atomicAdd(cnt_nxt, __popc(bf));
                                pos = __shfl_sync(kFull, pos, 0);
                                if (got_fg) {
                                    const int slot = pos + __popc(bf & lt);
                                    if (slot < QCAP) qnxt[slot] = ne; else gnxt[slot - QCAP] = ne;
                                    if (k + j + 1 < RW) my_set += probe_settle<WIDE>(p, hash, cand, ch, (uint32_t)idx, k, j + 1);
                                }
                            }
                            if (bb) {
                                int pos = 0;
                                if (lane == 0) pos = seed_total + atomicAdd(seed_nxt, __popc(bb));
                                pos = __shfl_sync(kFull, pos, 0);
                                if (got_bg) gs_nxt[pos + __popc(bb & lt)] = ne;
                            }
                            if (my_set) atomicAdd(set_nxt, my_set);
                        }
                        __syncthreads();
                        nlist = *cnt_nxt;
                        unsettled -= *set_nxt;
                        seed_total += *seed_nxt;           // CTA-uniform running count of the next level's seeds
                        if (tid == 0) { ctl.cnt[step % 3] = 0; ctl.settled[step % 3] = 0; ctl.seedadd[step % 3] = 0; cal_nxt[j] = seed_total; }
                        // one barrier per wave: every counter a wave appends to rotates over three slots
                        sp = send;
                        ++step;
                        ++j;
                    }
                    // ---- level end: complete the next level's calendar ------------------------------
                    __syncthreads();
                    const int total_next = seed_total;
                    for (int jj = tid; jj < kCal; jj += NT) {
                        if (jj >= j) cal_nxt[jj] = total_next;
                        else if (jj < jstart) cal_nxt[jj] = 0;
                    }
                    __syncthreads();
                    if (tid == 0) { ctl.seedadd[0] = 0; ctl.seedadd[1] = 0; ctl.seedadd[2] = 0; ctl.cnt[0] = 0; ctl.cnt[1] = 0; ctl.cnt[2] = 0; ctl.settled[0] = 0; ctl.settled[1] = 0; ctl.settled[2] = 0; }
                    __syncthreads();
                    if (hit) break;                    // k* = k: every deeper target stays None
                    n_seeds = total_next;
                }
            }
        }
        if (tt >= total_t) break;
    }
    return false;
}

}  // namespace

template <bool WIDE>
__device__ __forceinline__ void general_body(const K2Params& p, uint32_t* smem, Ctl& ctl) {
    uint32_t* const gscr = p.gen_scratch + (size_t)blockIdx.x * Geo<WIDE>::SCRATCH;
    const int total = p.counters[p.gen_len_ctr];
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) ctl.item = atomicAdd(&p.counters[p.gen_ticket_ctr], 1);
        __syncthreads();
        const int t = ctl.item;
        if (t >= total) break;
        const int item = p.gen_src[t];
        run_item<true, WIDE ? kK2Threads : kGenThreads, WIDE>(p, item, smem, gscr, ctl);
        if (p.trace_par) {             // one (source, target) request per pair: walk the entry directions back
            __syncthreads();
            if (threadIdx.x == 0) {
                const axb_pair_desc pd = p.pairs[item];
                const int sy = p.dets_yx[2 * (size_t)pd.src_off], sx = p.dets_yx[2 * (size_t)pd.src_off + 1];
                const int ty = p.dets_yx[2 * (size_t)pd.dst_off], tx = p.dets_yx[2 * (size_t)pd.dst_off + 1];
                const uint8_t* const par = p.trace_par + (size_t)blockIdx.x * Geo<WIDE>::PAR;
                int32_t* const out = p.trace_yx + (size_t)item * p.max_px * 2;
                int L = p.unreachable[pd.out_off] ? 0 : p.path_len[pd.out_off];
                const int x0 = sx & ~31, o = sx & 31;
                int dy = ty - sy, cx = tx - x0;
                for (int n = L - 1; n >= 1; --n) {
                    out[2 * n] = sy + dy; out[2 * n + 1] = x0 + cx;
                    const int d = par[fold_index<WIDE>(dy, cx, p.max_px)] & 3;
                    dy -= d == 0 ? -1 : d == 3 ? 1 : 0;
                    cx -= d == 1 ? -1 : d == 2 ? 1 : 0;
                }
                if (L >= 1) {
                    out[0] = sy + dy; out[1] = x0 + cx;
                    if (dy != 0 || cx != o) L = -1;      // cannot happen: the walk must end at the source
                }
                p.trace_len[item] = L;
            }
        }
    }
}

__global__ void __launch_bounds__(kGenThreads, 1) k2_general_kernel(const K2Params p) {
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ Ctl ctl;
    general_body<false>(p, smem, ctl);
}

// 512 <= max_px <= 1023: every source goes through this kernel (bitmaps in global scratch)
__global__ void __launch_bounds__(kK2Threads, kWideCtasPerSm) k2_general_wide_kernel(const K2Params p) {
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ Ctl ctl;
    general_body<true>(p, smem, ctl);
}

// gen_list = every work item (max_px beyond the tile-word kernels' window)
__global__ void __launch_bounds__(256) k2_all_to_general_kernel(const int* __restrict__ work_prefix, int n_pairs,
                                                                int* __restrict__ gen_list, int* __restrict__ counters) {
    const int total = work_prefix[n_pairs];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) gen_list[i] = i;
    if (blockIdx.x == 0 && threadIdx.x == 0) counters[1] = total;
}

// Chains + cover + prefix sums + counter reset, one CTA.
//   same key = same (mask_idx, src_off, n_src): the searches of the two pairs are identical, only
//   the targets differ.  prev/next link same-key pairs inside a +-kChainWindow index window; a
//   pair whose depth in its list is a multiple of kMaxChain leads a chain of <= kMaxChain pairs.
//   The first pair of a list is the HEAD of its detection set.
//
// Cover (lim.allow_reverse).  A frame pair (t, t_bef) asks for the path lengths from the detections of t_bef to
// those of t (AxonDetections.py:553-566).  On the mask the length does not depend on the direction (unit weights,
// astar.cpp:122-136), so a search from a detection of t settles the pair as well as one from t_bef (what it cannot
// settle falls back to a directed search, axb_components.cu).  The frames are the vertices and the pairs the edges
// of a graph, and it is enough to search from a vertex cover: with pairs at distance 1 and 2 two frames of three.
// Greedy, in the order of the heads: a set that no earlier decision forces to be searched and whose pairs all
// lead to searchable sets with room in their chains is NOT searched; its pairs become REVERSE members of the
// chains of their target sets (rev_head / rev_next), which are thereby forced to be searched.
constexpr int kChainWindow = 16;
constexpr int kPrepSmemPairs = 11000;           // cover decided in shared memory up to this many pairs (5 ints each: 215 KB; beyond: in global memory, ~3 us per pair)
struct PrepLimits {
    int n_masks, n_dets;            // mask_idx in [0, n_masks), detection rows in [0, n_dets]
    long long out_capacity;         // out_off + n_src * n_dst <= out_capacity
    int total_src;                  // capacity of the deferred-work lists (sum of n_src the caller sized them for)
    int allow_reverse;
};
__device__ __forceinline__ bool pair_ok(const axb_pair_desc& a, const PrepLimits& lim) {
    return a.n_src >= 0 && a.n_dst >= 0 && a.src_off >= 0 && a.dst_off >= 0 &&
           (long long)a.src_off + a.n_src <= lim.n_dets && (long long)a.dst_off + a.n_dst <= lim.n_dets &&
           a.mask_idx >= 0 && a.mask_idx < lim.n_masks && a.out_off >= 0 &&
           a.out_off + (long long)a.n_src * a.n_dst <= lim.out_capacity;
}
// A descriptor that points outside the caller's buffers gets no work (and status flag 32): the search
// kernels never see it.  Same when the pairs hold more sources than the work lists were sized for.
__global__ void __launch_bounds__(1024) k2_prep_kernel(const axb_pair_desc* __restrict__ pairs, int n_pairs, K2Prep o,
                                                         int* __restrict__ counters, const PrepLimits lim,
                                                         unsigned long long* __restrict__ err_flags) {
    extern __shared__ int s_dyn[];
    __shared__ int s_carry[3];
    __shared__ int s_warp[3][32];
    __shared__ int s_bad;
    const int tid = threadIdx.x;
    int* const work_prefix = o.work_prefix;
    int* const all_prefix = o.all_prefix;
    int* const next_same = o.next_same;
    int* const prev = work_prefix;            // scratch until the scan overwrites it
    int* const head_of = o.p1_prefix;         // scratch until the scan overwrites it: head pair of the pair's source set
    if (tid < 16) counters[tid] = 0;
    if (tid == 0) s_bad = 0;
    for (int i = tid; i < (lim.total_src + 31) / 32; i += blockDim.x) o.item_bits[i] = 0u;
    __syncthreads();
    for (int i = tid; i < n_pairs; i += blockDim.x) {
        const axb_pair_desc a = pairs[i];
        int pv = -1, nx = -1;
        const bool ok = pair_ok(a, lim);
        if (!ok) s_bad = 1;
        if (ok && a.n_src > 0) {
            for (int q = i - 1; q >= max(0, i - kChainWindow); --q) {
                const axb_pair_desc b = pairs[q];
                if (b.mask_idx == a.mask_idx && b.src_off == a.src_off && b.n_src == a.n_src && pair_ok(b, lim)) { pv = q; break; }
            }
            for (int q = i + 1; q <= min(n_pairs - 1, i + kChainWindow); ++q) {
                const axb_pair_desc b = pairs[q];
                if (b.mask_idx == a.mask_idx && b.src_off == a.src_off && b.n_src == a.n_src && pair_ok(b, lim)) { nx = q; break; }
            }
        }
        prev[i] = pv;
        next_same[i] = nx;
        o.rev_head[i] = -1;
        o.rev_next[i] = -1;
        o.state[i] = (ok && a.n_src > 0) ? 0 : -1;          // -1: no work; else bits 0-1 = 1 searched / 2 not searched, bit 2 forced, bits 4.. reverse members
    }
    __syncthreads();
    // depth in the list -> leader flag (kept in all_prefix until the scan), head of the set, leader of the pair's chain
    for (int i = tid; i < n_pairs; i += blockDim.x) {
        int d = 0, h = i;
        for (int q = prev[i]; q >= 0; q = prev[q]) { ++d; h = q; }
        all_prefix[i] = (d % kMaxChain) == 0;
        head_of[i] = o.state[i] < 0 ? -1 : h;
        int l = i;
        for (int k = d % kMaxChain; k > 0; --k) l = prev[l];
        o.lead_of[i] = l;
    }
    __syncthreads();
    // head of the set a pair's TARGETS are (-1: that set is never a source, it cannot be searched; -2: no targets)
    for (int i = tid; i < n_pairs; i += blockDim.x) {
        int dh = -1;
        if (o.state[i] >= 0) {
            const axb_pair_desc a = pairs[i];
            if (a.n_dst == 0) dh = -2;
            else {
                for (int q = max(0, i - kChainWindow); q <= min(n_pairs - 1, i + kChainWindow); ++q) {
                    const axb_pair_desc b = pairs[q];
                    if (o.state[q] >= 0 && b.mask_idx == a.mask_idx && b.src_off == a.dst_off && b.n_src == a.n_dst) { dh = head_of[q]; break; }
                }
            }
        }
        o.dst_head[i] = dh;
    }
    __syncthreads();
    {   // the cover: sequential over the heads, on shared-memory copies when they fit
        const bool in_smem = n_pairs <= kPrepSmemPairs;
        int* const nxt = in_smem ? s_dyn : next_same;
        int* const dsh = in_smem ? s_dyn + n_pairs : o.dst_head;
        int* const st = in_smem ? s_dyn + 2 * n_pairs : o.state;
        int* const rh = in_smem ? s_dyn + 3 * n_pairs : o.rev_head;
        int* const rn = in_smem ? s_dyn + 4 * n_pairs : o.rev_next;
        if (in_smem) {
            for (int i = tid; i < n_pairs; i += blockDim.x) { nxt[i] = next_same[i]; dsh[i] = o.dst_head[i]; st[i] = o.state[i]; rh[i] = -1; rn[i] = -1; }
            __syncthreads();
        }
        if (tid == 0) {
            for (int h = 0; h < n_pairs; ++h) {
                if (st[h] < 0 || head_of[h] != h) continue;
                bool skip = lim.allow_reverse && !(st[h] & 4);
                int len = 0;
                for (int q = h; q >= 0 && skip; q = nxt[q]) ++len;
                for (int q = h; q >= 0 && skip; q = nxt[q]) {
                    const int D = dsh[q];
                    if (D == -2) continue;
                    if (D < 0 || D == h || (st[D] & 3) == 2 || (st[D] >> 4) + len > kMaxRev) skip = false;
                }
                if (!skip) { st[h] = (st[h] & ~3) | 1; continue; }
                st[h] = (st[h] & ~3) | 2;
                for (int q = h; q >= 0; q = nxt[q]) {
                    const int D = dsh[q];
                    if (D == -2) continue;
                    st[D] = (st[D] | 4) + 16;
                    rn[q] = rh[D];
                    rh[D] = q;
                }
            }
        }
        __syncthreads();
        if (in_smem) {
            for (int i = tid; i < n_pairs; i += blockDim.x) { o.state[i] = st[i]; o.rev_head[i] = rh[i]; o.rev_next[i] = rn[i]; }
            __syncthreads();
        }
    }
    // searched flag of every pair (by its source set), kept in dst_head until the scan is done
    for (int i = tid; i < n_pairs; i += blockDim.x) o.dst_head[i] = (o.state[i] >= 0 && (o.state[head_of[i]] & 3) == 1) ? 1 : 0;
    __syncthreads();
    if (tid == 0) { s_carry[0] = 0; s_carry[1] = 0; s_carry[2] = 0; }
    __syncthreads();
    for (int base = 0; base < n_pairs; base += blockDim.x) {
        const int i = base + tid;
        int na = 0;
        if (i < n_pairs) { const axb_pair_desc a = pairs[i]; na = pair_ok(a, lim) ? a.n_src : 0; }
        const int nw = (i < n_pairs && all_prefix[i]) ? na : 0;
        const int np = (i < n_pairs && o.dst_head[i]) ? nw : 0;
        int xa = na, xw = nw, xp = np;
        for (int k = 1; k < 32; k <<= 1) {
            const int ya = __shfl_up_sync(kFull, xa, k), yw = __shfl_up_sync(kFull, xw, k), yp = __shfl_up_sync(kFull, xp, k);
            if ((tid & 31) >= k) { xa += ya; xw += yw; xp += yp; }
        }
        if ((tid & 31) == 31) { s_warp[0][tid >> 5] = xa; s_warp[1][tid >> 5] = xw; s_warp[2][tid >> 5] = xp; }
        __syncthreads();
        if (tid < 32) {
            int wa = tid < (int)(blockDim.x >> 5) ? s_warp[0][tid] : 0;
            int ww = tid < (int)(blockDim.x >> 5) ? s_warp[1][tid] : 0;
            int wp = tid < (int)(blockDim.x >> 5) ? s_warp[2][tid] : 0;
            for (int k = 1; k < 32; k <<= 1) {
                const int ya = __shfl_up_sync(kFull, wa, k), yw = __shfl_up_sync(kFull, ww, k), yp = __shfl_up_sync(kFull, wp, k);
                if (tid >= k) { wa += ya; ww += yw; wp += yp; }
            }
            s_warp[0][tid] = wa; s_warp[1][tid] = ww; s_warp[2][tid] = wp;
        }
        __syncthreads();
        const int offa = (tid >> 5) ? s_warp[0][(tid >> 5) - 1] : 0;
        const int offw = (tid >> 5) ? s_warp[1][(tid >> 5) - 1] : 0;
        const int offp = (tid >> 5) ? s_warp[2][(tid >> 5) - 1] : 0;
        const int incla = s_carry[0] + offa + xa, inclw = s_carry[1] + offw + xw, inclp = s_carry[2] + offp + xp;
        __syncthreads();
        if (i < n_pairs) { all_prefix[i] = incla - na; work_prefix[i] = inclw - nw; o.p1_prefix[i] = inclp - np; }
        if (tid == (int)blockDim.x - 1) { s_carry[0] = incla; s_carry[1] = inclw; s_carry[2] = inclp; }
        __syncthreads();
    }
    if (tid == 0) {
        const bool over = s_carry[1] > lim.total_src;        // the deferred-work lists would overflow: no work at all
        all_prefix[n_pairs] = s_carry[0];
        work_prefix[n_pairs] = over ? 0 : s_carry[1];
        o.p1_prefix[n_pairs] = over ? 0 : s_carry[2];
        if (s_bad || over) atomicOr(err_flags, 32ull);
    }
}

void launch_k2_prep(const axb_pair_desc* pairs, int n_pairs, const K2Prep& o, int* counters, int n_masks, int n_dets,
                    long long out_capacity, int total_src, int allow_reverse, unsigned long long* err_flags, cudaStream_t stream) {
    static unsigned long long attr_set = 0ull;
    if (first_use_on_device(attr_set))
        cudaFuncSetAttribute(k2_prep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 5 * kPrepSmemPairs * (int)sizeof(int));
    PrepLimits lim;
    lim.n_masks = n_masks; lim.n_dets = n_dets; lim.out_capacity = out_capacity; lim.total_src = total_src; lim.allow_reverse = allow_reverse;
    const size_t smem = n_pairs <= kPrepSmemPairs ? 5 * (size_t)n_pairs * sizeof(int) : 0;
    k2_prep_kernel<<<1, 1024, smem, stream>>>(pairs, n_pairs, o, counters, lim, err_flags);
}

// Work items whose detection lies in the grid but off the mask: level 0 is at most their four neighbours, the
// all-levels kernel takes them from the start (list `early`), concurrently with the level-0 kernels, which skip
// the forward members of such items and leave their rows alone.
__global__ void __launch_bounds__(256) k2_classify_kernel(const K2Params p, int* __restrict__ early) {
    const int total = p.work_prefix[p.n_pairs];
    for (int item = blockIdx.x * blockDim.x + threadIdx.x; item < total; item += gridDim.x * blockDim.x) {
        int lo = 0, hi = p.n_pairs - 1;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (p.work_prefix[mid] <= item) lo = mid; else hi = mid - 1;
        }
        const axb_pair_desc lead = p.pairs[lo];
        const size_t d = (size_t)lead.src_off + (item - p.work_prefix[lo]);
        const int y = p.dets_yx[2 * d], x = p.dets_yx[2 * d + 1];
        if (y < 0 || y >= p.H || x < 0 || x >= p.W) continue;
        if ((__ldg(p.mask_bits + ((size_t)lead.mask_idx * p.H + y) * p.Wp + (x >> 5)) >> (x & 31)) & 1u) continue;
        atomicOr(&p.item_bits[item >> 5], 1u << (item & 31));
        early[atomicAdd(&p.counters[kCtrEarlyLen], 1)] = item;
    }
}

// Matrices of the pairs that are answered from their target side: None everywhere, row-wise and coalesced (the
// searches then store only what they settle; their own default stores would run down the columns, which matters
// when the outputs live in another GPU's memory).  pair_searched[q] = 0 and state[q] >= 0: q is such a pair.
__global__ void __launch_bounds__(256) k2_fill_reverse_kernel(const K2Params p, const int* __restrict__ pair_searched,
                                                              const int* __restrict__ state) {
    for (int q = blockIdx.y; q < p.n_pairs; q += gridDim.y) {
        if (pair_searched[q] || state[q] < 0) continue;
        const axb_pair_desc pd = p.pairs[q];
        const long long n = (long long)pd.n_src * pd.n_dst;
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
            p.path_len[pd.out_off + i] = p.max_px;
            p.unreachable[pd.out_off + i] = 1;
            if (p.wcost) p.wcost[pd.out_off + i] = __int_as_float(0x7f800000);
        }
    }
}

void launch_k2_fill_reverse(const K2Params& p, const int* pair_searched, const int* state, cudaStream_t stream) {
    const dim3 grid(8, (unsigned)std::max(1, std::min(p.n_pairs, 148 * 4)));
    k2_fill_reverse_kernel<<<grid, 256, 0, stream>>>(p, pair_searched, state);
}

void launch_k2_classify(const K2Params& p, int* early, int total_src, cudaStream_t stream) {
    const int grid = std::max(1, std::min((total_src + 255) / 256, 148 * 4));
    k2_classify_kernel<<<grid, 256, 0, stream>>>(p, early);
}

void launch_k2_general(const K2Params& p, int grid, cudaStream_t stream) {
    if (p.max_px > kFastMaxPx) {
        k2_general_wide_kernel<<<grid, kK2Threads, k2_gen_wide_smem_bytes(), stream>>>(p);
        return;
    }
    static unsigned long long attr_set = 0ull;
    if (first_use_on_device(attr_set)) {
        cudaFuncSetAttribute(k2_general_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)k2_gen_smem_bytes(kFastMaxPx));
    }
    k2_general_kernel<<<grid, kGenThreads, k2_gen_smem_bytes(p.max_px), stream>>>(p);
}

void launch_k2_all_to_general(const int* work_prefix, int n_pairs, int* gen_list, int* counters, cudaStream_t stream) {
    k2_all_to_general_kernel<<<148, 256, 0, stream>>>(work_prefix, n_pairs, gen_list, counters);
}

}  // namespace axb
