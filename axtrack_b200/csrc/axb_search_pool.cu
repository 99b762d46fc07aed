// K2 pooled kernel (v9) -- kPoolSlots level-0 searches per CTA share ONE wavefront loop (sm_100a).
//
// Same search, window layout and semantics as k2_sparse_kernel (axb_search_fast.cu header; replaces
// pyastar2d/src/cpp/astar.cpp:46-172 driven by axtrack/AxonDetections.py:553-579,600-629).  What
// changes is the mapping to the SM.  A search is ~max_px DEPENDENT wavefronts of ~44 tile-word
// entries, i.e. 1.4 warps of work per wave: with one search per 3-warp CTA (round 1) two warps ran
// ~270 instructions each for 44 entries, the third only paid the ~80-instruction hand-over, and
// the kernel was bound by instruction issue (ncu r01l: 9.5e9 warp instructions per launch, issue
// slots 70 % busy, 26/32 lanes).  One warp per search halves the instructions but leaves 10 warps
// per SM, far too few to hide the ~7 cycles per dependent instruction (measured: same time).
// Here a CTA owns kPoolSlots searches ("slots", each with its own sparse window in shared memory)
// and ONE frontier queue: an entry is (32 frontier bits, slot << 16 | tile word key), every lane
// expands one entry of whichever slot, and a CTA wave advances every running slot by one hop.
// Entries of kPoolSlots searches fill the lanes of kPoolWarps warps (2 x 44 entries -> 3 warps),
// the per-wave hand-over is paid once per CTA, and every warp has work in (nearly) every wave.
//
// Slots are independent state machines: a slot that finishes (all targets reached, frontier dead,
// hop max_px reached, or window too dense) is finalised and refilled from the ticket counter in a
// CTA-wide "service" section between two waves while the other slots keep their frontier entries
// in the queue; its stale entries are scrubbed.  Per-slot frontier sizes are counted where the
// entries are CONSUMED (one ballot per iteration) and reach the control words one wave later.
//
// Targets: a word whose column pair holds a target (flag bits of the presence table) appends
// (key, frontier bits, hop) to the warp's pending-hit list in global scratch -- a store, no
// dependent load -- and the list is resolved against the slot's target hash 32 hits at a time
// (the global-memory probe chain cost 23 % of a wave when it ran inside the wave loop).
#include "axb_search_tiles.cuh"

namespace axb {

namespace {

constexpr int PS = kPoolSlots;
constexpr int PNW = kPoolWarps;
constexpr int PNT = 32 * PNW;
constexpr int PSEG = kPoolQCap / PNW;          // shared-memory frontier entries per warp segment and buffer
constexpr int PGSEG = kFastGQCap / PNW;        // global overflow entries per warp segment and buffer
constexpr int PQBUF = kPoolQCap + 2 * PS;      // entries of one queue buffer: PNW segments + the seed segment (PS live + PS scrubbed seeds)
constexpr int PHITS = 64;                      // pending target hits per warp
constexpr int PHASH = kSparseHash;
constexpr int kPoolPassLoad = kSparseHashMaxLoad > PNT ? kSparseHashMaxLoad - PNT : 0;
static_assert(PS >= 1 && PS <= 3, "a warp's wave record holds one (consumed | settled << 16) word per slot in y, z, w");
static_assert(PSEG + PGSEG < 0x8000, "the per-warp wave counter keeps 15 bits of entries");

enum { kSlotIdle = 0, kSlotRun = 1 };
enum { kFinNone = 0, kFinSettled = 1, kFinDead = 2, kFinHit = 3, kFinDefer = 4 };

struct __align__(16) SlotVar {     // a slot's search state between two runs of the wave loop (registers while it runs)
    int j;             // hop of the cells this slot's entries of the CURRENT wave sit at
    int state;         // kSlotIdle / kSlotRun
    int unsettled;     // candidates not yet reached
    int fresh;         // 1 until the slot's seed entry has been consumed
};

struct Slot {
    Chain ch;
    int item, tt, total_t, ncand, alloc;
    int sy, sx, mask_idx, si, lead;
};

struct __align__(16) PoolCtl {
    int4 wc[2][PNW + 1];       // per warp (+ the seed segment) and wave: x = entries appended | overflow << 15,
                               // y/z/w = slot 0/1/2: its entries this warp CONSUMED | its targets this warp settled << 16
    SlotVar sv[PS];
    Slot slot[PS];
    int fin[PS];               // service section: why the slot finished
    int ticket;                // broadcast of the ticket counter
    int drained;               // no work item left
    int nseed;                 // seed entries in the current buffer
};

struct PoolMem {               // CTA-uniform addresses
    uint32_t* win0;            // slot 0's packed words; slot s at win0 + s * slot_words
    int slot_words;            // words per slot: packed window words + presence table
    int sparse_words;
    uint2* q0;                 // [2][PQBUF]
    uint2* gq0;                // global overflow [2][PNW][PGSEG]
    uint32_t* cand0;           // [PS][kCandCap]
    uint32_t* hash0;           // [PS][PHASH]
    uint4* hits0;              // [PNW][PHITS]
};

__device__ __forceinline__ uint32_t* slot_win(const PoolMem& m, int s) { return m.win0 + (size_t)s * m.slot_words; }
__device__ __forceinline__ uint2* slot_pb(const PoolMem& m, int s) { return (uint2*)(slot_win(m, s) + m.sparse_words); }

// Resolve the warp's pending hits: (key | slot << 16, frontier bits, hop) against the slot's target hash.
// -> targets settled per slot, 16 bits each (warp-uniform).
__device__ __noinline__ unsigned long long resolve_hits(const K2Params& p, const PoolMem& m, PoolCtl& ctl, const int hcnt) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint4* const hits = m.hits0 + (size_t)warp * PHITS;
    unsigned long long mine = 0ull;
    for (int i = lane; i < hcnt; i += 32) {
        const uint4 h = hits[i];
        const int hs = (int)(h.x >> 16);
        const int ns = probe_settle<PHASH>(p, m.hash0 + (size_t)hs * PHASH, m.cand0 + (size_t)hs * kCandCap, ctl.slot[hs].ch,
                                           h.x & 0xFFFFu, h.y, (int)h.z);
        mine += (unsigned long long)ns << (16 * hs);
    }
    // at most kSparseHashMaxLoad (< 2^16) targets per slot: the 16-bit fields cannot carry into each other
    const uint32_t lo = __reduce_add_sync(kFull, (uint32_t)mine), hi = __reduce_add_sync(kFull, (uint32_t)(mine >> 32));
    return ((unsigned long long)hi << 32) | lo;
}

// Where frontier entry i of a wave lives: segment sg (warp that appended it, or PNW = seeds), offset off.
struct EntryRef { int sg, off; };
__device__ __forceinline__ EntryRef entry_ref(int i, const int* pre) {
    int sg = 0;
    #pragma unroll
    for (int w = 1; w <= PNW; ++w) sg += i >= pre[w];
    int base = 0;
    #pragma unroll
    for (int w = 1; w <= PNW; ++w) base = sg == w ? pre[w] : base;
    EntryRef r;
    r.sg = sg; r.off = i - base;
    return r;
}

// Set up slot s for the next pass of its item (a new item when it.tt == 0): target hash, candidates,
// staged window, target flags, seed entry.  CTA-wide; returns true when the slot now runs a search.
// When it returns false the item either needs nothing more (out.done) or goes to another kernel.
__device__ bool setup_slot(const K2Params& p, const PoolMem& m, PoolCtl& ctl, const int s, const int cur) {
    const int tid = threadIdx.x, lane = tid & 31;
    const uint32_t lt = lanemask_lt();
    const int RW = p.max_px;
    const int NR = (RW + 3) / 4 + 1;
    Slot& sl = ctl.slot[s];
    Chain& ch = sl.ch;
    uint32_t* const WIN = slot_win(m, s);
    uint2* const PB = slot_pb(m, s);
    uint32_t* const hash = m.hash0 + (size_t)s * PHASH;
    uint32_t* const cand = m.cand0 + (size_t)s * kCandCap;

    if (sl.tt == 0) {                  // new item: which chain / source
        const int item = sl.item;
        int lo = 0, hi = p.n_pairs - 1;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (p.work_prefix[mid] <= item) lo = mid; else hi = mid - 1;
        }
        if (tid == 0) {
            const int si = item - p.work_prefix[lo];
            const axb_pair_desc lead = p.pairs[lo];
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
            sl.total_t = ch.cum[nm];
            sl.sy = p.dets_yx[2 * ((size_t)lead.src_off + si)];
            sl.sx = p.dets_yx[2 * ((size_t)lead.src_off + si) + 1];
            sl.mask_idx = lead.mask_idx;
        }
        __syncthreads();
    }
    const int nm = ch.nm;
    const int total_t = sl.total_t;
    const int sy = sl.sy, sx = sl.sx;
    Win w;
    w.tiles = p.mask_tiles + (size_t)sl.mask_idx * p.TH * p.TWp;
    w.TH = p.TH; w.TWp = p.TWp; w.RW = RW; w.NR = NR;
    w.oy = sy & 3; w.ox = sx & 31;
    w.ty0 = sy >> 2; w.tx0 = (sx >> 5) * 4;
    const bool src_in = sy >= 0 && sy < p.H && sx >= 0 && sx < p.W;
    const long long r2 = (long long)RW * RW;

    for (;;) {                         // passes without a candidate are finished on the spot
        __syncthreads();               // thread 0's sl.tt of the previous pass
        int tt = sl.tt;
        if (tt >= total_t) return false;                   // nothing (more) to do for this item
        for (int i = tid; i < PHASH; i += PNT) hash[i] = kEmpty;
        if (tid == 0) { sl.ncand = 0; sl.alloc = 0; }
        for (;;) {                     // groups of PNT targets
            __syncthreads();
            if (tt >= total_t || sl.ncand > kPoolPassLoad) break;
            const int t = tt + tid;
            bool is_cand = false, self = false;
            int mm = 0, j = 0, dy = 0, dx = 0;
            if (t < total_t) {
                while (mm + 1 < nm && t >= ch.cum[mm + 1]) ++mm;
                j = t - ch.cum[mm];
                const int ty = p.dets_yx[2 * ((size_t)ch.dst_off[mm] + j)];
                const int tx = p.dets_yx[2 * ((size_t)ch.dst_off[mm] + j) + 1];
                dy = ty - sy; dx = tx - sx;
                const long long d2 = (long long)dy * dy + (long long)dx * dx;
                if (d2 < r2) {                                        // AxonDetections.py:620-623
                    const bool t_in = ty >= 0 && ty < p.H && tx >= 0 && tx < p.W;
                    const unsigned long long ek = ((unsigned long long)ch.item_orig[mm] << 26) | ((unsigned long long)j << 2);
                    if (!src_in) atomicMin(p.err_key, ek | 1ull);     // astar_wrapper.py:39-42
                    else if (!t_in) atomicMin(p.err_key, ek | 2ull);  // astar_wrapper.py:43-45
                    else if (dy == 0 && dx == 0) self = true;         // start == goal: L = 1 (astar.cpp:67,99)
                    else is_cand = abs(dy) + abs(dx) < RW;            // farther cells need > max_px pixels
                }
                const long long oo = ch.orow[mm] + j;
                p.path_len[oo] = self ? 1 : RW;                       // None -> max_px (AxonDetections.py:736)
                p.unreachable[oo] = self ? 0 : 1;
                if (p.wcost) p.wcost[oo] = self ? 0.f : __int_as_float(0x7f800000);
            }
            const uint32_t bal = __ballot_sync(kFull, is_cand);
            if (bal) {
                int base = 0;
                if (lane == 0) base = atomicAdd(&sl.ncand, __popc(bal));
                base = __shfl_sync(kFull, base, 0);
                if (is_cand) {
                    const uint32_t cs = (uint32_t)base + __popc(bal & lt);
                    AXB_ASSERT(cs < (uint32_t)kCandCap && cs < 2048u);
                    cand[cs] = ((uint32_t)mm << 30) | (uint32_t)j;
                    int bit;
                    const uint32_t key = cell_key(dy, dx, w.oy, w.ox, NR, bit);
                    const uint32_t entry = (key << 16) | ((uint32_t)bit << 11) | cs;
                    uint32_t h = hash_slot<PHASH>(key);
                    while (atomicCAS(&hash[h], kEmpty, entry) != kEmpty) h = (h + 1) & (PHASH - 1);
                }
            }
            tt += PNT;
        }
        if (tid == 0) sl.tt = tt;
        const int ncand = sl.ncand;
        if (ncand == 0) continue;      // no in-range target in this pass: next pass (or done)

        stage_window<true>(w, WIN, PB, &sl.alloc, m.sparse_words, tid >> 5, PNW);
        __syncthreads();
        if (sl.alloc > m.sparse_words) {                   // more non-zero tile words than the slot holds: dense kernel
            __syncthreads();
            if (tid == 0) { p.dense_list[atomicAdd(&p.counters[3], 1)] = sl.item; sl.tt = total_t; }
            __syncthreads();
            return false;
        }
        for (int i = tid; i < PHASH; i += PNT) {           // flag the tile-word pairs that hold a target: bits 16.. of PB[].y
            const uint32_t e = hash[i];
            if (e != kEmpty) {
                const uint32_t key = e >> 16;
                atomicOr(&PB[(key & 0x7FFFu) >> 5].y, 1u << (16u + ((key & 31u) >> 1)));
            }
        }
        if (tid == 0) {                // close the source cell, seed the search with it
            int bit;
            const uint32_t key = cell_key(0, 0, w.oy, w.ox, NR, bit);
            const int widx = (int)(key & 0x7FFFu);
            const uint2 pb = PB[widx >> 5];
            if ((pb.x >> (widx & 31)) & 1u) {
                const int wi = (int)(pb.y & 0xFFFFu) + __popc(pb.x & ((1u << (widx & 31)) - 1u));
                AXB_ASSERT(wi < m.sparse_words);
                WIN[wi] &= ~(1u << bit);
            }
            (m.q0 + (size_t)cur * PQBUF)[kPoolQCap + ctl.nseed] = make_uint2(1u << bit, ((uint32_t)s << 16) | key);
            ctl.nseed += 1;
            SlotVar v;
            v.j = 0; v.state = kSlotRun; v.unsettled = ncand; v.fresh = 1;
            ctl.sv[s] = v;
        }
        __syncthreads();
        return true;
    }
}

}  // namespace

__global__ void __launch_bounds__(PNT, kPoolCtasPerSm) k2_pool_kernel(const K2Params p) {
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ PoolCtl ctl;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lt = lanemask_lt();
    const int RW = p.max_px;
    const int NR = (RW + 3) / 4 + 1;
    const int total_items = p.work_prefix[p.n_pairs];

    PoolMem m;
    m.sparse_words = p.sparse_words;
    m.slot_words = p.sparse_words + NR * 8;
    m.win0 = smem;
    m.q0 = (uint2*)(smem + (size_t)PS * m.slot_words);
    uint32_t* const gscr = p.pool_scratch + (size_t)blockIdx.x * kPoolScratchWords;
    m.gq0 = (uint2*)gscr;
    m.cand0 = gscr + 4 * (size_t)kFastGQCap;
    m.hash0 = m.cand0 + (size_t)PS * kCandCap;
    m.hits0 = (uint4*)(m.hash0 + (size_t)PS * PHASH);

    if (tid == 0) {
        ctl.drained = 0; ctl.nseed = 0;
        for (int s = 0; s < PS; ++s) {
            SlotVar v; v.j = 0; v.state = kSlotIdle; v.unsettled = 0; v.fresh = 0;
            ctl.sv[s] = v; ctl.fin[s] = kFinNone; ctl.slot[s].tt = 0; ctl.slot[s].total_t = 0; ctl.slot[s].item = -1;
        }
        for (int b = 0; b < 2; ++b)
            for (int w = 0; w <= PNW; ++w) ctl.wc[b][w] = make_int4(0, 0, 0, 0);
    }
    __syncthreads();

    const uint32_t win_s0 = smem_u32(m.win0);
    const uint32_t slot_bytes = 4u * (uint32_t)m.slot_words;
    const uint32_t pb_off = 4u * (uint32_t)m.sparse_words;
    const uint32_t q_s = smem_u32(m.q0);
    const uint32_t lastU = (uint32_t)(NR - 1);
    int cur = 0;                       // queue buffer the next wave reads
    int hcnt = 0;                      // this warp's pending hits (warp-uniform)
    uint4* const myhits = m.hits0 + (size_t)warp * PHITS;

    for (;;) {
        // ================= service: finalise finished slots, refill idle ones =================
        if (hcnt) {
            const unsigned long long set = resolve_hits(p, m, ctl, hcnt);
            hcnt = 0;
            #pragma unroll
            for (int s = 0; s < PS; ++s) {
                const int n = (int)((set >> (16 * s)) & 0xFFFFull);
                if (lane == 0 && n) atomicSub(&ctl.sv[s].unsettled, n);
            }
        }
        __syncthreads();
        if (tid == 0) {
            for (int s = 0; s < PS; ++s) {
                const int f = ctl.fin[s];
                if (f == kFinNone) continue;
                // level 0 cannot answer: frontier dead (or a frontier list overflowed) with targets still open
                if (f == kFinDefer || (f == kFinDead && ctl.sv[s].unsettled > 0)) {
                    p.gen_list[atomicAdd(&p.counters[1], 1)] = ctl.slot[s].item;
                    ctl.slot[s].tt = ctl.slot[s].total_t;           // the all-levels kernel redoes the whole item
                }
                ctl.fin[s] = kFinNone;
                ctl.sv[s].state = kSlotIdle;
            }
            ctl.nseed = ctl.wc[cur][PNW].x;                         // seeds of an earlier service that no wave has consumed yet
            for (int w = 0; w < PNW; ++w) ctl.wc[cur][w].x &= 0x7FFF;   // an overflow has been answered (every slot deferred)
        }
        __syncthreads();
        for (int s = 0; s < PS; ++s) {
            for (;;) {                 // until slot s runs a search or no work is left
                if (ctl.sv[s].state == kSlotRun) break;
                if (ctl.slot[s].tt >= ctl.slot[s].total_t) {       // no further pass of the slot's item: next item
                    if (ctl.drained) break;
                    __syncthreads();
                    if (tid == 0) {
                        const int item = atomicAdd(&p.counters[0], 1);
                        if (item >= total_items) ctl.drained = 1;
                        else { ctl.slot[s].item = item; ctl.slot[s].tt = 0; ctl.slot[s].total_t = 1; }
                    }
                    __syncthreads();
                    if (ctl.drained) break;
                }
                if (setup_slot(p, m, ctl, s, cur)) break;
                __syncthreads();
            }
            __syncthreads();
        }
        if (tid == 0) ctl.wc[cur][PNW].x = ctl.nseed;
        __syncthreads();

        // the slots' state lives in registers while the waves run (identical in every thread)
        int jv[PS], uns[PS];
        uint32_t running = 0u, fresh = 0u;
        #pragma unroll
        for (int s = 0; s < PS; ++s) {
            const SlotVar v = ctl.sv[s];
            jv[s] = v.j; uns[s] = v.unsettled;
            running |= v.state == kSlotRun ? 1u << s : 0u;
            fresh |= v.fresh ? 1u << s : 0u;
        }
        if (!running) break;           // drained and every slot idle
        bool reentry = true;           // the control words of buffer `cur` have been applied already

        // ================= waves =================
        uint32_t finmask = 0u;
        for (;;) {
            // ---- control: segment sizes, per-slot frontier sizes and settled targets of the previous wave ----
            int pre[PNW + 2];
            int cons[PS], setl[PS];
            bool ovf = false;
            pre[0] = 0;
            #pragma unroll
            for (int s = 0; s < PS; ++s) { cons[s] = 0; setl[s] = 0; }
            #pragma unroll
            for (int w = 0; w <= PNW; ++w) {
                const int4 c = ctl.wc[cur][w];
                pre[w + 1] = pre[w] + (c.x & 0x7FFF);
                ovf |= (c.x & 0x8000) != 0;
                #pragma unroll
                for (int s = 0; s < PS; ++s) {
                    const int wd = s == 0 ? c.y : s == 1 ? c.z : c.w;
                    cons[s] += wd & 0xFFFF;
                    setl[s] += (wd >> 16) & 0xFFFF;
                }
            }
            const int total = pre[PNW + 1];
            uint32_t runmask = 0u;     // slots whose entries are expanded in this wave
            int fins[PS];
            #pragma unroll
            for (int s = 0; s < PS; ++s) {
                int f = kFinNone;
                if ((running >> s) & 1u) {
                    if (!reentry) uns[s] -= setl[s];
                    // cons[s] = this slot's entries of the PREVIOUS wave (the cells at hop jv[s] - 1)
                    if (ovf) f = kFinDefer;
                    else if (uns[s] <= 0) f = kFinSettled;
                    else if (!reentry && !((fresh >> s) & 1u) && cons[s] == 0) f = kFinDead;
                    else if (jv[s] > RW) f = kFinHit;                 // entries at hop max_px existed (counted, not expanded)
                    else if (jv[s] < RW) runmask |= 1u << s;          // at hop max_px the entries are only counted
                }
                fins[s] = f;
                finmask |= f != kFinNone ? 1u << s : 0u;
            }
            if (finmask) {             // -> service (every thread took the same decision from the same words)
                if (tid == 0) {
                    #pragma unroll
                    for (int s = 0; s < PS; ++s) {
                        ctl.fin[s] = fins[s];
                        SlotVar v;
                        v.j = jv[s]; v.state = ((running >> s) & 1u) ? kSlotRun : kSlotIdle; v.unsettled = uns[s]; v.fresh = (fresh >> s) & 1u;
                        ctl.sv[s] = v;
                    }
                }
                break;
            }
            reentry = false;

            const uint32_t qcur_s = q_s + (uint32_t)cur * (PQBUF * 8u);
            const uint32_t qnxt_s = q_s + (uint32_t)(cur ^ 1) * (PQBUF * 8u) + (uint32_t)warp * (PSEG * 8u);
            const uint2* const gcur = m.gq0 + (size_t)cur * kFastGQCap;
            uint2* const gnxt = m.gq0 + (size_t)(cur ^ 1) * kFastGQCap + (size_t)warp * PGSEG;
            int wn = 0;
            int ccnt[PS], wset[PS];
            #pragma unroll
            for (int s = 0; s < PS; ++s) { ccnt[s] = 0; wset[s] = 0; }
            bool wovf = false;
            for (int base = warp * 32; base < total; base += PNT) {
                const int i = base + lane;
                const bool have = i < total;
                const EntryRef er = entry_ref(i, pre);
                const bool seedseg = er.sg == PNW;
                const bool in_sm = have && (seedseg || er.off < PSEG);
                uint2 e = lds64_if(qcur_s + 8u * (uint32_t)(seedseg ? kPoolQCap + er.off : er.sg * PSEG + er.off), in_sm);
                if (__any_sync(kFull, have && !in_sm)) {          // entries that spilled to global memory
                    if (have && !in_sm) e = gcur[(size_t)er.sg * PGSEG + (er.off - PSEG)];
                }
                const uint32_t slot = e.y >> 16;                  // >= PS: scrubbed entry
                const uint32_t k0 = e.y & 0xFFFFu;
                // entries consumed per slot (whether expanded or only counted at hop max_px)
                #pragma unroll
                for (int s = 0; s < PS; ++s) ccnt[s] += __popc(__ballot_sync(kFull, have && slot == (uint32_t)s));
                const bool act = have && slot < (uint32_t)PS && ((runmask >> slot) & 1u);
                const uint32_t f = act ? e.x : 0u;
                AXB_ASSERT(!act || f != 0u);
                int j = 0;
                #pragma unroll
                for (int s = 0; s < PS; ++s) j = slot == (uint32_t)s ? jv[s] : j;
                const uint32_t win_s = win_s0 + (act ? slot : 0u) * slot_bytes;
                const uint32_t pb_s = win_s + pb_off;
                auto resolve = [&](uint32_t key, uint32_t& want, uint32_t& aux) -> uint32_t {
                    const uint32_t widx = key & 0x7FFFu;
                    AXB_ASSERT(!want || (widx >> 7) < (uint32_t)NR);
                    const uint2 pb = lds64_if(pb_s + 8u * (widx >> 5), want);
                    const uint32_t bit = 1u << (widx & 31u);
                    if (!(pb.x & bit)) want = 0u;
                    aux = pb.y;
                    AXB_ASSERT(!want || (pb.y & 0xFFFFu) + (uint32_t)__popc(pb.x & (bit - 1u)) < (uint32_t)m.sparse_words);
                    return win_s + 4u * ((pb.y & 0xFFFFu) + (uint32_t)__popc(pb.x & (bit - 1u)));
                };
                // candidates: cells of the word itself, of the right / left / lower / upper word
                uint32_t cd0 = ((f << 1) & 0xFEFEFEFEu) | ((f >> 1) & 0x7F7F7F7Fu) | (f << 8) | (f >> 8);
                uint32_t cd1 = (f & 0x80808080u) >> 7;
                uint32_t cd2 = (f & 0x01010101u) << 7;
                uint32_t cd3 = f >> 24, cd4 = f << 24;
                const uint32_t rowk = k0 & ~127u;
                const uint32_t k1 = rowk | ((k0 + 1u) & 127u);
                const uint32_t k2 = rowk | ((k0 - 1u) & 127u);
                uint32_t k3 = k0 + 128u, k4 = k0 - 128u;
                const uint32_t hr = k0 >> 7;                      // half << 8 | folded row
                if (hr == lastU) k3 = 0x8000u | ((k0 + 64u) & 127u);              // tile row 0 -> 1
                if (hr == 256u) k4 = (lastU << 7) | ((k0 + 64u) & 127u);         // tile row 1 -> 0
                if (hr == 255u + lastU) cd3 = 0u;                 // window borders (never open anyway)
                if (hr == 0u) cd4 = 0u;
                uint32_t aux0 = 0u, auxn = 0u;
                const uint32_t a0 = resolve(k0, cd0, aux0), a1 = resolve(k1, cd1, auxn), a2 = resolve(k2, cd2, auxn);
                const uint32_t a3 = resolve(k3, cd3, auxn), a4 = resolve(k4, cd4, auxn);
                const bool maybe = f != 0u && ((aux0 >> (16u + ((k0 & 31u) >> 1))) & 1u);
                const uint32_t o0 = lds_if(a0, cd0) & cd0;
                const uint32_t o1 = lds_if(a1, cd1) & cd1;
                const uint32_t o2 = lds_if(a2, cd2) & cd2;
                const uint32_t o3 = lds_if(a3, cd3) & cd3;
                const uint32_t o4 = lds_if(a4, cd4) & cd4;
                const uint32_t c0 = claim_if(a0, o0);
                const uint32_t c1 = claim_if(a1, o1);
                const uint32_t c2 = claim_if(a2, o2);
                const uint32_t c3 = claim_if(a3, o3);
                const uint32_t c4c = claim_if(a4, o4);
                // targets among the cells being expanded (hop j, path length j + 1): remember, resolve in batches
                const uint32_t mb = __ballot_sync(kFull, maybe);
                if (mb) {
                    if (maybe) myhits[hcnt + __popc(mb & lt)] = make_uint4(e.y, f, (uint32_t)j, 0u);
                    hcnt += __popc(mb);
                    if (hcnt > PHITS - 32) {
                        __syncwarp();
                        const unsigned long long set = resolve_hits(p, m, ctl, hcnt);
                        hcnt = 0;
                        #pragma unroll
                        for (int s = 0; s < PS; ++s) wset[s] += (int)((set >> (16 * s)) & 0xFFFFull);
                    }
                }
                // append the claims to this warp's segment of the next frontier, lane-major
                const uint32_t sbits = e.y & 0xFFFF0000u;
                const uint32_t b0 = __ballot_sync(kFull, c0 != 0u), b1 = __ballot_sync(kFull, c1 != 0u);
                const uint32_t b2 = __ballot_sync(kFull, c2 != 0u), b3 = __ballot_sync(kFull, c3 != 0u);
                const uint32_t b4 = __ballot_sync(kFull, c4c != 0u);
                const int tot = __popc(b0) + __popc(b1) + __popc(b2) + __popc(b3) + __popc(b4);
                if (tot) {
                    int slotq = wn + __popc(b0 & lt) + __popc(b1 & lt) + __popc(b2 & lt) + __popc(b3 & lt) + __popc(b4 & lt);
                    if (wn + tot <= PSEG) {
                        uint32_t sa = qnxt_s + 8u * (uint32_t)slotq;
                        sts64_if(sa, c0, sbits | k0, c0);   sa += c0 ? 8u : 0u;
                        sts64_if(sa, c1, sbits | k1, c1);   sa += c1 ? 8u : 0u;
                        sts64_if(sa, c2, sbits | k2, c2);   sa += c2 ? 8u : 0u;
                        sts64_if(sa, c3, sbits | k3, c3);   sa += c3 ? 8u : 0u;
                        sts64_if(sa, c4c, sbits | k4, c4c);
                    } else if (wn + tot <= PSEG + PGSEG) {
                        const uint32_t cc[5] = {c0, c1, c2, c3, c4c};
                        const uint32_t kk[5] = {k0, k1, k2, k3, k4};
                        #pragma unroll
                        for (int d = 0; d < 5; ++d) {
                            if (cc[d]) {
                                if (slotq < PSEG) sts64_if(qnxt_s + 8u * (uint32_t)slotq, cc[d], sbits | kk[d], 1u);
                                else gnxt[slotq - PSEG] = make_uint2(cc[d], sbits | kk[d]);
                                ++slotq;
                            }
                        }
                    } else {
                        wovf = true;
                    }
                    wn += tot;
                }
            }
            if (lane == 0) {
                int4 c = make_int4(wovf ? 0x8000 : wn, 0, 0, 0);
                #pragma unroll
                for (int s = 0; s < PS; ++s) {
                    const int wd = (ccnt[s] & 0xFFFF) | (wset[s] << 16);
                    if (s == 0) c.y = wd; else if (s == 1) c.z = wd; else c.w = wd;
                }
                ctl.wc[cur ^ 1][warp] = c;
            }
            if (tid == 0) ctl.wc[cur ^ 1][PNW] = make_int4(0, 0, 0, 0);   // seeds only sit in a buffer a service section filled
            #pragma unroll
            for (int s = 0; s < PS; ++s) jv[s] += (running >> s) & 1u;
            fresh = 0u;                // every seed has been consumed by this wave
            __syncthreads();
            cur ^= 1;
        }
        // ---- leaving the waves: scrub the queue entries of the slots that finished ----
        __syncthreads();
        {
            int pre[PNW + 2];
            pre[0] = 0;
            #pragma unroll
            for (int w = 0; w <= PNW; ++w) pre[w + 1] = pre[w] + (ctl.wc[cur][w].x & 0x7FFF);
            const int total = pre[PNW + 1];
            uint2* const qc = m.q0 + (size_t)cur * PQBUF;
            uint2* const gc = m.gq0 + (size_t)cur * kFastGQCap;
            for (int i = tid; i < total; i += PNT) {
                const EntryRef er = entry_ref(i, pre);
                uint2* const ep = er.sg == PNW ? qc + kPoolQCap + er.off
                                : er.off < PSEG ? qc + er.sg * PSEG + er.off : gc + (size_t)er.sg * PGSEG + (er.off - PSEG);
                const uint32_t y = ep->y;
                if ((y >> 16) < (uint32_t)PS && ((finmask >> (y >> 16)) & 1u)) ep->y = y | 0xFFFF0000u;
            }
        }
        __syncthreads();
    }
}

void launch_k2_pool(const K2Params& p, int grid, cudaStream_t stream) {
    static unsigned long long attr_set = 0ull;
    const size_t smem = k2_pool_smem_bytes(p.max_px);
    if (first_use_on_device(attr_set)) {
        cudaFuncSetAttribute(k2_pool_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)k2_pool_smem_bytes(kFastMaxPx));
        cudaFuncSetAttribute(k2_pool_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    }
    K2Params q = p;
    q.hash_smem = 0;
    q.sparse_words = kPoolWords;
    k2_pool_kernel<<<grid > 0 ? grid : 1, PNT, smem, stream>>>(q);
}

}  // namespace axb
