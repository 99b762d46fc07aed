// K2 fast kernel (v5) -- level-0 bounded multi-target search with WORD-PARALLEL wavefronts (sm_100a).
//
// Same contract as the per-cell kernel it replaces (see axb_search.cu header and DESIGN.md "K2
// semantics"): one search per SOURCE detection answers every in-range target of the source's
// chain of frame pairs; replaces pyastar2d/src/cpp/astar.cpp:46-172 driven by
// axtrack/AxonDetections.py:553-579,600-629 for every source whose mask component reaches hop
// max_px or contains all its in-range targets.  Sources it cannot answer (level 0 exhausted
// with targets open) are deferred to k2_general_kernel.
//
// What changed against v4 (profiles/r01a_*): v4 spent ~230 warp instructions per warp per
// wavefront on one lane per (frontier cell, direction) and was issue bound.  Here the window
// bitmap is stored as 8x4-pixel TILE WORDS (bit = (y&3)*8 + (x&7)); a frontier entry is a
// (tile word, 32 frontier bits) pair and ONE thread expands all cells of the word with five
// shift/mask candidates (same word, left, right, up, down word) and at most five shared-memory
// atomicAnd claims.  A corridor wavefront of ~15 cells is 3-5 entries whatever its orientation.
//
// Window layout.  The L1 ball of radius RW = max_px around the source, in tile rows of 4 image
// rows aligned to (sy & ~3) and tile columns of 8 pixels aligned to (sx & ~31), FOLDED on a
// circle of 128 tile columns (1024 px): tile rows tr <= 0 live in folded row tr+NR-1 with the
// source's aligned column at tile column 64, tile rows tr >= 1 live in folded row tr-1 with it
// at tile column 0; NR = ceil(RW/4)+1.  The half widths of the two tile rows that share a folded
// row add up to <= RW-1, so their pixel ranges never overlap on the 1024 px circle.  Only
// in-diamond, in-image mask pixels are staged as open, and cells at hop RW are never expanded,
// so every candidate bit of an expansion is a true 4-neighbour inside the same half: no bound
// checks in the wavefront.  64.5 KB at max_px = 500 -> 3 CTAs per SM.
#include "axb_common.cuh"

namespace axb {

namespace {

constexpr unsigned kFull = 0xFFFFFFFFu;
constexpr int kTC = 128;                  // tile columns per folded row

__device__ __forceinline__ uint32_t span_mask(int a, int b) {
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

// meta of a tile word: half << 15 | shared-memory word index, where the word index is
// folded row << 7 | physical column and the 128 words of folded row fr are rotated by kRot*fr
// (physical column = (tile column + kRot*fr) & 127) so that vertical runs of tile words spread
// over the banks.
constexpr int kRot = 0;
__device__ __forceinline__ uint32_t make_meta(int half, int fr, int col) {
    return ((uint32_t)half << 15) | ((uint32_t)fr << 7) | (uint32_t)((col + kRot * fr) & (kTC - 1));
}
__device__ __forceinline__ int phys_word(uint32_t meta) { return (int)(meta & 0x7FFFu); }
// cell (dy, dx) relative to the source -> meta and bit of its tile word
__device__ __forceinline__ uint32_t cell_meta(int dy, int dx, int oy, int ox, int NR, int& bit) {
    const int cy = dy + oy, cx = dx + ox;
    const int tr = cy >> 2;
    const int half = tr > 0;
    const int fr = half ? tr - 1 : tr + NR - 1;
    const int col = ((cx >> 3) + (half ? 0 : 64)) & (kTC - 1);
    bit = ((cy & 3) << 3) | (cx & 7);
    return make_meta(half, fr, col);
}

__device__ __forceinline__ uint32_t hash_slot(uint32_t key) { return (key * 0x9E3779B1u) >> 22; }
static_assert(kHashSize == 1024, "hash_slot keeps 10 bits");
static_assert(kCandCap <= 2048, "hash entries keep 11 bits of candidate slot");

struct Chain {
    long long orow[kMaxChain];      // first output element of this source's row, per member pair
    long long item_orig[kMaxChain]; // work index in the reference's iteration order (errors)
    int dst_off[kMaxChain];
    int cum[kMaxChain + 1];         // flattened target index space
    int nm;
};

struct __align__(16) Ctl {
    // per-warp (entries appended | overflow << 15 | targets reached << 16) of the last wave, double
    // buffered: wave s reads wc[s&1], every warp stores its own word of wc[(s+1)&1] at its end
    int wc[2][4];
    Chain ch;
    int ncand;
    int item;
};

struct Win {
    const uint32_t* mbits;
    int H, W, Wp, sy, sx, RW, NR, oy, ox, iw0;
};

// one aligned word of the packed mask of image row y, restricted to the diamond row span
__device__ __forceinline__ uint32_t row_word(const Win& w, int dy, int iw) {
    const int hw = w.RW - abs(dy);
    const int y = w.sy + dy;
    if (hw < 0 || y < 0 || y >= w.H || iw < 0 || iw >= w.Wp) return 0u;
    const int xlo = max(0, w.sx - hw), xhi = min(w.W - 1, w.sx + hw);
    const uint32_t valid = span_mask(xlo - 32 * iw, xhi - 32 * iw);
    if (!valid) return 0u;
    return __ldg(w.mbits + (size_t)y * w.Wp + iw) & valid;
}

// the four tile words (tile columns 4q..4q+3 of the quad) of tile row tr and image word iw
__device__ __forceinline__ void quad_tiles(const Win& w, int tr, int iw, uint32_t t[4]) {
    const int dy0 = 4 * tr - w.oy;
    const uint32_t r0 = row_word(w, dy0, iw), r1 = row_word(w, dy0 + 1, iw);
    const uint32_t r2 = row_word(w, dy0 + 2, iw), r3 = row_word(w, dy0 + 3, iw);
    #pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t lo = __byte_perm(r0, r1, k | ((4 + k) << 4));
        const uint32_t hi = __byte_perm(r2, r3, k | ((4 + k) << 4));
        t[k] |= __byte_perm(lo, hi, 0x5410);
    }
}

// Stage the folded window: one warp per folded row, lane = quad of 4 tile columns = one aligned
// word of the packed mask per image row.
template <int NT>
__device__ __forceinline__ void stage_window(const Win& w, uint32_t* OF) {
    const int lane = threadIdx.x & 31;
    for (int fr = threadIdx.x >> 5; fr < w.NR; fr += NT / 32) {
        uint32_t t[4] = {0u, 0u, 0u, 0u};
        {   // upper half: tile row fr-NR+1 <= 0, quad q holds cx in [32(q-16), 32(q-16)+31]
            const int tr = fr - w.NR + 1;
            quad_tiles(w, tr, w.iw0 - 16 + lane, t);
            if (lane == 0) quad_tiles(w, tr, w.iw0 + 16, t);
        }
        if (fr + 1 <= w.NR - 1) {   // lower half: tile row fr+1, quad q holds cx in [32q, ..] (q<16) or [32(q-32), ..]
            const int tr = fr + 1;
            quad_tiles(w, tr, w.iw0 + (lane < 16 ? lane : lane - 32), t);
            if (lane == 16) quad_tiles(w, tr, w.iw0 + 16, t);
        }
        static_assert(kRot == 0, "the quad is stored with one 16-byte store");
        *reinterpret_cast<uint4*>(OF + fr * kTC + 4 * lane) = make_uint4(t[0], t[1], t[2], t[3]);
    }
}

// Slow path of the target probe: the hash slot of tile word `key` is occupied.  The cells
// `claimed` of the word were reached with j mask pixels on the path (hop j, path length j+1).
__device__ __noinline__ int probe_settle(const K2Params& p, const uint32_t* hash, const uint32_t* cand,
                                         const Chain& ch, uint32_t key, uint32_t claimed, int j) {
    int n = 0;
    uint32_t h = hash_slot(key);
    for (;;) {
        const uint32_t e = hash[h];
        if (e == kEmpty) break;
        if ((e >> 16) == key && ((claimed >> ((e >> 11) & 31u)) & 1u)) {
            const uint32_t c = cand[e & 0x7FFu];
            const long long o = ch.orow[c >> 30] + (long long)(c & 0x3FFFFFFFu);
            p.path_len[o] = j + 1;
            p.unreachable[o] = 0;
            if (p.wcost) p.wcost[o] = (float)j;
            ++n;
        }
        h = (h + 1) & (kHashSize - 1);
    }
    return n;
}

}  // namespace




#ifdef AXB_TIMING
__device__ unsigned long long g_dbg[8];   // cycles: [0] setup+candidates [1] staging [2] waves; [3] #waves [4] #items [5] entries [6] warp-0 wave body [7] barrier
#define AXB_T(k, t0) do { if (threadIdx.x == 0) { const long long t1_ = clock64(); atomicAdd(&g_dbg[k], (unsigned long long)(t1_ - t0)); t0 = t1_; } } while (0)
#define AXB_C(k, v) do { if (threadIdx.x == 0) atomicAdd(&g_dbg[k], (unsigned long long)(v)); } while (0)
#else
#define AXB_T(k, t0) do { } while (0)
#define AXB_C(k, v) do { } while (0)
#endif

namespace {

constexpr int kNW = kFastThreads / 32;            // warps per search
constexpr int kSeg = kFastQCap / kNW;             // shared-memory frontier entries per warp per buffer
constexpr int kGSeg = kFastGQCap / kNW;           // global overflow entries per warp per buffer
static_assert(kNW == 4, "Ctl::wc holds one word per warp in one 16-byte row");
static_assert(kSeg + kGSeg < 0x8000, "the per-warp wave counter keeps 15 bits of entries");

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// Shared-memory operations of the wavefront as predicated PTX, so that the five directions of an
// entry issue back to back instead of through five branches.
// value at shared address `a`, or 0 when want == 0 (the address may then be invalid)
__device__ __forceinline__ uint32_t lds_if(uint32_t a, uint32_t want) {
    uint32_t v;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, 0;\n\tmov.u32 %0, 0;\n\t@p ld.shared.u32 %0, [%1];\n\t}"
                 : "=r"(v) : "r"(a), "r"(want) : "memory");
    return v;
}
// clear the bits `o` of the word at shared address `a`; returns those of them this thread cleared
__device__ __forceinline__ uint32_t claim_if(uint32_t a, uint32_t o) {
    uint32_t old;
    asm volatile("{\n\t.reg .pred p;\n\t.reg .b32 n;\n\tsetp.ne.u32 p, %2, 0;\n\tnot.b32 n, %2;\n\tmov.u32 %0, 0;\n\t"
                 "@p atom.shared.and.b32 %0, [%1], n;\n\t}"
                 : "=r"(old) : "r"(a), "r"(o) : "memory");
    return old & o;
}
__device__ __forceinline__ void sts64_if(uint32_t a, uint32_t x, uint32_t y, uint32_t want) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %3, 0;\n\t@p st.shared.v2.u32 [%0], {%1, %2};\n\t}"
                 :: "r"(a), "r"(x), "r"(y), "r"(want) : "memory");
}

// One work item = one source of a chain leader, processed by the whole CTA.  Returns true when
// the item must be redone by the general kernel.  All control flow is CTA-uniform.
__device__ bool run_item_fast(const K2Params& p, const int item, uint32_t* smem, uint32_t* gscr, Ctl& ctl) {
    constexpr int NT = kFastThreads;
    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const uint32_t lt = lanemask_lt();
    const int RW = p.max_px;
    const int NR = (RW + 3) / 4 + 1;

    uint32_t* const OF = smem;
    uint2* const q0 = (uint2*)(OF + NR * kTC);            // [2][kNW][kSeg] (bits, meta)
    uint32_t* const hash = (uint32_t*)(q0 + 2 * kFastQCap);
    uint32_t* const cand = gscr;
    uint2* const gq0 = (uint2*)(gscr + kCandCap);         // [2][kNW][kGSeg] global overflow
    Chain& ch = ctl.ch;

    long long tclk = clock64();
    (void)tclk;
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
            ch.item_orig[nm] = (long long)p.all_prefix[q] + si;
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

    Win w;
    w.sy = p.dets_yx[2 * ((size_t)lead.src_off + si)];
    w.sx = p.dets_yx[2 * ((size_t)lead.src_off + si) + 1];
    w.H = p.H; w.W = p.W; w.Wp = p.Wp; w.RW = RW; w.NR = NR;
    w.mbits = p.mask_bits + (size_t)lead.mask_idx * p.H * p.Wp;
    w.oy = w.sy & 3; w.ox = w.sx & 31; w.iw0 = w.sx >> 5;     // only used when the source is in the grid
    const bool src_in = w.sy >= 0 && w.sy < p.H && w.sx >= 0 && w.sx < p.W;
    const long long r2 = (long long)RW * RW;
    constexpr int kPassLoad = kHashMaxLoad > NT ? kHashMaxLoad - NT : 0;

    int tt = 0;                        // next target of the flattened chain target list
    for (;;) {                         // passes of <= kHashMaxLoad candidates (one on real data)
        for (int i = tid; i < kHashSize; i += NT) hash[i] = kEmpty;
        if (tid == 0) ctl.ncand = 0;
        for (;;) {                     // groups of NT targets
            __syncthreads();
            if (tt >= total_t || ctl.ncand > kPassLoad) break;
            const int t = tt + tid;
            bool is_cand = false, self = false;
            int m = 0, j = 0, dy = 0, dx = 0;
            if (t < total_t) {
                while (m + 1 < nm && t >= ch.cum[m + 1]) ++m;
                j = t - ch.cum[m];
                const int ty = p.dets_yx[2 * ((size_t)ch.dst_off[m] + j)];
                const int tx = p.dets_yx[2 * ((size_t)ch.dst_off[m] + j) + 1];
                dy = ty - w.sy; dx = tx - w.sx;
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
                    int bit;
                    const uint32_t key = cell_meta(dy, dx, w.oy, w.ox, NR, bit);
                    const uint32_t entry = (key << 16) | ((uint32_t)bit << 11) | slot;
                    uint32_t h = hash_slot(key);
                    while (atomicCAS(&hash[h], kEmpty, entry) != kEmpty) h = (h + 1) & (kHashSize - 1);
                }
            }
            tt += NT;
        }
        const int ncand = ctl.ncand;
        if (ncand > 0) {
            AXB_T(0, tclk);
            stage_window<NT>(w, OF);
            __syncthreads();
            AXB_T(1, tclk);
            if (tid == 0) {            // close the source cell, seed the search with it
                int bit;
                const uint32_t key = cell_meta(0, 0, w.oy, w.ox, NR, bit);
                OF[phys_word(key)] &= ~(1u << bit);
                q0[0] = make_uint2(1u << bit, key << 2);
                ctl.wc[0][0] = 1; ctl.wc[0][1] = 0; ctl.wc[0][2] = 0; ctl.wc[0][3] = 0;
            }
            __syncthreads();

            // ---- wavefronts: one thread per frontier entry (tile word, frontier bits) -----------
            // queue metas are BYTE offsets of the tile word in the window | half << 17 (= key << 2)
            const uint32_t of_s = smem_u32(OF);
            const uint32_t hash_s = smem_u32(hash);
            const uint32_t lastU = (uint32_t)(NR - 1);
            int unsettled = ncand;     // CTA-uniform: updated from ctl.wc[] after each barrier
            bool hit = false, defer = false;
            for (int j = 0;; ++j) {    // wave j: probe + expand the cells at hop j
                const int4 c4 = *reinterpret_cast<const int4*>(ctl.wc[j & 1]);
                const int p1 = c4.x & 0x7FFF, p2 = p1 + (c4.y & 0x7FFF), p3 = p2 + (c4.z & 0x7FFF);
                const int total = p3 + (c4.w & 0x7FFF);
                unsettled -= (c4.x >> 16) + (c4.y >> 16) + (c4.z >> 16) + (c4.w >> 16);
                if ((c4.x | c4.y | c4.z | c4.w) & 0x8000) { defer = true; break; }   // a frontier list overflowed
                if (total == 0 || unsettled <= 0) break;
                if (j >= RW) { hit = true; break; }                   // cells at hop == max_px exist
                AXB_C(3, 1); AXB_C(5, total);
#ifdef AXB_TIMING
                long long tw = clock64();
#endif
                const uint2* const qcur = q0 + (j & 1) * kFastQCap;
                const uint32_t qnxt_s = smem_u32(q0 + ((j + 1) & 1) * kFastQCap + warp * kSeg);
                const uint2* const gcur = gq0 + (size_t)(j & 1) * kFastGQCap;
                uint2* const gnxt = gq0 + (size_t)((j + 1) & 1) * kFastGQCap + (size_t)warp * kGSeg;
                int wn = 0, ws = 0;
                bool wovf = false;
                for (int base = warp * 32; base < total; base += NT) {
                    const int i = base + lane;
                    const int sg = (i >= p1) + (i >= p2) + (i >= p3);
                    const int off = i - (sg == 0 ? 0 : sg == 1 ? p1 : sg == 2 ? p2 : p3);
                    uint32_t f = 0u, mb = 0u;
                    if (i < total) {
                        const uint2 e = off < kSeg ? qcur[sg * kSeg + off] : gcur[(size_t)sg * kGSeg + (off - kSeg)];
                        f = e.x; mb = e.y;
                    }
                    // candidates: cells of the word itself, of the right / left / lower / upper word
                    const uint32_t cd0 = ((f << 1) & 0xFEFEFEFEu) | ((f >> 1) & 0x7F7F7F7Fu) | (f << 8) | (f >> 8);
                    const uint32_t cd1 = (f & 0x80808080u) >> 7;
                    const uint32_t cd2 = (f & 0x01010101u) << 7;
                    uint32_t cd3 = f >> 24, cd4 = f << 24;
                    const uint32_t rowm = mb & ~0x1FCu;
                    const uint32_t m1 = rowm | ((mb + 4u) & 0x1FCu);
                    const uint32_t m2 = rowm | ((mb - 4u) & 0x1FCu);
                    uint32_t m3 = mb + 512u, m4 = mb - 512u;
                    const uint32_t hr = mb >> 9;                      // half << 8 | folded row
                    if (hr == lastU) m3 = 0x20000u | ((mb + 256u) & 0x1FCu);          // tile row 0 -> 1
                    if (hr == 256u) m4 = (lastU << 9) | ((mb + 256u) & 0x1FCu);      // tile row 1 -> 0
                    if (hr == 255u + lastU) cd3 = 0u;                 // window borders (never open anyway)
                    if (hr == 0u) cd4 = 0u;
                    const uint32_t a0 = of_s + (mb & 0x1FFFFu), a1 = of_s + (m1 & 0x1FFFFu), a2 = of_s + (m2 & 0x1FFFFu);
                    const uint32_t a3 = of_s + (m3 & 0x1FFFFu), a4 = of_s + (m4 & 0x1FFFFu);
                    const uint32_t hv = lds_if(hash_s + 4u * hash_slot(mb >> 2), f);
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
                    // targets among the cells being expanded (hop j, path length j + 1)
                    if (__any_sync(kFull, f != 0u && hv != kEmpty)) {
                        int ns = 0;
                        if (f != 0u && hv != kEmpty) ns = probe_settle(p, hash, cand, ch, mb >> 2, f, j);
                        ws += __reduce_add_sync(kFull, ns);
                    }
                    // append the claims to this warp's segment of the next frontier, lane-major
                    const uint32_t b0 = __ballot_sync(kFull, c0 != 0u), b1 = __ballot_sync(kFull, c1 != 0u);
                    const uint32_t b2 = __ballot_sync(kFull, c2 != 0u), b3 = __ballot_sync(kFull, c3 != 0u);
                    const uint32_t b4 = __ballot_sync(kFull, c4c != 0u);
                    const int tot = __popc(b0) + __popc(b1) + __popc(b2) + __popc(b3) + __popc(b4);
                    if (tot) {
                        int slot = wn + __popc(b0 & lt) + __popc(b1 & lt) + __popc(b2 & lt) + __popc(b3 & lt) + __popc(b4 & lt);
                        if (wn + tot <= kSeg) {
                            uint32_t sa = qnxt_s + 8u * (uint32_t)slot;
                            sts64_if(sa, c0, mb, c0);   sa += c0 ? 8u : 0u;
                            sts64_if(sa, c1, m1, c1);   sa += c1 ? 8u : 0u;
                            sts64_if(sa, c2, m2, c2);   sa += c2 ? 8u : 0u;
                            sts64_if(sa, c3, m3, c3);   sa += c3 ? 8u : 0u;
                            sts64_if(sa, c4c, m4, c4c);
                        } else if (wn + tot <= kSeg + kGSeg) {
                            const uint32_t cc[5] = {c0, c1, c2, c3, c4c};
                            const uint32_t mm[5] = {mb, m1, m2, m3, m4};
                            #pragma unroll
                            for (int d = 0; d < 5; ++d) {
                                if (cc[d]) {
                                    if (slot < kSeg) sts64_if(qnxt_s + 8u * (uint32_t)slot, cc[d], mm[d], 1u);
                                    else gnxt[slot - kSeg] = make_uint2(cc[d], mm[d]);
                                    ++slot;
                                }
                            }
                        } else {
                            wovf = true;
                        }
                        wn += tot;
                    }
                }
                if (lane == 0) ctl.wc[(j + 1) & 1][warp] = wovf ? 0x8000 : (wn | (ws << 16));
                AXB_T(6, tw);
                __syncthreads();
#ifdef AXB_TIMING
                if (ctl.wc[(j + 1) & 1][0] != -12345) AXB_T(7, tw);
#endif
            }
            AXB_T(2, tclk); AXB_C(4, 1);
            if (defer || (!hit && unsettled > 0)) return true;        // level 0 cannot answer: defer
        }
        if (tt >= total_t) break;
    }
    return false;
}

}  // namespace

__global__ void __launch_bounds__(kFastThreads, kFastCtasPerSm) k2_fast_kernel(const K2Params p) {
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ Ctl ctl;
    uint32_t* const gscr = p.fast_scratch + (size_t)blockIdx.x * kFastScratchWords;
    const int total = p.work_prefix[p.n_pairs];
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) ctl.item = atomicAdd(&p.counters[0], 1);
        __syncthreads();
        const int item = ctl.item;
        if (item >= total) break;
        const bool defer = run_item_fast(p, item, smem, gscr, ctl);
        if (defer && threadIdx.x == 0) p.gen_list[atomicAdd(&p.counters[1], 1)] = item;
    }
}

void launch_k2_fast(const K2Params& p, int grid, cudaStream_t stream) {
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(k2_fast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)k2_fast_smem_bytes(AXB_MAX_PX));
        cudaFuncSetAttribute(k2_fast_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        attr_set = true;
    }
    k2_fast_kernel<<<grid, kFastThreads, k2_fast_smem_bytes(p.max_px), stream>>>(p);
}

#ifdef AXB_TIMING
extern "C" void axb_dbg_fetch(unsigned long long* out, int reset) {
    cudaMemcpyFromSymbol(out, g_dbg, sizeof(g_dbg));
    if (reset) { unsigned long long z[8] = {0}; cudaMemcpyToSymbol(g_dbg, z, sizeof(z)); }
}
#endif

}  // namespace axb
