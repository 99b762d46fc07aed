// Shared device helpers of the tile-word search kernels (axb_search_fast.cu):
// window geometry, staging of the folded sparse / dense window, predicated shared-memory
// operations, the target probe.  See the header of axb_search_fast.cu for the layout.
#pragma once
#include "axb_common.cuh"
#include <cstdlib>
#ifdef AXB_CHECK       // `make check`: index asserts in the search kernels (compute-sanitizer stand-in), never shipped
#undef NDEBUG
#include <cassert>
#define AXB_ASSERT(c) assert(c)
#else
#define AXB_ASSERT(c) ((void)0)
#endif

namespace axb {

namespace {

constexpr unsigned kFull = 0xFFFFFFFFu;
constexpr int kTC = 128;                  // tile columns per folded row (max_px <= 511)

// Geometry of the folded window for a circle of 2^LOGC tile columns: LOGC = 7 (1024 px, max_px <= 511) and
// LOGC = 8 (2048 px, max_px <= 1023; the reference honours any max_px_assoc_dist, AxonDetections.py:623-625,736).
template <int LOGC> struct TW {
    static constexpr int TC = 1 << LOGC;                 // tile columns of a folded row
    static constexpr int HALF = TC / 2;                  // where the upper half's source column sits
    static constexpr int FRB = LOGC + 1;                 // bits of a folded row index (NR <= 129 / 257)
    static constexpr int WB = LOGC + FRB;                // bits of a window word index = folded row << LOGC | column
    static constexpr uint32_t WMASK = (1u << WB) - 1u;
    static constexpr uint32_t HALFBIT = 1u << WB;        // key of a tile word = half << WB | word index
    static constexpr int KB = WB + 1;                    // key bits: 16 / 18
    static constexpr int ESH = 32 - KB;                  // target-hash entry = key << ESH | bit << SLOTB | candidate slot (128 columns) ...
    static constexpr bool BIT_IN_ENTRY = ESH >= 16;      // ... or key << ESH | candidate slot, the cell's bit in a byte table of the CTA's scratch (256 columns)
    static constexpr int SLOTB = BIT_IN_ENTRY ? ESH - 5 : ESH;   // 11 / 14 bits of candidate slot
    static constexpr int GROUPS = TC / 32;               // presence-table entries per folded row
};
static_assert(TW<7>::HALFBIT == 0x8000u && TW<7>::ESH == 16 && TW<7>::SLOTB == 11, "the 128-column layout of round 1");

__device__ __forceinline__ uint32_t lanemask_lt() {
    uint32_t m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

template <int LOGC = 7>
__device__ __forceinline__ uint32_t make_key(int half, int fr, int col) {
    return ((uint32_t)half << TW<LOGC>::WB) | ((uint32_t)fr << LOGC) | (uint32_t)col;
}
// cell (dy, dx) relative to the source -> key and bit of its tile word
template <int LOGC = 7>
__device__ __forceinline__ uint32_t cell_key(int dy, int dx, int oy, int ox, int NR, int& bit) {
    const int cy = dy + oy, cx = dx + ox;
    const int tr = cy >> 2;
    const int half = tr > 0;
    const int fr = half ? tr - 1 : tr + NR - 1;
    const int col = ((cx >> 3) + (half ? 0 : TW<LOGC>::HALF)) & (TW<LOGC>::TC - 1);
    bit = ((cy & 3) << 3) | (cx & 7);
    return make_key<LOGC>(half, fr, col);
}

template <int HASH> __device__ __forceinline__ uint32_t hash_slot(uint32_t key);
template <> __device__ __forceinline__ uint32_t hash_slot<2048>(uint32_t key) { return (key * 0x9E3779B1u) >> 21; }
template <> __device__ __forceinline__ uint32_t hash_slot<1024>(uint32_t key) { return (key * 0x9E3779B1u) >> 22; }
template <> __device__ __forceinline__ uint32_t hash_slot<512>(uint32_t key) { return (key * 0x9E3779B1u) >> 23; }
static_assert(kCandCap <= 2048 && kSparseHashMaxLoad <= 512, "hash entries keep 11 (wide layout: 14) bits of candidate slot");

// Members of a chain: forward pairs (this source's row of the pair's matrix, targets = the pair's dst detections)
// and reverse pairs (k2_prep_kernel "cover": this detection is a TARGET of the pair; the search answers its column,
// the flattened "targets" of the member are the pair's SOURCE detections).
struct Chain {
    long long orow[kMaxMem];        // output element of flattened target 0 of the member ...
    int ebase[kMaxMem];             // first work index of the pair in the reference's iteration order (errors; + K2Params::orig_base)
    int ostride[kMaxMem];           // ... and the step to the next one: 1 (forward: along the row) or n_dst (reverse: down the column)
    int dst_off[kMaxMem];           // detections the flattened targets are
    int ditem[kMaxMem];             // reverse: work item of flattened target 0 (its own directed search, if it comes to that)
    int cum[kMaxMem + 1];           // flattened target index space
    int nm;
    int si;                         // index of this detection in its set
    uint32_t revmask;               // bit m: member m is a reverse member
    int mask_idx;
};

struct __align__(16) Ctl {
    // per-warp (entries appended | overflow << 15 | targets reached << 16) of the last wave, double
    // buffered: wave s reads wc[s&1], every warp stores its own word of wc[(s+1)&1] at its end
    int wc[2][8];
    Chain ch;
    int ncand;
    int item;
    int alloc;         // sparse: packed tile words allocated so far
    int nchk;          // reverse candidates whose None needs the directed search unless this search settles them
};

struct Win {
    const uint32_t* tiles;       // tiled mask of this source's mask
    int TH, TWp, ty0, tx0;       // tile grid, global tile of the window's (tile row 0, tile column 0)
    int RW, NR, oy, ox;
};

// Four adjacent tiles (a 16-byte aligned quad: 32 x 4 pixels) of window tile row tr starting at
// tile column tcq (a multiple of 4): where they live in the tiled mask, or nullptr when no pixel
// of the quad lies inside the diamond or the image.  x0 = dx of the quad's first pixel.
__device__ __forceinline__ const uint4* window_quad(const Win& w, int tr, int tcq, int hw_out, int& x0) {
    x0 = 8 * tcq - w.ox;
    if (x0 + 31 < -hw_out || x0 > hw_out) return nullptr;
    const int ty = w.ty0 + tr, tx = w.tx0 + tcq;        // tx0, tcq and TWp are multiples of 4
    if (ty < 0 || ty >= w.TH || tx < 0 || tx >= w.TWp) return nullptr;
    return reinterpret_cast<const uint4*>(w.tiles + (size_t)ty * w.TWp + tx);
}

// Clear the pixels of a loaded quad that lie outside the diamond (only when its edge crosses the quad).
__device__ __forceinline__ void clip_quad(const Win& w, int tr, int x0, int hw_in, uint4& v) {
    if (x0 >= -hw_in && x0 + 31 <= hw_in) return;
    const int dy0 = 4 * tr - w.oy;
    uint32_t rm[4];                                     // the quad's 32 pixels of each of its 4 image rows
    #pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int hw = w.RW - abs(dy0 + r);
        const int lo = max(-hw - x0, 0), hi = min(hw - x0, 31);
        rm[r] = (hw >= 0 && lo <= hi) ? ((0xFFFFFFFFu >> (31 - hi)) & (0xFFFFFFFFu << lo)) : 0u;
    }
    uint32_t m[4];                                      // tile k = byte k of the four rows
    #pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t lo = __byte_perm(rm[0], rm[1], k | ((4 + k) << 4));
        const uint32_t hi = __byte_perm(rm[2], rm[3], k | ((4 + k) << 4));
        m[k] = __byte_perm(lo, hi, 0x5410);
    }
    v.x &= m[0]; v.y &= m[1]; v.z &= m[2]; v.w &= m[3];
}

__device__ __forceinline__ void row_half_widths(const Win& w, int tr, int& hw_in, int& hw_out) {
    const int a = 4 * tr - w.oy, b = a + 3;             // dy of the first / last image row
    const int far = max(abs(a), abs(b));
    const int near = (a <= 0 && b >= 0) ? 0 : min(abs(a), abs(b));
    hw_in = w.RW - far;                                 // may be negative: no tile is interior
    hw_out = w.RW - near;
}

}  // namespace

#ifdef AXB_TIMING
__device__ unsigned long long g_dbg[24];       // (this header has one includer)   // cycles: [0] setup+candidates [1] staging [2] waves; [3] #waves [4] #items [5] entries [6] warp-0 wave body [7] barrier
#define AXB_T(k, t0) do { if (threadIdx.x == 0) { const long long t1_ = clock64(); atomicAdd(&g_dbg[k], (unsigned long long)(t1_ - t0)); t0 = t1_; } } while (0)
#define AXB_C(k, v) do { if (threadIdx.x == 0) atomicAdd(&g_dbg[k], (unsigned long long)(v)); } while (0)
// phase timer of the per-warp search: the clock is read after `dep` is available (scoreboard wait)
#define AXB_P(k, dep) do { unsigned long long t1_; asm volatile("{\n\t.reg .u32 t;\n\tadd.u32 t, %1, 0;\n\tmov.u64 %0, %%clock64;\n\t}" : "=l"(t1_) : "r"((uint32_t)(dep)) : "memory"); ph[k] += t1_ - tph; tph = t1_; } while (0)
#else
#define AXB_P(k, dep) do { } while (0)
#define AXB_T(k, t0) do { } while (0)
#define AXB_C(k, v) do { } while (0)
#endif

namespace {

constexpr int kNW = kFastThreads / 32;            // warps per search
constexpr int kGSeg = kFastGQCap / kNW;           // global overflow entries per warp per buffer
constexpr int kWideNW = 8;                        // warps per search of the 256-column kernels (about 100 entries per wave, 2 CTAs per SM)
static_assert(kNW >= 1 && kNW <= 4 && kWideNW <= 8, "Ctl::wc holds one word per warp in one / two 16-byte rows");

template <bool SPARSE, int LOGC = 7> struct Cfg;
template <> struct Cfg<false, 7> {
    static constexpr int kQCap = kFastQCap, kHash = kHashSize, kMaxLoad = kHashMaxLoad;
};
template <> struct Cfg<true, 7> {
    static constexpr int kQCap = kSparseQCap, kHash = kSparseHash, kMaxLoad = kSparseHashMaxLoad;
};
template <> struct Cfg<true, 8> {      // max_px 512 .. 1023: about twice the frontier entries per wave and four times the in-range targets
    static constexpr int kQCap = kWideQCap, kHash = kWideHash, kMaxLoad = kWideHashMaxLoad;
};
static_assert(kFastQCap / kNW + kGSeg < 0x8000 && kWideQCap / kWideNW + kFastGQCap / kWideNW < 0x8000, "the per-warp wave counter keeps 15 bits of entries");
static_assert(kWideHashMaxLoad <= (1 << TW<8>::SLOTB) && kWideHashMaxLoad <= kCandCap, "wide hash entries keep 14 bits of candidate slot");

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// Shared-memory operations of the wavefront as predicated PTX, so that the five directions of an
// entry issue back to back.
// value at shared address `a`, or 0 when want == 0 (the address may then be invalid)
__device__ __forceinline__ uint32_t lds_if(uint32_t a, uint32_t want) {
    uint32_t v;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, 0;\n\tmov.u32 %0, 0;\n\t@p ld.shared.u32 %0, [%1];\n\t}"
                 : "=r"(v) : "r"(a), "r"(want) : "memory");
    return v;
}
__device__ __forceinline__ uint2 lds64_if(uint32_t a, uint32_t want) {
    uint2 v;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %3, 0;\n\tmov.u32 %0, 0;\n\tmov.u32 %1, 0;\n\t@p ld.shared.v2.u32 {%0, %1}, [%2];\n\t}"
                 : "=r"(v.x), "=r"(v.y) : "r"(a), "r"(want) : "memory");
    return v;
}
// clear the bits `o` of the word at shared address `a`; returns those of them this thread cleared.  With o == 0 the
// atomic goes to `idle` (any valid shared word: AND with all ones changes nothing) -- ptxas turns a predicated
// atom.shared into a branch with a reconvergence point, and a lone branch costs a warp ~13 issue cycles
// (ncu source view, stall_branch_resolving), five of them per frontier entry.
__device__ __forceinline__ uint32_t claim_if(uint32_t a, uint32_t o, uint32_t idle) {
    uint32_t old;
    asm volatile("{\n\t.reg .pred p;\n\t.reg .b32 n, b;\n\tsetp.ne.u32 p, %2, 0;\n\tnot.b32 n, %2;\n\tselp.b32 b, %1, %3, p;\n\t"
                 "atom.shared.and.b32 %0, [b], n;\n\t}"
                 : "=r"(old) : "r"(a), "r"(o), "r"(idle) : "memory");
    return old & o;
}
// value at global address `p`, or 0 when want == 0
__device__ __forceinline__ uint2 ldg64_if(const uint2* p, uint32_t want) {
    uint2 v;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %3, 0;\n\tmov.u32 %0, 0;\n\tmov.u32 %1, 0;\n\t@p ld.global.v2.u32 {%0, %1}, [%2];\n\t}"
                 : "=r"(v.x), "=r"(v.y) : "l"(p), "r"(want) : "memory");
    return v;
}
__device__ __forceinline__ void sts64_if(uint32_t a, uint32_t x, uint32_t y, uint32_t want) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %3, 0;\n\t@p st.shared.v2.u32 [%0], {%1, %2};\n\t}"
                 :: "r"(a), "r"(x), "r"(y), "r"(want) : "memory");
}

// Does the mask component of pixel (y, x) hold a pixel at L1 distance >= RW from it?  Then the level-0 wavefront of
// a search from (y, x) reaches hop RW (axb_components.cu) and every target it does not settle is None.
// 1: yes, 0: no, -1: (y, x) is off the mask.
__device__ __forceinline__ int component_reaches(const K2Params& p, int mask_idx, int y, int x, int RW) {
    const size_t base = (size_t)mask_idx * p.H * p.W;
    AXB_ASSERT(y >= 0 && y < p.H && x >= 0 && x < p.W);
    const int r = __ldg(p.comp_label + base + (size_t)y * p.W + x);
    AXB_ASSERT(r < p.H * p.W);
    if (r < 0) return -1;
    const int4 e = __ldg(p.comp_ext + base + r);
    return max(max(e.x - (y + x), e.y - (y - x)), max(e.z - (x - y), e.w + (y + x))) >= RW ? 1 : 0;
}

// A detection of a set that is not searched needs its own search after all: once per work item.  On the mask: the
// level-0 kernels again (directed list); off the mask its level 0 is at most its four neighbours, so straight to
// the all-levels kernel.
__device__ __forceinline__ void push_directed(const K2Params& p, int item, bool off_mask) {
    AXB_ASSERT(item >= 0 && item < p.work_prefix[p.n_pairs]);
    const uint32_t bit = 1u << (item & 31);
    if (atomicOr(&p.item_bits[item >> 5], bit) & bit) return;
    if (off_mask) p.gen_list[atomicAdd(&p.counters[kCtrGenLen], 1)] = item;
    else p.dir_list[atomicAdd(&p.counters[kCtrDirLen], 1)] = item;
}

// Slow path of the target probe: the hash slot of tile word `key` is occupied.  The cells
// `claimed` of the word were reached with j mask pixels on the path (hop j, path length j+1).
template <int HASH, int LOGC = 7>
__device__ __noinline__ int probe_settle(const K2Params& p, const uint32_t* hash, const uint32_t* cand,
                                         const Chain& ch, uint32_t key, uint32_t claimed, int j) {
    using G = TW<LOGC>;
    int n = 0;
    uint32_t h = hash_slot<HASH>(key);
    for (;;) {
        const uint32_t e = hash[h];
        if (e == kEmpty) break;
        const uint32_t slot = e & ((1u << G::SLOTB) - 1u);
        if ((e >> G::ESH) == key &&
            ((claimed >> (G::BIT_IN_ENTRY ? (e >> G::SLOTB) & 31u : (uint32_t)reinterpret_cast<const uint8_t*>(cand + kCandCap + 4 * (size_t)kFastGQCap + kSparseHash)[slot])) & 1u)) {
            const uint32_t c = cand[slot];
            const long long o = ch.orow[c >> 29] + (long long)(c & 0x0FFFFFFFu) * ch.ostride[c >> 29];
            p.path_len[o] = j + 1;
            p.unreachable[o] = 0;
            if (p.wcost) p.wcost[o] = (float)j;
            ++n;
        }
        h = (h + 1) & (HASH - 1);
    }
    return n;
}

// Stage the folded window from the tiled mask: one warp per folded row, lane = a quad of four
// folded tile columns (4*lane .. 4*lane+3).  A folded column holds tile column col-64 or col+64
// of the upper tile row and tile column col or col-128 of the lower one: the four candidate quads
// are loaded as 16-byte words, clipped to the diamond and ORed.  (Up to max_px = 504 at most one
// candidate per half is inside the row's span and at most one per folded word; above that a row
// can wrap onto its own first tile and the two halves can meet inside one word, but the pixels of
// the candidates stay disjoint, so the OR is exact at every radius.)
// Dense: WIN[fr*128 + col].  Sparse: the non-zero words of each group of 32 columns are packed
// behind each other in column order, PB[fr*4 + g] = (presence bits, index of the first one);
// the groups of a folded row are allocated together (one atomicAdd per row), rows in any order.
template <bool SPARSE>
__device__ __noinline__ void stage_window(const Win w, uint32_t* WIN, uint2* PB, int* alloc, const int cap_words,
                                          const int row0, const int row_step) {
    const int lane = threadIdx.x & 31;
    for (int fr = row0; fr < w.NR; fr += row_step) {
        const int tr_u = fr - w.NR + 1, tr_l = fr + 1;
        int hwi_u, hwo_u, hwi_l, hwo_l;
        row_half_widths(w, tr_u, hwi_u, hwo_u);
        row_half_widths(w, tr_l, hwi_l, hwo_l);
        if (tr_l > w.NR - 1) hwo_l = -64;               // no lower tile row: nothing in span
        uint32_t t[4] = {0u, 0u, 0u, 0u};
        #pragma unroll
        for (int h = 0; h < 2; ++h) {                   // upper, lower: addresses, loads back to back, clipping
            const int tr = h ? tr_l : tr_u, hwi = h ? hwi_l : hwi_u, hwo = h ? hwo_l : hwo_u;
            int xa, xb;
            const uint4* sa = window_quad(w, tr, h ? 4 * lane : 4 * lane - 64, hwo, xa);
            const uint4* sb = window_quad(w, tr, h ? 4 * lane - 128 : 4 * lane + 64, hwo, xb);
            uint4 va = sa ? __ldg(sa) : make_uint4(0u, 0u, 0u, 0u);
            uint4 vb = sb ? __ldg(sb) : make_uint4(0u, 0u, 0u, 0u);
            if (sa) clip_quad(w, tr, xa, hwi, va);
            if (sb) clip_quad(w, tr, xb, hwi, vb);
            t[0] |= va.x | vb.x; t[1] |= va.y | vb.y; t[2] |= va.z | vb.z; t[3] |= va.w | vb.w;
        }
        if (!SPARSE) {
            *reinterpret_cast<uint4*>(WIN + fr * kTC + 4 * lane) = make_uint4(t[0], t[1], t[2], t[3]);
        } else {
            // presence bits of my group of 32 columns (8 lanes x 4 columns)
            const int sh = 4 * (lane & 7), grp = lane >> 3;
            const uint32_t nib = (t[0] != 0u ? 1u : 0u) | (t[1] != 0u ? 2u : 0u) | (t[2] != 0u ? 4u : 0u) | (t[3] != 0u ? 8u : 0u);
            uint32_t pr[4];
            #pragma unroll
            for (int g = 0; g < 4; ++g) pr[g] = __reduce_or_sync(kFull, grp == g ? nib << sh : 0u);
            const int c0 = __popc(pr[0]), c1 = __popc(pr[1]), c2 = __popc(pr[2]), c3 = __popc(pr[3]);
            int base = 0;
            if (lane == 0) base = atomicAdd(alloc, c0 + c1 + c2 + c3);
            base = __shfl_sync(kFull, base, 0);
            const uint32_t mypr = grp == 0 ? pr[0] : grp == 1 ? pr[1] : grp == 2 ? pr[2] : pr[3];
            const int gbase = base + (grp > 0 ? c0 : 0) + (grp > 1 ? c1 : 0) + (grp > 2 ? c2 : 0);
            if ((lane & 7) == 0) PB[fr * 4 + grp] = make_uint2(mypr, (uint32_t)gbase);
            int idx = gbase + __popc(mypr & ((1u << sh) - 1u));
            #pragma unroll
            for (int k = 0; k < 4; ++k)
                if (t[k] != 0u) {
                    if (idx < cap_words) WIN[idx] = t[k];
                    ++idx;
                }
        }
    }
}


// The same for the 256-column circle (max_px 512 .. 1023), sparse layout only: one warp per folded row, each lane
// TWO quads of four folded tile columns (4*lane .. and 128 + 4*lane ..), eight presence groups per row.  A folded
// column holds tile column col-128 or col+128 of the upper tile row and col or col-256 of the lower one.
__device__ __noinline__ void stage_window_wide(const Win w, uint32_t* WIN, uint2* PB, int* alloc, const int cap_words,
                                               const int row0, const int row_step) {
    constexpr int TC = TW<8>::TC, HALF = TW<8>::HALF;
    const int lane = threadIdx.x & 31;
    for (int fr = row0; fr < w.NR; fr += row_step) {
        const int tr_u = fr - w.NR + 1, tr_l = fr + 1;
        int hwi_u, hwo_u, hwi_l, hwo_l;
        row_half_widths(w, tr_u, hwi_u, hwo_u);
        row_half_widths(w, tr_l, hwi_l, hwo_l);
        if (tr_l > w.NR - 1) hwo_l = -64;               // no lower tile row: nothing in span
        uint32_t t[2][4] = {{0u, 0u, 0u, 0u}, {0u, 0u, 0u, 0u}};
        #pragma unroll
        for (int part = 0; part < 2; ++part) {
            const int q4 = 4 * lane + part * HALF;      // first folded column of this lane's quad
            #pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int tr = h ? tr_l : tr_u, hwi = h ? hwi_l : hwi_u, hwo = h ? hwo_l : hwo_u;
                int xa, xb;
                const uint4* sa = window_quad(w, tr, h ? q4 : q4 - HALF, hwo, xa);
                const uint4* sb = window_quad(w, tr, h ? q4 - TC : q4 + HALF, hwo, xb);
                uint4 va = sa ? __ldg(sa) : make_uint4(0u, 0u, 0u, 0u);
                uint4 vb = sb ? __ldg(sb) : make_uint4(0u, 0u, 0u, 0u);
                if (sa) clip_quad(w, tr, xa, hwi, va);
                if (sb) clip_quad(w, tr, xb, hwi, vb);
                t[part][0] |= va.x | vb.x; t[part][1] |= va.y | vb.y; t[part][2] |= va.z | vb.z; t[part][3] |= va.w | vb.w;
            }
        }
        // presence bits: group = part * 4 + lane / 8 (8 lanes x 4 columns = 32 columns)
        const int sh = 4 * (lane & 7), grp = lane >> 3;
        uint32_t pr[8];
        int cnt[8];
        #pragma unroll
        for (int part = 0; part < 2; ++part) {
            const uint32_t nib = (t[part][0] != 0u ? 1u : 0u) | (t[part][1] != 0u ? 2u : 0u) | (t[part][2] != 0u ? 4u : 0u) | (t[part][3] != 0u ? 8u : 0u);
            #pragma unroll
            for (int g = 0; g < 4; ++g) {
                pr[part * 4 + g] = __reduce_or_sync(kFull, grp == g ? nib << sh : 0u);
                cnt[part * 4 + g] = __popc(pr[part * 4 + g]);
            }
        }
        int total = 0;
        #pragma unroll
        for (int g = 0; g < 8; ++g) total += cnt[g];
        int base = 0;
        if (lane == 0) base = atomicAdd(alloc, total);
        base = __shfl_sync(kFull, base, 0);
        #pragma unroll
        for (int part = 0; part < 2; ++part) {
            const int mg = part * 4 + grp;
            int gbase = base;
            uint32_t mypr = 0u;
            #pragma unroll
            for (int g = 0; g < 8; ++g) { gbase += g < mg ? cnt[g] : 0; mypr = g == mg ? pr[g] : mypr; }
            if ((lane & 7) == 0) PB[fr * 8 + mg] = make_uint2(mypr, (uint32_t)gbase);
            int idx = gbase + __popc(mypr & ((1u << sh) - 1u));
            #pragma unroll
            for (int k = 0; k < 4; ++k)
                if (t[part][k] != 0u) {
                    if (idx < cap_words) WIN[idx] = t[part][k];
                    ++idx;
                }
        }
    }
}

}  // namespace

}  // namespace axb
