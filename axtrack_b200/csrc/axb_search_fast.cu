// K2 fast kernels (v8) -- level-0 bounded multi-target search with word-parallel wavefronts (sm_100a).
//
// Contract (axb_search.cu header, DESIGN.md "K2 semantics"): one search per SOURCE detection
// answers every in-range target of the source's chain of frame pairs; replaces
// pyastar2d/src/cpp/astar.cpp:46-172 driven by axtrack/AxonDetections.py:553-579,600-629 for every
// source whose mask component reaches hop max_px or contains all its in-range targets.  Sources
// level 0 cannot answer are deferred to k2_general_kernel.
//
// Window = the L1 ball of radius RW = max_px around the source as 8x4-pixel TILE WORDS
// (bit = (y&3)*8 + (x&7)), read from the tiled copy of the packed mask that k2_tile_mask_kernel
// builds once per call.  Tile rows of 4 image rows are aligned to (sy & ~3), tile columns of 8
// pixels to (sx & ~31), and the diamond is FOLDED on a circle of 128 tile columns: tile rows
// tr <= 0 live in folded row tr+NR-1 with the source's aligned column at tile column 64, tile
// rows tr >= 1 in folded row tr-1 with it at tile column 0; NR = ceil(RW/4)+1.  The half widths of
// the two tile rows sharing a folded row add up to <= 2*RW - 4*NR + 3 <= 509, so their pixels
// never meet on the 1024 px circle (above RW = 504 they can share a tile WORD, and a row can wrap
// onto its own first tile: stage_window ORs the disjoint pixels of all candidate tiles).
// Only in-diamond mask pixels are staged as open and cells at hop RW are never
// expanded, so every candidate bit of an expansion is a true 4-neighbour: no bound checks.
//   key of a tile word = half << 15 | folded row << 7 | tile column.
//
// A frontier entry is (tile word key, 32 frontier bits); ONE thread expands all its cells with
// five shift/mask candidates (word itself, right, left, below, above) and at most five
// predicated shared-memory atomicAnd claims.  Each warp appends its claims to its own segment
// of the next frontier (ballot ranks, no counter atomics); one __syncthreads per wavefront.
//
// Two instantiations of the same search:
//   k2_sparse_kernel  window = only the NON-ZERO tile words (<= kSparseWords) + a (presence
//                     bits, base index | target flags) pair per 32 tile columns, target hash
//                     in the CTA's global scratch: 21 KB and 3 warps per search -> 10 CTAs per
//                     SM.  A search is ~max_px dependent wavefronts (latency bound), so the
//                     throughput comes from the number of resident searches.  A launch too
//                     small to fill the slots gets larger windows instead (launch_k2_sparse).
//   k2_fast_kernel    dense window (64.5 KB, 3 CTAs per SM) for the sources whose window has
//                     more non-zero tile words than the sparse layout holds (dense masks).
#include "axb_search_tiles.cuh"
#include <cstdio>

namespace axb {

// Tiled copy of the bit-packed mask: one thread per (tile row, packed word) turns four row words
// into the four 8x4 tile words of the same 32 pixels (byte k of the four rows).
__global__ void __launch_bounds__(256) k2_tile_mask_kernel(const uint32_t* __restrict__ bits, int n_masks, int H, int Wp,
                                                            uint32_t* __restrict__ tiles) {
    const int TH = (H + 3) / 4;
    const long long n = (long long)n_masks * TH * Wp;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int iw = (int)(i % Wp);
        const long long rest = i / Wp;
        const int ty = (int)(rest % TH);
        const int m = (int)(rest / TH);
        const uint32_t* src = bits + ((size_t)m * H + 4 * (size_t)ty) * Wp + iw;
        uint32_t r[4];
        #pragma unroll
        for (int k = 0; k < 4; ++k) r[k] = 4 * ty + k < H ? __ldg(src + (size_t)k * Wp) : 0u;
        uint4 t;
        uint32_t* tp = &t.x;
        #pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t lo = __byte_perm(r[0], r[1], k | ((4 + k) << 4));
            const uint32_t hi = __byte_perm(r[2], r[3], k | ((4 + k) << 4));
            tp[k] = __byte_perm(lo, hi, 0x5410);
        }
        *reinterpret_cast<uint4*>(tiles + ((size_t)m * TH + ty) * (4 * (size_t)Wp) + 4 * (size_t)iw) = t;
    }
}

void launch_k2_tiles(const uint32_t* mask_bits, int n_masks, int H, int Wp, uint32_t* tiles, cudaStream_t stream) {
    const long long n = (long long)n_masks * ((H + 3) / 4) * Wp;
    const int grid = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
    k2_tile_mask_kernel<<<grid > 0 ? grid : 1, 256, 0, stream>>>(mask_bits, n_masks, H, Wp, tiles);
}


namespace {

enum ItemResult { kItemDone = 0, kItemToGeneral = 1, kItemToDense = 2 };

// One work item = one source of a chain leader, processed by the whole CTA.  All control flow is
// CTA-uniform.  kItemToGeneral: level 0 cannot answer; kItemToDense (sparse only): the window has
// more non-zero tile words than the sparse layout holds.
template <bool SPARSE, int LOGC = 7, int NW = kNW>
__device__ int run_item_fast(const K2Params& p, const int item, uint32_t* smem, uint32_t* gscr, Ctl& ctl) {
    using C = Cfg<SPARSE, LOGC>;
    using G = TW<LOGC>;
    static_assert(SPARSE || LOGC == 7, "the dense window exists for the 128-column circle only");
    constexpr int NT = 32 * NW;
    constexpr int kSeg = C::kQCap / NW;             // frontier entries per warp per buffer in shared memory
    constexpr int GSEG = kFastGQCap / NW;           // ... and in the CTA's global scratch
    constexpr int HASH = C::kHash;
    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const uint32_t lt = lanemask_lt();
    const int RW = p.max_px;
    const int NR = (RW + 3) / 4 + 1;

    // shared memory: window words | (sparse) presence/base table | two frontier buffers | target hash
    uint32_t* const WIN = smem;
    uint2* const PB = (uint2*)(WIN + (SPARSE ? p.sparse_words : NR * G::TC));
    uint2* const q0 = PB + (SPARSE ? NR * G::GROUPS : 0);  // [2][kNW][kSeg] (bits, key)
    // target hash: probed only for the few words flagged as holding a target, so the sparse
    // kernel keeps it in the CTA's global scratch (L2) and spends the shared memory on residency
    // (unless the launch is too small to fill the SMs: then latency counts, p.hash_smem)
    uint32_t* const hash = SPARSE && !p.hash_smem ? gscr + kCandCap + 4 * (size_t)kFastGQCap : (uint32_t*)(q0 + 2 * C::kQCap);
    uint32_t* const cand = gscr;
    uint2* const gq0 = (uint2*)(gscr + kCandCap);          // [2][kNW][GSEG] global overflow
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
            ch.ostride[nm] = 1;
            ch.ebase[nm] = p.all_prefix[q];
            ch.dst_off[nm] = pd.dst_off;
            ch.ditem[nm] = 0;
            ch.cum[nm + 1] = ch.cum[nm] + max(pd.n_dst, 0);
            ++nm;
            q = p.next_same[q];
        }
        uint32_t rm = 0u;
        if (p.reverse) {               // pairs this detection is a target of and whose sources are not searched
            for (q = p.rev_head[lo]; q >= 0 && nm < kMaxMem; q = p.rev_next[q]) {
                const axb_pair_desc pd = p.pairs[q];
                ch.orow[nm] = pd.out_off + si;
                ch.ostride[nm] = pd.n_dst;
                ch.ebase[nm] = p.all_prefix[q];
                ch.dst_off[nm] = pd.src_off;
                ch.ditem[nm] = p.work_prefix[p.lead_of[q]];
                ch.cum[nm + 1] = ch.cum[nm] + max(pd.n_src, 0);
                rm |= 1u << nm;
                ++nm;
            }
        }
        ch.nm = nm;
        ch.si = si;
        ch.revmask = rm;
        ch.mask_idx = lead.mask_idx;
    }
    __syncthreads();
    const int nm = ch.nm;
    const int total_t = ch.cum[nm];
    const uint32_t revmask = ch.revmask;
    if (total_t <= 0) return kItemDone;

    const int sy = p.dets_yx[2 * ((size_t)lead.src_off + si)];
    const int sx = p.dets_yx[2 * ((size_t)lead.src_off + si) + 1];
    Win w;
    w.tiles = p.mask_tiles + (size_t)lead.mask_idx * p.TH * p.TWp;
    w.TH = p.TH; w.TWp = p.TWp; w.RW = RW; w.NR = NR;
    w.oy = sy & 3; w.ox = sx & 31;                         // only used when the source is in the grid
    w.ty0 = sy >> 2; w.tx0 = (sx >> 5) * 4;
    const bool src_in = sy >= 0 && sy < p.H && sx >= 0 && sx < p.W;
    const bool src_on = src_in && ((__ldg(p.mask_bits + ((size_t)lead.mask_idx * p.H + sy) * p.Wp + (sx >> 5)) >> (sx & 31)) & 1u);
    const bool pre = p.early && src_in && !src_on;          // the all-levels kernel has this item's forward members already (k2_classify_kernel)
    if (pre && !revmask) return kItemDone;
    const long long r2 = (long long)RW * RW;
    constexpr int kPassLoad = C::kMaxLoad > NT ? C::kMaxLoad - NT : 0;

    int tt = 0;                        // next target of the flattened chain target list
    bool to_gen = false;
    for (;;) {                         // passes of <= kMaxLoad candidates (one on real data)
        for (int i = tid; i < HASH; i += NT) hash[i] = kEmpty;
        if (tid == 0) { ctl.ncand = 0; ctl.alloc = 0; ctl.nchk = 0; }
        for (;;) {                     // groups of NT targets
            __syncthreads();
            if (tt >= total_t || ctl.ncand > kPassLoad) break;
            const int t = tt + tid;
            bool is_cand = false, self = false, chk = false;
            int m = 0, j = 0, dy = 0, dx = 0;
            if (t < total_t) {
                while (m + 1 < nm && t >= ch.cum[m + 1]) ++m;
                j = t - ch.cum[m];
            }
            const bool rev = (revmask >> m) & 1u;
            if (t < total_t && (rev || !pre)) {
                const int ty = p.dets_yx[2 * ((size_t)ch.dst_off[m] + j)];
                const int tx = p.dets_yx[2 * ((size_t)ch.dst_off[m] + j) + 1];
                const bool t_in = ty >= 0 && ty < p.H && tx >= 0 && tx < p.W;
                // reverse member whose other detection is off the mask: its whole row belongs to the all-levels kernel
                const bool foreign = rev && p.early && t_in &&
                                     !((__ldg(p.mask_bits + ((size_t)ch.mask_idx * p.H + ty) * p.Wp + (tx >> 5)) >> (tx & 31)) & 1u);
                dy = ty - sy; dx = tx - sx;
                const long long d2 = (long long)dy * dy + (long long)dx * dx;
                if (d2 < r2 && !foreign) {                            // AxonDetections.py:620-623
                    // the reference's start is this detection (forward) or the other one (reverse member)
                    const bool start_in = rev ? t_in : src_in, goal_in = rev ? src_in : t_in;
                    const unsigned long long ek = rev ? (((unsigned long long)(p.orig_base + ch.ebase[m] + j) << 26) | ((unsigned long long)si << 2))
                                                      : (((unsigned long long)(p.orig_base + ch.ebase[m] + si) << 26) | ((unsigned long long)j << 2));
                    if (!start_in) atomicMin(p.err_key, ek | 1ull);    // astar_wrapper.py:39-42
                    else if (!goal_in) atomicMin(p.err_key, ek | 2ull);  // astar_wrapper.py:43-45
                    else if (dy == 0 && dx == 0) self = true;         // start == goal: L = 1 (astar.cpp:67,99)
                    else is_cand = abs(dy) + abs(dx) < RW;            // farther cells need > max_px pixels
                    // reverse member: "not reached from here" is the reference's None only if the other detection's own
                    // level 0 is known to reach hop max_px (axb_components.cu); else it gets its own search afterwards
                    if (rev && is_cand) {
                        const int reach = component_reaches(p, ch.mask_idx, ty, tx, RW);
                        chk = reach <= 0;
                        // nothing to settle from here when the other detection is off the mask (it is no open cell of
                        // this window) or this one is (a path INTO an off-mask cell is not a level-0 path)
                        if (reach < 0 || !src_on) {
                            is_cand = false;
                            if (chk) push_directed(p, ch.ditem[m] + j, reach < 0);
                            chk = false;
                        }
                    }
                }
                if (!foreign && (!rev || self)) {                     // reverse members: None was filled in row-wise (k2_fill_reverse_kernel)
                    const long long oo = ch.orow[m] + (long long)j * ch.ostride[m];
                    AXB_ASSERT(oo >= 0 && m < nm && j >= 0 && j < ch.cum[m + 1] - ch.cum[m]);
                    p.path_len[oo] = self ? 1 : RW;                   // None -> max_px (AxonDetections.py:736)
                    p.unreachable[oo] = self ? 0 : 1;
                    if (p.wcost) p.wcost[oo] = self ? 0.f : __int_as_float(0x7f800000);
                }
            }
            if (revmask) {
                const uint32_t cb = __ballot_sync(kFull, chk);
                if (cb && lane == 0) atomicAdd(&ctl.nchk, __popc(cb));
            }
            const uint32_t bal = __ballot_sync(kFull, is_cand);
            if (bal) {
                int base = 0;
                if (lane == 0) base = atomicAdd(&ctl.ncand, __popc(bal));
                base = __shfl_sync(kFull, base, 0);
                if (is_cand) {
                    const uint32_t slot = (uint32_t)base + __popc(bal & lt);
                    AXB_ASSERT(slot < (uint32_t)kCandCap && slot < 2048u);
                    cand[slot] = ((uint32_t)m << 29) | (chk ? 1u << 28 : 0u) | (uint32_t)j;
                    int bit;
                    const uint32_t key = cell_key<LOGC>(dy, dx, w.oy, w.ox, NR, bit);
                    AXB_ASSERT(((key & G::WMASK) >> LOGC) < (uint32_t)NR && bit >= 0 && bit < 32 && slot < (1u << G::SLOTB));
                    uint32_t entry = (key << G::ESH) | slot;
                    if (G::BIT_IN_ENTRY) entry |= (uint32_t)bit << G::SLOTB;
                    else reinterpret_cast<uint8_t*>(gscr + kCandCap + 4 * (size_t)kFastGQCap + kSparseHash)[slot] = (uint8_t)bit;
                    uint32_t h = hash_slot<HASH>(key);
                    while (atomicCAS(&hash[h], kEmpty, entry) != kEmpty) h = (h + 1) & (HASH - 1);
                }
            }
            tt += NT;
        }
        const int ncand = ctl.ncand;
        if (ncand > 0) {
            AXB_T(0, tclk);
            if (LOGC == 7) stage_window<SPARSE>(w, WIN, PB, &ctl.alloc, p.sparse_words, threadIdx.x >> 5, NW);
            else stage_window_wide(w, WIN, PB, &ctl.alloc, p.sparse_words, threadIdx.x >> 5, NW);
            __syncthreads();
            AXB_T(1, tclk);
            if (SPARSE && ctl.alloc > p.sparse_words) return kItemToDense;
            if (SPARSE) {              // flag the tile-word pairs that hold a target: bits 16.. of PB[].y
                for (int i = tid; i < HASH; i += NT) {
                    const uint32_t e = hash[i];
                    if (e != kEmpty) {
                        const uint32_t key = e >> G::ESH;
                        atomicOr(&PB[(key & G::WMASK) >> 5].y, 1u << (16u + ((key & 31u) >> 1)));
                    }
                }
            }
            if (tid == 0) {            // close the source cell, seed the search with it
                int bit;
                const uint32_t key = cell_key<LOGC>(0, 0, w.oy, w.ox, NR, bit);
                int widx = (int)(key & G::WMASK);
                bool there = true;
                if (SPARSE) {
                    const uint2 pb = PB[widx >> 5];
                    there = (pb.x >> (widx & 31)) & 1u;
                    widx = (int)(pb.y & 0xFFFFu) + __popc(pb.x & ((1u << (widx & 31)) - 1u));
                }
                AXB_ASSERT(!there || widx < (SPARSE ? p.sparse_words : NR * G::TC));
                if (there) WIN[widx] &= ~(1u << bit);
                q0[0] = make_uint2(1u << bit, key);
                #pragma unroll
                for (int k = 0; k < 8; ++k) { ctl.wc[0][k] = k == 0 ? 1 : 0; ctl.wc[1][k] = 0; }   // words of absent warps stay 0
            }
            __syncthreads();

            // ---- wavefronts: one thread per frontier entry (tile word key, frontier bits) ---------
            const uint32_t win_s = smem_u32(WIN);
            const uint32_t pb_s = smem_u32(PB);
            const uint32_t hash_s = SPARSE ? 0u : smem_u32(hash);
            const uint32_t q_s = smem_u32(q0);
            const uint32_t lastU = (uint32_t)(NR - 1);
            const uint32_t idle_s = win_s + 4u * (uint32_t)lane;   // where a lane's claims go when it has nothing to claim (one bank per lane)
            // shared address of tile word `key` (only when want != 0); want is cleared when the
            // sparse window holds no such word
            // (sparse: PB[].y = index of the group's first word | target flags << 16, returned in aux)
            auto resolve = [&](uint32_t key, uint32_t& want, uint32_t& aux) -> uint32_t {
                const uint32_t widx = key & G::WMASK;
                if (!SPARSE) return win_s + 4u * widx;
                AXB_ASSERT(!want || (widx >> LOGC) < (uint32_t)NR);
                const uint2 pb = lds64_if(pb_s + 8u * (widx >> 5), want);
                const uint32_t bit = 1u << (widx & 31u);
                if (!(pb.x & bit)) want = 0u;
                aux = pb.y;
                AXB_ASSERT(!want || (pb.y & 0xFFFFu) + (uint32_t)__popc(pb.x & (bit - 1u)) < (uint32_t)p.sparse_words);
                return win_s + 4u * ((pb.y & 0xFFFFu) + (uint32_t)__popc(pb.x & (bit - 1u)));
            };
            int unsettled = ncand;     // CTA-uniform: updated from ctl.wc[] after each barrier
            bool hit = false, defer = false;
            for (int j = 0;; ++j) {    // wave j: probe + expand the cells at hop j
                constexpr int NWC = NW > 4 ? 8 : 4;                    // control words read: one or two 16-byte rows
                int wcv[NWC];
                *reinterpret_cast<int4*>(wcv) = *reinterpret_cast<const int4*>(ctl.wc[j & 1]);
                if (NW > 4) *reinterpret_cast<int4*>(wcv + 4) = *reinterpret_cast<const int4*>(ctl.wc[j & 1] + 4);
                int pfx[NWC + 1];                                      // pfx[k] = first entry of warp k's segment in the wave's numbering
                pfx[0] = 0;
                int any_ovf = 0;
                #pragma unroll
                for (int k = 0; k < NWC; ++k) {
                    pfx[k + 1] = pfx[k] + (wcv[k] & 0x7FFF);
                    unsettled -= wcv[k] >> 16;
                    any_ovf |= wcv[k];
                }
                const int total = pfx[NWC];
                if (any_ovf & 0x8000) { defer = true; break; }         // a frontier list overflowed
                if (total == 0 || unsettled <= 0) break;
                if (j >= RW) { hit = true; break; }                   // cells at hop == max_px exist
                AXB_C(3, 1); AXB_C(5, total);
#ifdef AXB_TIMING
                long long tw = clock64();
#endif
                const uint32_t qcur_s = q_s + (uint32_t)(j & 1) * (C::kQCap * 8u);
                const uint32_t qnxt_s = q_s + (uint32_t)((j + 1) & 1) * (C::kQCap * 8u) + (uint32_t)warp * (kSeg * 8u);
                int wn = 0, ws = 0;
                bool wovf = false;
                for (int base = warp * 32; base < total; base += NT) {
                    const int i = base + lane;
                    int sg = 0, sbase = 0;                                    // whose segment entry i is in, without branches
                    #pragma unroll
                    for (int k = 1; k < NW; ++k) {
                        const bool ge = i >= pfx[k];
                        sg += (int)ge;
                        sbase = ge ? pfx[k] : sbase;
                    }
                    const int off = i - sbase;
                    const bool act = i < total;
                    AXB_ASSERT(!act || (sg < NW && off >= 0 && off < kSeg + GSEG));
                    uint2 e = lds64_if(qcur_s + 8u * (uint32_t)(sg * kSeg + off), act && off < kSeg);
                    {                                                         // entries that spilled to global memory
                        const uint2 eg = ldg64_if(gq0 + (size_t)(j & 1) * kFastGQCap + (size_t)sg * GSEG + (off - kSeg), act && off >= kSeg);
                        e.x |= eg.x; e.y |= eg.y;
                    }
                    const uint32_t f = e.x, k0 = e.y;
                    // candidates: cells of the word itself, of the right / left / lower / upper word
                    uint32_t cd0 = ((f << 1) & 0xFEFEFEFEu) | ((f >> 1) & 0x7F7F7F7Fu) | (f << 8) | (f >> 8);
                    uint32_t cd1 = (f & 0x80808080u) >> 7;
                    uint32_t cd2 = (f & 0x01010101u) << 7;
                    uint32_t cd3 = f >> 24, cd4 = f << 24;
                    constexpr uint32_t CM = (uint32_t)G::TC - 1u;
                    const uint32_t rowk = k0 & ~CM;
                    const uint32_t k1 = rowk | ((k0 + 1u) & CM);
                    const uint32_t k2 = rowk | ((k0 - 1u) & CM);
                    uint32_t k3 = k0 + (uint32_t)G::TC, k4 = k0 - (uint32_t)G::TC;
                    const uint32_t hr = k0 >> LOGC;                   // half << FRB | folded row
                    if (hr == lastU) k3 = G::HALFBIT | ((k0 + (uint32_t)G::HALF) & CM);           // tile row 0 -> 1
                    if (hr == (1u << G::FRB)) k4 = (lastU << LOGC) | ((k0 + (uint32_t)G::HALF) & CM);   // tile row 1 -> 0
                    if (hr == (1u << G::FRB) - 1u + lastU) cd3 = 0u;  // window borders (never open anyway)
                    if (hr == 0u) cd4 = 0u;
                    uint32_t aux0 = 0u, auxn = 0u;
                    const uint32_t a0 = resolve(k0, cd0, aux0), a1 = resolve(k1, cd1, auxn), a2 = resolve(k2, cd2, auxn);
                    const uint32_t a3 = resolve(k3, cd3, auxn), a4 = resolve(k4, cd4, auxn);
                    // may the word hold a target?  sparse: flag of its column pair; dense: its hash slot is in use
                    bool maybe;
                    AXB_ASSERT(!act || f != 0u);
                    if (SPARSE) maybe = f != 0u && ((aux0 >> (16u + ((k0 & 31u) >> 1))) & 1u);
                    else maybe = f != 0u && lds_if(hash_s + 4u * hash_slot<HASH>(k0), f) != kEmpty;
                    const uint32_t o0 = lds_if(a0, cd0) & cd0;
                    const uint32_t o1 = lds_if(a1, cd1) & cd1;
                    const uint32_t o2 = lds_if(a2, cd2) & cd2;
                    const uint32_t o3 = lds_if(a3, cd3) & cd3;
                    const uint32_t o4 = lds_if(a4, cd4) & cd4;
                    const uint32_t c0 = claim_if(a0, o0, idle_s);
                    const uint32_t c1 = claim_if(a1, o1, idle_s);
                    const uint32_t c2 = claim_if(a2, o2, idle_s);
                    const uint32_t c3 = claim_if(a3, o3, idle_s);
                    const uint32_t c4c = claim_if(a4, o4, idle_s);
                    // targets among the cells being expanded (hop j, path length j + 1)
                    if (__any_sync(kFull, maybe)) {
                        int ns = 0;
                        if (maybe) ns = probe_settle<HASH, LOGC>(p, hash, cand, ch, k0, f, j);
                        ws += __reduce_add_sync(kFull, ns);
                    }
                    // append the claims to this warp's segment of the next frontier, lane-major
                    const uint32_t b0 = __ballot_sync(kFull, c0 != 0u), b1 = __ballot_sync(kFull, c1 != 0u);
                    const uint32_t b2 = __ballot_sync(kFull, c2 != 0u), b3 = __ballot_sync(kFull, c3 != 0u);
                    const uint32_t b4 = __ballot_sync(kFull, c4c != 0u);
                    const int tot = __popc(b0) + __popc(b1) + __popc(b2) + __popc(b3) + __popc(b4);
                    int slot = wn + __popc(b0 & lt) + __popc(b1 & lt) + __popc(b2 & lt) + __popc(b3 & lt) + __popc(b4 & lt);
                    AXB_ASSERT(slot >= 0 && (wn + tot > kSeg + GSEG || slot + (c0 != 0u) + (c1 != 0u) + (c2 != 0u) + (c3 != 0u) + (c4c != 0u) <= wn + tot));
                    const uint32_t fits = wn + tot <= kSeg ? 1u : 0u;     // warp-uniform; the stores are predicated, not branched over
                    {
                        uint32_t sa = qnxt_s + 8u * (uint32_t)slot;
                        sts64_if(sa, c0, k0, fits ? c0 : 0u);   sa += c0 ? 8u : 0u;
                        sts64_if(sa, c1, k1, fits ? c1 : 0u);   sa += c1 ? 8u : 0u;
                        sts64_if(sa, c2, k2, fits ? c2 : 0u);   sa += c2 ? 8u : 0u;
                        sts64_if(sa, c3, k3, fits ? c3 : 0u);   sa += c3 ? 8u : 0u;
                        sts64_if(sa, c4c, k4, fits ? c4c : 0u);
                    }
                    if (!fits) {                                              // rare: the tail of the segment lives in global memory
                        if (wn + tot <= kSeg + GSEG) {
                            const uint32_t cc[5] = {c0, c1, c2, c3, c4c};
                            const uint32_t kk[5] = {k0, k1, k2, k3, k4};
                            uint2* const gnxt = gq0 + (size_t)((j + 1) & 1) * kFastGQCap + (size_t)warp * GSEG;
                            #pragma unroll
                            for (int d = 0; d < 5; ++d) {
                                if (cc[d]) {
                                    if (slot < kSeg) sts64_if(qnxt_s + 8u * (uint32_t)slot, cc[d], kk[d], 1u);
                                    else gnxt[slot - kSeg] = make_uint2(cc[d], kk[d]);
                                    ++slot;
                                }
                            }
                        } else {
                            wovf = true;
                        }
                    }
                    wn += tot;
                }
                if (lane == 0) ctl.wc[(j + 1) & 1][warp] = wovf ? 0x8000 : (wn | (ws << 16));
                AXB_T(6, tw);
                __syncthreads();
#ifdef AXB_TIMING
                if (ctl.wc[(j + 1) & 1][0] != -12345) AXB_T(7, tw);
#endif
            }
            AXB_T(2, tclk); AXB_C(4, 1);
            if (revmask && (defer || ctl.nchk > 0)) {
                // reverse members: what this search did not reach stays None only where that is certain (nothing is
                // certain after a frontier overflow); the others get their own directed search in a later launch
                for (int i = tid; i < ncand; i += NT) {
                    const uint32_t c = cand[i];
                    const int m = (int)(c >> 29);
                    if (!((revmask >> m) & 1u) || !(defer || ((c >> 28) & 1u))) continue;
                    const int jj = (int)(c & 0x0FFFFFFFu);
                    const long long oo = ch.orow[m] + (long long)jj * ch.ostride[m];
                    if (*(volatile const uint8_t*)(p.unreachable + oo)) {
                        push_directed(p, ch.ditem[m] + jj, false);
                    }
                }
            }
            if (defer || (!hit && unsettled > 0)) {                   // level 0 cannot answer the forward members
                if (!revmask) return kItemToGeneral;
                to_gen = true;                                        // the reverse members of later passes are still due
            }
        }
        if (tt >= total_t) break;
    }
    return to_gen ? kItemToGeneral : kItemDone;
}

}  // namespace

// Next work item of this launch (-1: none left).  First pass: tickets over the sources of the SEARCHED sets
// (p1_prefix); later passes: a list a previous launch wrote.
__device__ __forceinline__ int fetch_item(const K2Params& p, Ctl& ctl) {
    __syncthreads();
    if (threadIdx.x == 0) {
        const int t = atomicAdd(&p.counters[p.src_ticket_ctr], 1);
        int item = -1;
        if (p.src_list) {
            if (t < p.counters[p.src_len_ctr]) item = p.src_list[t];
        } else if (t < p.p1_prefix[p.n_pairs]) {
            int lo = 0, hi = p.n_pairs - 1;
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (p.p1_prefix[mid] <= t) lo = mid; else hi = mid - 1;
            }
            item = p.work_prefix[lo] + (t - p.p1_prefix[lo]);
            AXB_ASSERT(t - p.p1_prefix[lo] < p.pairs[lo].n_src && item < p.work_prefix[p.n_pairs]);
        }
        ctl.item = item;
    }
    __syncthreads();
    return ctl.item;
}

__device__ __forceinline__ void defer_item(const K2Params& p, int r, int item) {
    if (threadIdx.x != 0) return;
    if (r == kItemToGeneral) p.gen_list[atomicAdd(&p.counters[kCtrGenLen], 1)] = item;
    else if (r == kItemToDense) p.dense_out[atomicAdd(&p.counters[p.dense_out_ctr], 1)] = item;
}

// Persistent CTAs over a ticket counter.
__global__ void __launch_bounds__(kFastThreads, kSparseCtasPerSm) k2_sparse_kernel(const K2Params p) {
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ Ctl ctl;
    uint32_t* const gscr = p.fast_scratch + (size_t)blockIdx.x * kFastScratchWords;
    for (;;) {
        const int item = fetch_item(p, ctl);
        if (item < 0) break;
        defer_item(p, run_item_fast<true>(p, item, smem, gscr, ctl), item);
    }
}

// Dense windows for the items the sparse layout could not hold.
__global__ void __launch_bounds__(kFastThreads, kFastCtasPerSm) k2_fast_kernel(const K2Params p) {
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ Ctl ctl;
    uint32_t* const gscr = p.fast_scratch + (size_t)blockIdx.x * kFastScratchWords;
    for (;;) {
        const int item = fetch_item(p, ctl);
        if (item < 0) break;
        defer_item(p, run_item_fast<false>(p, item, smem, gscr, ctl), item);
    }
}

// max_px 512 .. 1023: the same sparse-window search on the 256-column circle (257 folded rows, 18-bit word keys).
// Windows with more non-zero tile words than the layout holds, and everything level 0 cannot answer, go to the
// all-levels kernel (k2_general_wide_kernel).
// First kind (SECOND = false): windows of up to kWideWords non-zero tile words, 2 CTAs per SM; items whose window holds
// more go to a dense list.  Second kind (SECOND = true): those items with windows of as many words as one CTA per SM
// can hold.  What neither holds (e.g. an all-ones mask: 65 792 words) goes to the all-levels kernel.
template <bool SECOND>
__global__ void __launch_bounds__(32 * kWideNW, SECOND ? 1 : kWideSparseCtasPerSm) k2_sparse_wide_kernel(const K2Params p) {
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ Ctl ctl;
    uint32_t* const gscr = p.fast_scratch + (size_t)blockIdx.x * kFastScratchWords;
    for (;;) {
        const int item = fetch_item(p, ctl);
        if (item < 0) break;
        int r = run_item_fast<true, 8, kWideNW>(p, item, smem, gscr, ctl);
        if (SECOND && r == kItemToDense) r = kItemToGeneral;
        defer_item(p, r, item);
    }
}

void launch_k2_sparse_wide(const K2Params& p, int grid, bool second, cudaStream_t stream) {
    static unsigned long long attr_set = 0ull;
    const size_t fixed = k2_sparse_wide_smem_bytes(p.max_px) - (size_t)kWideWords * sizeof(uint32_t);
    if (first_use_on_device(attr_set)) {
        cudaFuncSetAttribute(k2_sparse_wide_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(k2_sparse_wide_smem_bytes(AXB_MAX_PX) + kWideHash * sizeof(uint32_t)));
        cudaFuncSetAttribute(k2_sparse_wide_kernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        cudaFuncSetAttribute(k2_sparse_wide_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
        cudaFuncSetAttribute(k2_sparse_wide_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    }
    K2Params q = p;
    q.hash_smem = 1;               // two CTAs per SM leave room for the target hash in shared memory (the probe is on the wave's critical path)
    const size_t hash_bytes = kWideHash * sizeof(uint32_t);
    if (!second) {
        q.sparse_words = kWideWords;
        k2_sparse_wide_kernel<false><<<grid, 32 * kWideNW, k2_sparse_wide_smem_bytes(p.max_px) + hash_bytes, stream>>>(q);
    } else {                       // as many words as one CTA per SM can hold next to the tables, queues and hash
        q.sparse_words = (int)(((size_t)226 * 1024 - fixed - hash_bytes) / sizeof(uint32_t)) & ~3;
        k2_sparse_wide_kernel<true><<<grid, 32 * kWideNW, fixed + hash_bytes + (size_t)q.sparse_words * sizeof(uint32_t), stream>>>(q);
    }
}

void launch_k2_sparse(const K2Params& p, int grid, cudaStream_t stream) {
    static unsigned long long attr_set = 0ull;
    if (first_use_on_device(attr_set)) {
        cudaFuncSetAttribute(k2_sparse_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
        cudaFuncSetAttribute(k2_sparse_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    }
    // A launch that does not fill the CTA slots has shared memory to spare: the window capacity
    // grows to what the launch's residency leaves (up to the whole folded diamond, so that fewer or
    // no windows need the dense kernel's second pass), and when at least half of the slots stay
    // empty (latency counts, not residency) the target hash moves from the CTA's global scratch
    // into shared memory as well.
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    K2Params q = p;
    q.hash_smem = 0;
    q.sparse_words = kSparseWords;
    size_t smem = k2_sparse_smem_bytes(p.max_px);
    const int per_sm = (grid + sms - 1) / sms;
    if (per_sm < kSparseCtasPerSm) {
        q.hash_smem = 2 * per_sm <= kSparseCtasPerSm ? 1 : 0;
        const size_t fixed = smem - (size_t)kSparseWords * sizeof(uint32_t) + (q.hash_smem ? kSparseHash * sizeof(uint32_t) : 0);
        const size_t budget = (size_t)227 * 1024 / per_sm - 1024;
        const size_t full = (size_t)k2_tile_rows(p.max_px) * 128;
        size_t words = (budget - fixed) / sizeof(uint32_t) & ~(size_t)3;
        if (words > full) words = full;
        if (words > (size_t)kSparseWords) {
            q.sparse_words = (int)words;
            smem = fixed + words * sizeof(uint32_t);
        } else {
            q.hash_smem = 0;
        }
    }
    static const bool dbg = getenv("AXB_K2_DEBUG") != nullptr;
    if (dbg) {
        int nb = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k2_sparse_kernel, kFastThreads, smem);
        fprintf(stderr, "[axb] k2_sparse_kernel: grid %d, %zu B dynamic shared memory, %d window words, %d CTAs per SM\n", grid, smem, q.sparse_words, nb);
    }
    k2_sparse_kernel<<<grid, kFastThreads, smem, stream>>>(q);
}

void launch_k2_fast(const K2Params& p, int grid, cudaStream_t stream) {
    static unsigned long long attr_set = 0ull;
    if (first_use_on_device(attr_set)) {
        cudaFuncSetAttribute(k2_fast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)k2_fast_smem_bytes(kFastMaxPx));
        cudaFuncSetAttribute(k2_fast_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    }
    k2_fast_kernel<<<grid, kFastThreads, k2_fast_smem_bytes(p.max_px), stream>>>(p);
}

#ifdef AXB_TIMING
extern "C" void axb_dbg_fetch(unsigned long long* out, int reset) {
    cudaMemcpyFromSymbol(out, g_dbg, sizeof(g_dbg));
    if (reset) { unsigned long long z[24] = {0}; cudaMemcpyToSymbol(g_dbg, z, sizeof(z)); }
}
#endif

}  // namespace axb
