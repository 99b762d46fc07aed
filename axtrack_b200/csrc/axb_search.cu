// K2 -- batched, radius-bounded, multi-target grid shortest-path search (sm_100a).
//
// Replaces the per-pair A* of the reference (pyastar2d/src/cpp/astar.cpp:46-172 driven by
// axtrack/AxonDetections.py:553-579,600-629): ONE bounded search per SOURCE detection answers
// every target of that source.
//
// Semantics (DESIGN.md "K2 semantics"; validated against the compiled reference in
// tests/test_oracle_golden.py).  Weights are 1 on the mask and 65536 off it
// (AxonDetections.py:598), so the exact cost of a path is n_fg + 65536*n_bg over the pixels it
// enters and its hop count is n_fg + n_bg.  Cells are settled in lexicographic (n_bg, n_fg)
// order: "level" k = n_bg, and inside a level a breadth-first wave over mask pixels indexed by
// j = n_fg.  Off-mask neighbours of level-k cells seed level k+1 with the j of the cell that
// touched them first.  A cell at hop == max_px is never expanded (the reference pushes it with
// path_length max_px+1, astar.cpp:136, and aborts when it reaches the top of the heap,
// astar.cpp:88-92).  The search stops after the first level k* that contains such a cell:
// every target of a higher level is "None" because that heap entry has a smaller priority than
// any goal of a higher level (65536 >> 2*max_px).  Targets settled at (k, j), k <= k*, get
// path_length = k + j + 1 if that is <= max_px.
//
// Mapping to the SM: one CTA per source, persistent over a ticket counter.  The source's
// window (L1 ball of radius max_px) is staged from the bit-packed frame mask in L2/HBM into
// two shared-memory bitmaps in the folded-diamond layout of axb_common.cuh -- "open mask
// pixel" and "open off-mask pixel"; claiming a cell is one shared-memory atomicAnd.  The
// frontier lives in shared memory (global overflow for pathological masks); each wavefront is
// expanded by four lanes per frontier cell and compacted with warp ballots.
#include "axb_common.cuh"

namespace axb {

namespace {

__device__ __forceinline__ uint32_t span_mask(int a, int b) {
    // bits i in [a, b] clipped to [0, 31]
    a = a < 0 ? 0 : a;
    b = b > 31 ? 31 : b;
    if (a > b) return 0u;
    return (0xFFFFFFFFu >> (31 - b)) & (0xFFFFFFFFu << a);
}

// 32 mask bits of image row `row` starting at pixel x0 (any alignment, may start left of the
// image or run past it: missing words read as 0).
__device__ __forceinline__ uint32_t load_bits(const uint32_t* __restrict__ row, int Wp, int x0) {
    const int wi = x0 >> 5;            // arithmetic shift: floor
    const int sh = x0 & 31;
    const uint32_t lo = (wi >= 0 && wi < Wp) ? __ldg(row + wi) : 0u;
    const uint32_t hi = (sh != 0 && wi + 1 >= 0 && wi + 1 < Wp) ? __ldg(row + wi + 1) : 0u;
    return __funnelshift_r(lo, hi, sh);
}

__device__ __forceinline__ int fold_index(int dy, int dx, int RW) {
    const int row = dy <= 0 ? dy + RW : dy - 1;
    const int col = (dx + (dy <= 0 ? 512 : 0)) & (kRowBits - 1);
    return row * kRowBits + col;
}

__device__ __forceinline__ uint32_t pack_cell(int dy, int dx) {
    return ((uint32_t)(dy + 512) << 10) | (uint32_t)(dx + 512);
}

}  // namespace

template <int NT>
__global__ void __launch_bounds__(NT, 1) k2_search_kernel(const K2Params p) {
    extern __shared__ __align__(16) uint32_t smem[];
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int RW = p.max_px;
    const int nrows = RW + 1;
    const int nwords = nrows * kRowWords;

    uint32_t* const OF = smem;                 // open mask pixels      [nrows][32]
    uint32_t* const OB = OF + nwords;          // open off-mask pixels  [nrows][32]
    uint32_t* const q0 = OB + nwords;          // frontier buffers      [2][kQCap]
    uint32_t* const send0 = q0 + 2 * kQCap;    // seed calendars        [2][kSeedEnd]
    uint32_t* const tcell = send0 + 2 * kSeedEnd;   // in-range targets: bit31 = on mask, low = fold index
    uint32_t* const tidx = tcell + kTCap;           // target index j
    uint32_t* const tres = tidx + kTCap;            // (k << 16 | j) once settled

    __shared__ uint32_t s_cnt[3];
    __shared__ uint32_t s_seedcnt;
    __shared__ uint32_t s_ntgt;
    __shared__ int s_item;

    uint32_t* const gq0 = p.scratch + (size_t)blockIdx.x * 4 * kGCap;   // frontier overflow [2][kGCap]
    uint32_t* const gs0 = gq0 + 2 * kGCap;                              // seeds            [2][kGCap]

    const long long r2 = (long long)p.max_px * p.max_px;

    for (;;) {
        __syncthreads();
        if (tid == 0) s_item = atomicAdd(p.work_counter, 1);
        __syncthreads();
        const int item = s_item;
        if (item >= p.work_prefix[p.n_pairs]) break;

        // ---- which pair / source ------------------------------------------------------
        // work_prefix = exclusive prefix of n_src over the pairs (launch_work_prefix)
        int lo = 0, hi = p.n_pairs - 1;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (p.work_prefix[mid] <= item) lo = mid; else hi = mid - 1;
        }
        const axb_pair_desc pd = p.pairs[lo];
        const int si = item - p.work_prefix[lo];
        const int n_dst = pd.n_dst;
        if (n_dst <= 0) continue;
        const int sy = p.dets_yx[2 * (pd.src_off + si)];
        const int sx = p.dets_yx[2 * (pd.src_off + si) + 1];
        const bool src_in = sy >= 0 && sy < p.H && sx >= 0 && sx < p.W;
        const int32_t* __restrict__ dst = p.dets_yx + 2 * (size_t)pd.dst_off;
        const uint32_t* __restrict__ mbits = p.mask_bits + (size_t)pd.mask_idx * p.H * p.Wp;
        const long long orow = pd.out_off + (long long)si * n_dst;

        for (int tb = 0; tb < n_dst; tb += kTCap) {
            const int tend = min(n_dst, tb + kTCap);
            __syncthreads();
            if (tid == 0) s_ntgt = 0;
            __syncthreads();

            // ---- defaults + in-range target list -----------------------------------------
            for (int j0 = tb; j0 < tend; j0 += NT) {
                const int j = j0 + tid;
                bool listed = false;
                int dy = 0, dx = 0;
                if (j < tend) {
                    p.path_len[orow + j] = p.max_px;
                    p.unreachable[orow + j] = 1;
                    if (p.wcost) p.wcost[orow + j] = __int_as_float(0x7f800000);
                    const int ty = dst[2 * j], tx = dst[2 * j + 1];
                    dy = ty - sy; dx = tx - sx;
                    const long long d2 = (long long)dy * dy + (long long)dx * dx;
                    if (d2 < r2) {                                   // AxonDetections.py:620-623
                        const bool t_in = ty >= 0 && ty < p.H && tx >= 0 && tx < p.W;
                        if (!src_in)
                            atomicMin(p.err_key, ((unsigned long long)item << 26) | ((unsigned long long)j << 2) | 1ull);
                        else if (!t_in)
                            atomicMin(p.err_key, ((unsigned long long)item << 26) | ((unsigned long long)j << 2) | 2ull);
                        else
                            listed = (abs(dy) + abs(dx) < RW);       // farther cells cannot be reached in < max_px hops
                    }
                }
                const uint32_t bal = __ballot_sync(0xFFFFFFFFu, listed);
                if (bal) {
                    uint32_t base = 0;
                    if (lane == 0) base = atomicAdd(&s_ntgt, __popc(bal));
                    base = __shfl_sync(0xFFFFFFFFu, base, 0);
                    if (listed) {
                        const uint32_t slot = base + __popc(bal & ((1u << lane) - 1));
                        tcell[slot] = (uint32_t)fold_index(dy, dx, RW);
                        tidx[slot] = (uint32_t)j;
                        tres[slot] = (dy == 0 && dx == 0) ? 0u : kUnset;   // target == source: (k, j) = (0, 0)
                    }
                }
            }
            __syncthreads();
            const int ntgt = (int)s_ntgt;
            if (ntgt == 0) continue;      // uniform: nothing in range (or input error) -> all None

            // ---- stage the window ---------------------------------------------------------
            for (int i = tid; i < nwords; i += NT) {
                const int r = i >> 5, w = i & 31;
                uint32_t fg = 0, bg = 0;
                {   // own half: dy = r - RW <= 0, centred at column 512
                    const int y = sy + r - RW;
                    const int hw = r;                    // RW - |dy|
                    const int x0 = sx + 32 * w - 512;    // pixel of bit 0
                    uint32_t valid = span_mask(512 - hw - 32 * w, 512 + hw - 32 * w) &
                                     span_mask(-x0, p.W - 1 - x0);
                    if (y < 0 || y >= p.H) valid = 0;
                    if (valid) {
                        const uint32_t m = load_bits(mbits + (size_t)y * p.Wp, p.Wp, x0);
                        fg |= m & valid; bg |= ~m & valid;
                    }
                }
                if (r + 1 <= RW) {   // partner half: dy = r + 1 > 0, centred at column 0 (mod 1024)
                    const int y = sy + r + 1;
                    const int hw = RW - r - 1;
                    const int off = w < 16 ? 0 : 1024;
                    const int x0 = sx + 32 * w - off;
                    uint32_t valid = span_mask(off - hw - 32 * w, off + hw - 32 * w) &
                                     span_mask(-x0, p.W - 1 - x0);
                    if (y < 0 || y >= p.H) valid = 0;
                    if (valid) {
                        const uint32_t m = load_bits(mbits + (size_t)y * p.Wp, p.Wp, x0);
                        fg |= m & valid; bg |= ~m & valid;
                    }
                }
                OF[i] = fg;
                OB[i] = bg;
            }
            __syncthreads();

            // ---- classify targets (on / off mask), then close the source cell --------------
            for (int t = tid; t < ntgt; t += NT) {
                const uint32_t c = tcell[t];
                if (OF[c >> 5] & (1u << (c & 31))) tcell[t] = c | 0x80000000u;
            }
            __syncthreads();
            if (tid == 0) {
                const int c = fold_index(0, 0, RW);
                OF[c >> 5] &= ~(1u << (c & 31));
                OB[c >> 5] &= ~(1u << (c & 31));
                gs0[0] = pack_cell(0, 0);                 // level 0 is seeded by the source at j = 0
            }
            for (int j = tid; j < kSeedEnd; j += NT) send0[j] = 1u;
            __syncthreads();

            // ---- levels ---------------------------------------------------------------------
            int kstar = -1;
            bool my_pending = false;               // does this thread still own an unsettled target?
            for (int t = tid; t < ntgt; t += NT) my_pending |= (tres[t] == kUnset);
            bool all_settled = false;
            uint32_t seeds_total = 1;
            int sb = 0;
            int jscan = 0;
            for (int k = 0; k <= RW && seeds_total > 0 && !all_settled; ++k) {
                const uint32_t* const send_cur = send0 + sb * kSeedEnd;
                uint32_t* const send_nxt = send0 + (sb ^ 1) * kSeedEnd;
                const uint32_t* const gs_cur = gs0 + (size_t)sb * kGCap;
                uint32_t* const gs_nxt = gs0 + (size_t)(sb ^ 1) * kGCap;
                if (tid == 0) { s_seedcnt = 0; s_cnt[0] = 0; s_cnt[1] = 0; s_cnt[2] = 0; }
                int j = jscan;
                while (send_cur[j] == 0) ++j;          // first seeded j of this level
                jscan = j;
                uint32_t sp = 0;                       // seeds consumed
                uint32_t nlist = 0;                    // mask pixels claimed by the previous wave
                int step = 0;
                bool hit = false;
                __syncthreads();
                for (;;) {
                    const uint32_t send = send_cur[j];
                    const uint32_t ncur = nlist + (send - sp);
                    if (ncur == 0) {
                        if (sp >= seeds_total) break;          // level exhausted
                        if (tid == 0) send_nxt[j] = s_seedcnt;
                        ++j;
                        continue;
                    }
                    if (k + j >= RW) { hit = true; break; }    // cells at hop == max_px exist

                    const uint32_t* const qcur = q0 + (step & 1) * kQCap;
                    const uint32_t* const gqcur = gq0 + (size_t)(step & 1) * kGCap;
                    uint32_t* const qnxt = q0 + ((step + 1) & 1) * kQCap;
                    uint32_t* const gqnxt = gq0 + (size_t)((step + 1) & 1) * kGCap;
                    uint32_t* const cnt_nxt = &s_cnt[(step + 1) % 3];

                    const uint32_t ntask = ncur * 4;
                    for (uint32_t t0 = 0; t0 < ntask; t0 += NT) {
                        const uint32_t task = t0 + tid;
                        bool got_fg = false, got_bg = false;
                        uint32_t ncell = 0;
                        if (task < ntask) {
                            const uint32_t ci = task >> 2;
                            uint32_t e;
                            if (ci < nlist) e = ci < (uint32_t)kQCap ? qcur[ci] : gqcur[ci - kQCap];
                            else            e = gs_cur[sp + (ci - nlist)];
                            int dy = (int)((e >> 10) & 1023u) - 512;
                            int dx = (int)(e & 1023u) - 512;
                            const int dir = task & 3;
                            dy += (dir == 0) ? -1 : (dir == 3) ? 1 : 0;
                            dx += (dir == 1) ? -1 : (dir == 2) ? 1 : 0;
                            if (abs(dy) + abs(dx) <= RW) {
                                const int c = fold_index(dy, dx, RW);
                                const uint32_t bit = 1u << (c & 31);
                                const int w = c >> 5;
                                if (OF[w] & bit) {
                                    got_fg = (atomicAnd(&OF[w], ~bit) & bit) != 0;
                                } else if (OB[w] & bit) {
                                    got_bg = (atomicAnd(&OB[w], ~bit) & bit) != 0;
                                }
                                ncell = pack_cell(dy, dx);
                            }
                        }
                        const uint32_t bf = __ballot_sync(0xFFFFFFFFu, got_fg);
                        const uint32_t bb = __ballot_sync(0xFFFFFFFFu, got_bg);
                        if (bf) {
                            uint32_t base = 0;
                            if (lane == 0) base = atomicAdd(cnt_nxt, __popc(bf));
                            base = __shfl_sync(0xFFFFFFFFu, base, 0);
                            if (got_fg) {
                                const uint32_t slot = base + __popc(bf & ((1u << lane) - 1));
                                if (slot < (uint32_t)kQCap) qnxt[slot] = ncell; else gqnxt[slot - kQCap] = ncell;
                            }
                        }
                        if (bb) {
                            uint32_t base = 0;
                            if (lane == 0) base = atomicAdd(&s_seedcnt, __popc(bb));
                            base = __shfl_sync(0xFFFFFFFFu, base, 0);
                            if (got_bg) gs_nxt[base + __popc(bb & ((1u << lane) - 1))] = ncell;
                        }
                    }
                    // barrier + "is any target still unsettled" (as of the previous wave's poll);
                    // once every in-range target is settled no later level can change the answer
                    if (!__syncthreads_or(my_pending)) { all_settled = true; break; }
                    nlist = *cnt_nxt;
                    if (tid == 0) { s_cnt[step % 3] = 0; send_nxt[j] = s_seedcnt; }
                    sp = send;
                    // targets settled by this wave: mask pixels at (k, j+1), off-mask pixels at (k+1, j)
                    my_pending = false;
                    for (int t = tid; t < ntgt; t += NT) {
                        if (tres[t] == kUnset) {
                            const uint32_t c = tcell[t];
                            const uint32_t ci = c & 0x7FFFFFFFu;
                            const uint32_t bit = 1u << (ci & 31);
                            if (c & 0x80000000u) {
                                if (!(OF[ci >> 5] & bit)) tres[t] = ((uint32_t)k << 16) | (uint32_t)(j + 1);
                                else my_pending = true;
                            } else {
                                if (!(OB[ci >> 5] & bit)) tres[t] = ((uint32_t)(k + 1) << 16) | (uint32_t)j;
                                else my_pending = true;
                            }
                        }
                    }
                    ++j;
                    ++step;
                }
                // ---- level end ------------------------------------------------------------
                __syncthreads();
                const uint32_t total_next = s_seedcnt;
                for (int jj = tid; jj < kSeedEnd; jj += NT) {
                    if (jj >= j) send_nxt[jj] = total_next;
                    else if (jj < jscan) send_nxt[jj] = 0;
                }
                if (hit) kstar = k;
                seeds_total = (hit || all_settled) ? 0 : total_next;
                sb ^= 1;
                __syncthreads();
            }

            // ---- resolve ----------------------------------------------------------------------
            for (int t = tid; t < ntgt; t += NT) {
                const uint32_t r = tres[t];
                if (r == kUnset) continue;
                const int kT = (int)(r >> 16), jT = (int)(r & 0xFFFFu);
                if (kT + jT + 1 > p.max_px) continue;
                if (kstar >= 0 && kT > kstar) continue;
                const long long o = orow + tidx[t];
                p.path_len[o] = kT + jT + 1;
                p.unreachable[o] = 0;
                if (p.wcost) p.wcost[o] = (float)kT * 65536.0f + (float)jT;
            }
        }
    }
}

// work_prefix[i] = sum of n_src over pairs < i, work_prefix[n_pairs] = total (one CTA, tiny)
__global__ void work_prefix_kernel(const axb_pair_desc* __restrict__ pairs, int n_pairs, int* __restrict__ prefix) {
    __shared__ int s_carry;
    __shared__ int s_warp[32];
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < n_pairs; base += blockDim.x) {
        const int i = base + threadIdx.x;
        const int v = i < n_pairs ? max(pairs[i].n_src, 0) : 0;
        int x = v;
        for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(0xFFFFFFFFu, x, o);
            if ((threadIdx.x & 31) >= o) x += y;
        }
        if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = x;
        __syncthreads();
        if (threadIdx.x < 32) {
            int w = threadIdx.x < (blockDim.x >> 5) ? s_warp[threadIdx.x] : 0;
            for (int o = 1; o < 32; o <<= 1) {
                const int y = __shfl_up_sync(0xFFFFFFFFu, w, o);
                if (threadIdx.x >= o) w += y;
            }
            s_warp[threadIdx.x] = w;
        }
        __syncthreads();
        const int warp_off = (threadIdx.x >> 5) ? s_warp[(threadIdx.x >> 5) - 1] : 0;
        const int incl = s_carry + warp_off + x;
        if (i < n_pairs) prefix[i] = incl - v;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) s_carry = incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) prefix[n_pairs] = s_carry;
}

void launch_k2(const K2Params& p, int grid, cudaStream_t stream) {
    const size_t smem = k2_smem_bytes(p.max_px);
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(k2_search_kernel<kK2Threads>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)k2_smem_bytes(AXB_MAX_PX));
        attr_set = true;
    }
    k2_search_kernel<kK2Threads><<<grid, kK2Threads, smem, stream>>>(p);
}

void launch_work_prefix(const axb_pair_desc* pairs, int n_pairs, int* prefix, cudaStream_t stream) {
    work_prefix_kernel<<<1, 256, 0, stream>>>(pairs, n_pairs, prefix);
}

}  // namespace axb
