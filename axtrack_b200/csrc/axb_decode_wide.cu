// K1 wide -- the fused decode + threshold + stitch + NMS of axb_decode_nms.cu for launches of FEW frames (sm_100a).
//
// Same contract and the same results, bit for bit (AxonDetections._yolo_Y2pandas_det, AxonDetections.py:178-248;
// Timelapse.stitch_tiles coordinates, Timelapse.py:166-197; AxonDetections._non_max_supression, :250-278).  With one
// CTA per frame a launch of 1 .. 8 frames of an 8192 x 8192 image (BASELINE configs[4]: 256 tiles, 36 864 grid cells,
// up to ~12 000 candidates per frame) used 1 .. 8 of 148 SMs.  Here a frame is spread over G = ~num_sms / F CTAs of
// ONE cooperative launch; the phases are separated by grid.sync():
//   1. decode + threshold, one WARP per tile, order-preserving ballot compaction inside the tile's own slot range;
//   2. tile offsets (every CTA scans the per-tile counts itself), pre-NMS order = tile-major, conf ascending inside
//      a tile (rank by counting over the tile's <= Sx*Sy candidates), keys;
//   3. global order (conf descending, pre-NMS index ascending) as a RANK BY COUNTING: the keys are distinct, so
//      rank(i) = #{keys > key_i}; tasks of 256 candidates x 1024 keys staged in shared memory, partial counts added
//      with atomics -- O(n^2) compares spread over the whole GPU instead of a 105-pass bitonic sort on one SM;
//   4. scatter into rank order, spatial hash;  5. hash insert;
//   6. the greedy NMS fixpoint of the one-CTA kernel, rounds separated by grid.sync() (states only ever go
//      undecided -> kept / suppressed, so concurrent readers are safe);
//   7. survivors compacted in rank order: per-slice counts, then each slice writes at its offset.
#include "axb_decode.cuh"
#include <cooperative_groups.h>
#include <algorithm>
#include <cstdlib>

namespace cg = cooperative_groups;

namespace axb {

namespace {
constexpr int kWT = 256;                 // threads per CTA
constexpr int kWW = kWT / 32;
constexpr int kRankKeys = 1024;          // keys per rank task (8 KB of shared memory)
constexpr int kWideMaxTiles = 4096;      // tile offsets live in shared memory
}  // namespace

__global__ void __launch_bounds__(kWT) k1_wide_kernel(const K1Params p, const int G) {
    cg::grid_group grid = cg::this_grid();
    __shared__ int s_off[kWideMaxTiles + 1];
    __shared__ unsigned long long s_keys[kRankKeys];
    __shared__ int s_base;
    __shared__ int s_wc[kWW];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int f = blockIdx.x / G, g = blockIdx.x % G;
    const int cpt = p.Sx * p.Sy;
    const int ncell = p.ntiles * cpt;
    const K1Work w = carve(p.scratch + (size_t)f * p.scratch_per_frame, ncell, p.npad, p.ntiles);
    const float* __restrict__ Y = p.Y + (size_t)f * ncell * 3;
    const float ts = (float)p.tilesize, fSx = (float)p.Sx, fSy = (float)p.Sy;
    const uint32_t lt = (1u << lane) - 1u;
    volatile int* const ctl = p.wide_ctl;            // [0..2] undecided counters of the NMS rounds, rotating

    // ---- 1. decode + threshold: one warp per tile, candidates of tile t at c_*[t*cpt ...] in cell order ----
    for (int t = g * kWW + warp; t < p.ntiles; t += G * kWW) {
        int cnt = 0;
        for (int c0 = 0; c0 < cpt; c0 += 32) {
            const int c = c0 + lane;
            bool keep = false;
            float conf = 0.f; int xi = 0, yi = 0;
            const int gi = t * cpt + c;
            if (c < cpt) {
                conf = __ldg(Y + 3 * (size_t)gi);
                const float x = __ldg(Y + 3 * (size_t)gi + 1), y = __ldg(Y + 3 * (size_t)gi + 2);
                const int ix = c / p.Sy, iy = c % p.Sy;
                float xp = rintf(__fdiv_rn(__fmul_rn(__fadd_rn(x, (float)ix), ts), fSx));      // AxonDetections.py:203-207
                float yp = rintf(__fdiv_rn(__fmul_rn(__fadd_rn(y, (float)iy), ts), fSy));
                if (conf == 0.f && x == 0.f && y == 0.f) { xp = 0.f; yp = 0.f; }              // noanchor_mask, :194,209
                keep = conf >= p.conf_thr;                                                     // :220
                xi = __float2int_rn(xp); yi = __float2int_rn(yp);
            }
            const uint32_t bal = __ballot_sync(0xFFFFFFFFu, keep);
            if (keep) {
                const int slot = t * cpt + cnt + __popc(bal & lt);
                w.c_conf[slot] = conf; w.c_x[slot] = xi; w.c_y[slot] = yi; w.c_g[slot] = gi;
            }
            cnt += __popc(bal);
        }
        if (lane == 0) w.tile_cnt[t] = cnt;
    }
    grid.sync();

    // ---- 2. tile offsets (each CTA for itself), pre-NMS order, keys ----
    if (tid == 0) {
        int acc = 0;
        for (int t = 0; t < p.ntiles; ++t) { s_off[t] = acc; acc += w.tile_cnt[t]; }
        s_off[p.ntiles] = acc;
    }
    __syncthreads();
    const int n_found = s_off[p.ntiles];
    const bool too_many = n_found > p.cap;                       // the outputs cannot hold this frame
    const int n_pre = too_many ? 0 : n_found;
    if (g == 0 && tid == 0) {
        p.n_pre[f] = n_pre;
        if (too_many) { atomicOr(p.err_flags, 4ull); p.n_out[f] = 0; }
    }
    int nh = 2;
    while (nh < 2 * n_pre) nh <<= 1;
    const uint32_t hmask = (uint32_t)nh - 1;
    for (int i = g * kWT + tid; i < nh; i += G * kWT) w.head[i] = -1;
    for (int i = g * kWT + tid; i < n_pre; i += G * kWT) w.rank[i] = 0;
    if (!too_many) {
        for (int t = g * kWW + warp; t < p.ntiles; t += G * kWW) {
            const int cnt = s_off[t + 1] - s_off[t];
            const int base = t * cpt;
            for (int k = lane; k < cnt; k += 32) {
                const float conf = w.c_conf[base + k];
                int rank = 0;                                    // conf ascending inside the tile, ties in cell order (:235)
                for (int q = 0; q < cnt; ++q) {
                    const float cq = w.c_conf[base + q];
                    rank += (cq < conf || (cq == conf && q < k)) ? 1 : 0;
                }
                const int pi = s_off[t] + rank;
                const int fx = w.c_x[base + k] + __ldg(p.tile_yx + 2 * t + 1) * p.tilesize;     // Timelapse.py:191-192
                const int fy = w.c_y[base + k] + __ldg(p.tile_yx + 2 * t) * p.tilesize;
                w.p_conf[pi] = conf; w.p_x[pi] = fx; w.p_y[pi] = fy; w.p_tile[pi] = t;
                w.key[pi] = ((unsigned long long)ordered_bits(conf) << 32) | (unsigned long long)(0xFFFFFFFFu - (uint32_t)pi);
                if (p.pre_conf) p.pre_conf[(size_t)f * p.cap + pi] = conf;
                if (p.pre_yx) { p.pre_yx[((size_t)f * p.cap + pi) * 2] = fy; p.pre_yx[((size_t)f * p.cap + pi) * 2 + 1] = fx; }
                if (p.pre_cell) p.pre_cell[(size_t)f * p.cap + pi] = w.c_g[base + k];
            }
        }
    }
    grid.sync();

    // ---- 3. rank by counting: rank(i) = #{j : key_j > key_i} (keys are distinct) ----
    {
        const int cb = (n_pre + kWT - 1) / kWT, kb = (n_pre + kRankKeys - 1) / kRankKeys;
        for (int task = g; task < cb * kb; task += G) {
            const int ci = task / kb, ki = task % kb;
            const int k0 = ki * kRankKeys, kn = min(kRankKeys, n_pre - k0);
            __syncthreads();
            for (int i = tid; i < kn; i += kWT) s_keys[i] = w.key[k0 + i];
            __syncthreads();
            const int i = ci * kWT + tid;
            if (i < n_pre) {
                const unsigned long long mine = w.key[i];
                int cnt = 0;
                #pragma unroll 8
                for (int q = 0; q < kn; ++q) cnt += s_keys[q] > mine ? 1 : 0;
                if (cnt) atomicAdd(&w.rank[i], cnt);
            }
        }
    }
    grid.sync();

    // ---- 4. rank order ----
    int shift = 1;
    while ((1 << shift) * (1 << shift) < p.nms_d2) ++shift;     // bin edge >= nms_dist
    for (int pi = g * kWT + tid; pi < n_pre; pi += G * kWT) {
        const int r = w.rank[pi];
        w.s_p[r] = pi; w.s_x[r] = w.p_x[pi]; w.s_y[r] = w.p_y[pi];
        w.state[r] = 0;
    }
    grid.sync();

    // ---- 5. spatial hash ----
    for (int r = g * kWT + tid; r < n_pre; r += G * kWT) {
        const uint32_t h = bin_hash(w.s_x[r] >> shift, w.s_y[r] >> shift, hmask);
        w.next[r] = atomicExch(&w.head[h], r);
    }
    grid.sync();

    // ---- 6. NMS fixpoint, rounds over the whole grid ----
    volatile int32_t* const state = w.state;
    for (int round = 0;; ++round) {
        bool undecided = false;
        for (int r = g * kWT + tid; r < n_pre; r += G * kWT) {
            if (state[r] != 0) continue;
            const int x = w.s_x[r], y = w.s_y[r];
            const int bx = x >> shift, by = y >> shift;
            bool sup = false, wait = false;
            for (int db = 0; db < 9 && !sup; ++db) {
                const uint32_t h = bin_hash(bx + db % 3 - 1, by + db / 3 - 1, hmask);
                for (int q = w.head[h]; q >= 0; q = w.next[q]) {
                    if (q >= r) continue;                        // only better-ranked candidates matter
                    const long long ddx = (long long)w.s_x[q] - x, ddy = (long long)w.s_y[q] - y;
                    if (ddx * ddx + ddy * ddy < (long long)p.nms_d2) {
                        const int st = state[q];
                        if (st == 1) { sup = true; break; }
                        if (st == 0) wait = true;
                    }
                }
            }
            if (sup) state[r] = 2;
            else if (!wait) state[r] = 1;
            else undecided = true;
        }
        if (__syncthreads_or(undecided) && tid == 0) atomicAdd((int*)&ctl[round % 3], 1);
        if (blockIdx.x == 0 && tid == 0) ctl[(round + 1) % 3] = 0;      // read by everybody one round ago, written next round
        __threadfence();
        grid.sync();
        if (ctl[round % 3] == 0) break;                                  // grid-uniform: nobody is undecided any more
    }

    // ---- 7. survivors in rank order: slice g of the ranks, counts, offsets, outputs ----
    const int chunk = (n_pre + G - 1) / G;
    const int r_lo = min(g * chunk, n_pre), r_hi = min(r_lo + chunk, n_pre);
    {
        int mine = 0;
        for (int r = r_lo + tid; r < r_hi; r += kWT) mine += state[r] == 1 ? 1 : 0;
        mine = __reduce_add_sync(0xFFFFFFFFu, mine);
        if (lane == 0) s_wc[warp] = mine;
        __syncthreads();
        if (tid == 0) {
            int tot = 0;
            for (int v = 0; v < kWW; ++v) tot += s_wc[v];
            w.slice_keep[g] = tot;
        }
    }
    grid.sync();
    {
        int off = 0, tot = 0;
        for (int v = 0; v < G; ++v) { const int c = w.slice_keep[v]; if (v < g) off += c; tot += c; }
        if (g == 0 && tid == 0 && !too_many) p.n_out[f] = tot;
        __syncthreads();
        if (tid == 0) s_base = off;
        __syncthreads();
        for (int r0 = r_lo; r0 < r_hi; r0 += kWT) {
            const int r = r0 + tid;
            const bool keep = r < r_hi && state[r] == 1;
            const int slot = block_compact_slot<kWT>(keep, &s_base, s_wc);
            if (slot >= 0) {
                const int pi = w.s_p[r];
                const size_t o = (size_t)f * p.cap + slot;
                p.conf[o] = w.p_conf[pi];
                p.yx[2 * o] = w.s_y[r]; p.yx[2 * o + 1] = w.s_x[r];
                p.keep_idx[o] = pi;
            }
        }
    }
}

bool launch_k1_wide(const K1Params& p, cudaStream_t stream) {
    // AXB_K1_NO_WIDE (read per call: the tests toggle it) forces the one-CTA-per-frame kernel.  From ~num_sms / 4 frames
    // per launch on, a CTA per frame keeps the GPU busy without the grid-wide barriers.
    if (getenv("AXB_K1_NO_WIDE") != nullptr || p.F <= 0 || 4 * p.F > p.num_sms || p.ntiles > kWideMaxTiles) return false;
    static int coop = -1, per_sm = 0;
    if (coop < 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k1_wide_kernel, kWT, 0);
    }
    if (!coop || per_sm < 1) return false;
    // a frame's CTAs: enough warps for its tiles and rank tasks, at most what is co-resident, at most kK1MaxSlices
    int G = std::min(p.num_sms * std::min(per_sm, 2) / p.F, kK1MaxSlices);
    const long long ncell = (long long)p.ntiles * p.Sx * p.Sy;
    G = std::min(G, (int)std::max<long long>((p.ntiles + kWW - 1) / kWW, ncell / 256));
    if (G < 2) return false;
    cudaMemsetAsync(p.wide_ctl, 0, 8 * sizeof(int), stream);
    K1Params q = p;
    int Gv = G;
    void* args[] = {(void*)&q, (void*)&Gv};
    return cudaLaunchCooperativeKernel((const void*)k1_wide_kernel, dim3((unsigned)(p.F * G)), dim3(kWT), args, 0, stream) == cudaSuccess;
}

}  // namespace axb
