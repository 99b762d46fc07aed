// K1 -- fused detector-head decode + confidence threshold + stitch + distance NMS (sm_100a),
// and K0 -- mask bit-packing.
//
// K1 replaces, per frame, AxonDetections._yolo_Y2pandas_det (axtrack/AxonDetections.py:178-248),
// the coordinate part of Timelapse.stitch_tiles (axtrack/Timelapse.py:166-197) and
// AxonDetections._non_max_supression (AxonDetections.py:250-278).  One CTA per frame:
//   a. decode every grid cell with the reference's float32 operation order
//      x_px = rint(((x + ix) * tilesize) / Sx)  (AxonDetections.py:203-207; separate add, mul,
//      div, round-half-even), keep conf >= thr (f32 compare, :220), compact in cell order
//      (tile-major, cell = ix*Sy + iy, :215-220) with warp ballots;
//   b. pre-NMS index = position in the reference's concat: tile-major, inside a tile by conf
//      ascending (:235, Timelapse.py:194-196); ties keep cell order;
//   c. order all candidates by (conf descending, pre-NMS index ascending) -- bitonic sort of
//      64-bit keys in shared memory (global scratch when a frame has > 4096 candidates);
//   d. greedy NMS (suppress every later candidate with integer d^2 < nms_dist^2, :264-274) as a
//      fixpoint over a spatial hash: a candidate is suppressed once a KEPT better candidate is
//      in range, kept once every better candidate in range is decided; identical to the
//      sequential greedy result, needs only a few rounds on real data;
//   e. compact the survivors in rank order.
#include "axb_decode.cuh"

namespace axb {


template <int NT>
__global__ void __launch_bounds__(NT) k1_decode_nms_kernel(const K1Params p) {
    __shared__ unsigned long long s_keys[kSortSmem];
    __shared__ int s_base;
    __shared__ int s_wc[NT / 32];
    const int tid = threadIdx.x;
    const int f = blockIdx.x;
    const int cpt = p.Sx * p.Sy;
    const int ncell = p.ntiles * cpt;
    const K1Work w = carve(p.scratch + (size_t)f * p.scratch_per_frame, ncell, p.npad, p.ntiles);
    const float* __restrict__ Y = p.Y + (size_t)f * ncell * 3;
    const float ts = (float)p.tilesize, fSx = (float)p.Sx, fSy = (float)p.Sy;

    // ---- a. decode + threshold + compaction ------------------------------------------------
    if (tid == 0) s_base = 0;
    __syncthreads();
    for (int g0 = 0; g0 < ncell; g0 += NT) {
        const int g = g0 + tid;
        bool keep = false;
        float conf = 0.f; int xi = 0, yi = 0;
        if (g < ncell) {
            conf = __ldg(Y + 3 * (size_t)g);
            const float x = __ldg(Y + 3 * (size_t)g + 1), y = __ldg(Y + 3 * (size_t)g + 2);
            const int c = g % cpt;
            const int ix = c / p.Sy, iy = c % p.Sy;
            float xp = rintf(__fdiv_rn(__fmul_rn(__fadd_rn(x, (float)ix), ts), fSx));
            float yp = rintf(__fdiv_rn(__fmul_rn(__fadd_rn(y, (float)iy), ts), fSy));
            if (conf == 0.f && x == 0.f && y == 0.f) { xp = 0.f; yp = 0.f; }   // noanchor_mask, :194,209
            keep = conf >= p.conf_thr;
            xi = __float2int_rn(xp); yi = __float2int_rn(yp);
        }
        const int slot = block_compact_slot<NT>(keep, &s_base, s_wc);
        if (slot >= 0) { w.c_conf[slot] = conf; w.c_x[slot] = xi; w.c_y[slot] = yi; w.c_g[slot] = g; }
    }
    __syncthreads();
    const int n_pre = s_base;
    if (tid == 0) p.n_pre[f] = n_pre;
    if (n_pre > p.cap) {      // outputs cannot hold this frame
        if (tid == 0) { atomicOr(p.err_flags, 4ull); p.n_out[f] = 0; p.n_pre[f] = 0; }
        return;
    }

    // ---- b. pre-NMS order -------------------------------------------------------------------
    for (int c = tid; c < n_pre; c += NT) {
        const int g = w.c_g[c];
        const int tile = g / cpt;
        int lo, hi;
        {   // candidates of this tile are contiguous: [lower_bound(tile*cpt), lower_bound((tile+1)*cpt))
            int a = 0, b = n_pre;
            const int v0 = tile * cpt, v1 = v0 + cpt;
            while (a < b) { const int m = (a + b) >> 1; if (w.c_g[m] < v0) a = m + 1; else b = m; }
            lo = a; b = n_pre;
            while (a < b) { const int m = (a + b) >> 1; if (w.c_g[m] < v1) a = m + 1; else b = m; }
            hi = a;
        }
        const float conf = w.c_conf[c];
        int rank = 0;
        for (int q = lo; q < hi; ++q) {
            const float cq = w.c_conf[q];
            rank += (cq < conf || (cq == conf && q < c)) ? 1 : 0;
        }
        const int pi = lo + rank;
        const int fx = w.c_x[c] + __ldg(p.tile_yx + 2 * tile + 1) * p.tilesize;
        const int fy = w.c_y[c] + __ldg(p.tile_yx + 2 * tile) * p.tilesize;
        w.p_conf[pi] = conf; w.p_x[pi] = fx; w.p_y[pi] = fy; w.p_tile[pi] = tile;
        w.key[pi] = ((unsigned long long)ordered_bits(conf) << 32) | (unsigned long long)(0xFFFFFFFFu - (uint32_t)pi);
        if (p.pre_conf) p.pre_conf[(size_t)f * p.cap + pi] = conf;
        if (p.pre_yx) { p.pre_yx[((size_t)f * p.cap + pi) * 2] = fy; p.pre_yx[((size_t)f * p.cap + pi) * 2 + 1] = fx; }
        if (p.pre_cell) p.pre_cell[(size_t)f * p.cap + pi] = g;
    }
    int nsort = 1;
    while (nsort < n_pre) nsort <<= 1;
    for (int i = n_pre + tid; i < nsort; i += NT) w.key[i] = 0ull;     // below every real key
    __syncthreads();

    // ---- c. sort by (conf desc, pre-NMS index asc) -----------------------------------------------
    unsigned long long* K = w.key;
    if (nsort <= kSortSmem) {
        for (int i = tid; i < nsort; i += NT) s_keys[i] = w.key[i];
        K = s_keys;
        __syncthreads();
    }
    for (int k = 2; k <= nsort; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < nsort; i += NT) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const unsigned long long a = K[i], b = K[ixj];
                    const bool desc = (i & k) == 0;
                    if (desc ? (a < b) : (a > b)) { K[i] = b; K[ixj] = a; }
                }
            }
            __syncthreads();
        }
    }
    int shift = 1;
    while ((1 << shift) * (1 << shift) < p.nms_d2) ++shift;     // bin edge >= nms_dist
    int nh = 2;
    while (nh < 2 * n_pre) nh <<= 1;
    const uint32_t hmask = (uint32_t)nh - 1;
    for (int i = tid; i < nh; i += NT) w.head[i] = -1;
    for (int r = tid; r < n_pre; r += NT) {
        const int pi = (int)(0xFFFFFFFFu - (uint32_t)(K[r] & 0xFFFFFFFFull));
        w.s_p[r] = pi; w.s_x[r] = w.p_x[pi]; w.s_y[r] = w.p_y[pi];
        w.state[r] = 0;
    }
    __syncthreads();

    // ---- d. NMS fixpoint over a spatial hash ------------------------------------------------------
    for (int r = tid; r < n_pre; r += NT) {
        const uint32_t h = bin_hash(w.s_x[r] >> shift, w.s_y[r] >> shift, hmask);
        w.next[r] = atomicExch(&w.head[h], r);
    }
    __syncthreads();
    volatile int32_t* const state = w.state;
    bool undecided;
    do {
        undecided = false;
        for (int r = tid; r < n_pre; r += NT) {
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
    } while (__syncthreads_or(undecided));

    // ---- e. survivors in rank order -----------------------------------------------------------------
    if (tid == 0) s_base = 0;
    __syncthreads();
    for (int r0 = 0; r0 < n_pre; r0 += NT) {
        const int r = r0 + tid;
        const bool keep = r < n_pre && w.state[r] == 1;
        const int slot = block_compact_slot<NT>(keep, &s_base, s_wc);
        if (slot >= 0) {
            const int pi = w.s_p[r];
            const size_t o = (size_t)f * p.cap + slot;
            p.conf[o] = w.p_conf[pi];
            p.yx[2 * o] = w.s_y[r]; p.yx[2 * o + 1] = w.s_x[r];
            p.keep_idx[o] = pi;
        }
    }
    __syncthreads();
    if (tid == 0) p.n_out[f] = s_base;
}

size_t k1_scratch_per_frame(int ncell, int npad, int ntiles) {
    auto al = [](size_t b) { return (b + 15) & ~(size_t)15; };
    return al((size_t)npad * 8) + 14 * al((size_t)ncell * 4) + al((size_t)npad * 2 * 4) + al((size_t)ntiles * 4) + al((size_t)kK1MaxSlices * 4) + 256;
}

void launch_k1(const K1Params& p, cudaStream_t stream) {
    // few frames per launch (configs[4] sweeps 1 .. 8): one CTA per frame leaves the GPU empty; a frame is spread over
    // many CTAs of a cooperative launch instead (axb_decode_wide.cu)
    if (launch_k1_wide(p, stream)) return;
    // one CTA per frame; frames with many grid cells (configs[4]: 36864) get 32 warps for the
    // decode pass, the bitonic sort and the NMS rounds
    if ((long long)p.ntiles * p.Sx * p.Sy > 4096) k1_decode_nms_kernel<1024><<<p.F, 1024, 0, stream>>>(p);
    else k1_decode_nms_kernel<kK1Threads><<<p.F, kK1Threads, 0, stream>>>(p);
}

// ---------------------------------------------------------------------------------------------
// K0: mask bit-packing.  One thread per output word: 32 pixels -> 32 bits, bit i = pixel
// 32*word+i.  HBM-bound: 1 B (uint8) or 4 B (float32) read per pixel, 1/8 B written.
// ---------------------------------------------------------------------------------------------
namespace {
__device__ __forceinline__ uint32_t eq1_nibble(uint32_t v) {
    // 4 bytes -> 4 bits (byte == 1)
    const uint32_t e = __vcmpeq4(v, 0x01010101u) & 0x01010101u;
    return ((e * 0x01020408u) >> 24) & 0xFu;
}
}  // namespace

__global__ void pack_mask_u8_kernel(const uint8_t* __restrict__ mask, long long rows, int W, int Wp,
                                    uint32_t* __restrict__ bits, int vec_ok) {
    const long long nwords = rows * Wp;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nwords;
         i += (long long)gridDim.x * blockDim.x) {
        const long long y = i / Wp;
        const int wx = (int)(i - y * Wp);
        const int x0 = wx * 32;
        const uint8_t* src = mask + y * (long long)W + x0;
        uint32_t out = 0;
        if (vec_ok) {      // W % 16 == 0 and 16-byte aligned base: rows start on 16 B boundaries
            const uint4 a = __ldg(reinterpret_cast<const uint4*>(src));
            out = eq1_nibble(a.x) | (eq1_nibble(a.y) << 4) | (eq1_nibble(a.z) << 8) | (eq1_nibble(a.w) << 12);
            if (x0 + 16 < W) {
                const uint4 b = __ldg(reinterpret_cast<const uint4*>(src) + 1);
                out |= (eq1_nibble(b.x) | (eq1_nibble(b.y) << 4) | (eq1_nibble(b.z) << 8) | (eq1_nibble(b.w) << 12)) << 16;
            }
        } else {
            const int n = min(32, W - x0);
            for (int b = 0; b < n; ++b) out |= (uint32_t)(__ldg(src + b) == 1) << b;
        }
        bits[i] = out;
    }
}

__global__ void pack_mask_f32_kernel(const float* __restrict__ wts, long long rows, int W, int Wp,
                                     uint32_t* __restrict__ bits, unsigned long long* err_flags) {
    // one warp per output word: lane b reads pixel 32*word+b (coalesced 128 B), ballot packs
    const long long nwords = rows * Wp;
    const int lane = threadIdx.x & 31;
    const long long warp0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    unsigned flags = 0;
    for (long long i = warp0; i < nwords; i += nwarps) {
        const long long y = i / Wp;
        const int x = (int)(i - y * Wp) * 32 + lane;
        float v = 65536.f;
        if (x < W) {
            v = __ldg(wts + y * (long long)W + x);
            if (v < 1.f) flags |= 1u;                                  // astar_wrapper.py:35
            else if (!(v == 1.f || v == 65536.f)) flags |= 2u;
        }
        const uint32_t word = __ballot_sync(0xFFFFFFFFu, x < W && v == 1.f);
        if (lane == 0) bits[i] = word;
    }
    if (flags) atomicOr(err_flags, (unsigned long long)flags);
}

void launch_pack_mask_u8(const uint8_t* mask, int n_masks, int H, int W, uint32_t* bits, cudaStream_t s) {
    const int Wp = (W + 31) / 32;
    const long long rows = (long long)n_masks * H;
    const long long nwords = rows * Wp;
    if (nwords == 0) return;
    const int vec_ok = (W % 16 == 0) && ((reinterpret_cast<uintptr_t>(mask) & 15) == 0);
    const int grid = (int)std::min<long long>((nwords + 255) / 256, 148 * 16);
    pack_mask_u8_kernel<<<grid, 256, 0, s>>>(mask, rows, W, Wp, bits, vec_ok);
}

void launch_pack_mask_f32(const float* w, int n_masks, int H, int W, uint32_t* bits,
                          unsigned long long* err_flags, cudaStream_t s) {
    const int Wp = (W + 31) / 32;
    const long long rows = (long long)n_masks * H;
    const long long nwords = rows * Wp;
    if (nwords == 0) return;
    const int grid = (int)std::min<long long>((nwords * 32 + 255) / 256, 148 * 16);
    pack_mask_f32_kernel<<<grid, 256, 0, s>>>(w, rows, W, Wp, bits, err_flags);
}

}  // namespace axb
