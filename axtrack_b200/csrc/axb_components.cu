// K2c -- connected components of the bit-packed masks and the L1 extents of each component (sm_100a).
//
// Why: a level-0 search from detection s settles every detection u of its mask component within max_px - 1 hops
// with the SAME path length the reference's A* finds from u to s (unit weights on the mask, astar.cpp:122-136: the
// cost of a mask path does not depend on its direction).  K2 uses that to answer a frame pair (t, t_bef) from the
// side of frame t when frame t_bef is not searched (axb_search.cu: k2_prep_kernel, "cover").  What the search from
// s cannot tell is the answer for a u it did NOT reach: the reference returns None for it iff u's own level 0
// contains a cell at hop max_px (axb_search.cu header), and that is certain when u's component holds a pixel at
// L1 distance >= max_px from u (hops along a path grow by one, so some pixel sits at hop max_px exactly).  The four
// extents max(y+x), max(y-x), max(-y+x), max(-y-x) of a component answer that with four subtractions.
//
// Union-find over the horizontal runs of the packed words (Playne & Hawick style atomicMin hooks):
//   k2c_init_kernel     label of a mask pixel = pixel index of the first pixel of its run inside its 32-bit word
//   k2c_merge_kernel    hooks runs across word boundaries and to the runs of the row above they touch
//   k2c_roots_kernel    clears the extents of the pixels that are still their own label (roots)
//   k2c_flatten_kernel  label = root for every mask pixel; extents by warp-aggregated atomicMax
// Labels: int32 [n_masks][H][W], -1 off the mask.  Extents: int4 [n_masks][H * W], valid at root pixels only.
#include "axb_common.cuh"
#include <climits>

namespace axb {

namespace {

constexpr unsigned kFullMask = 0xFFFFFFFFu;

__device__ __forceinline__ int find_root(const int* L, int a) {
    int l = L[a];
    while (l != a) { a = l; l = L[a]; }
    return a;
}

__device__ __forceinline__ void unite(int* L, int a, int b) {
    for (;;) {
        a = find_root(L, a);
        b = find_root(L, b);
        if (a == b) return;
        if (a < b) { const int t = a; a = b; b = t; }        // hook the larger root under the smaller
        const int old = atomicMin(&L[a], b);
        if (old == a) return;
        a = old;
    }
}

// lowest run of set bits of m: its mask; m loses it
__device__ __forceinline__ uint32_t pop_run(uint32_t& m) {
    const uint32_t low = m & (0u - m);
    const uint32_t t = m + low;                 // carries through the run (wraps to the bits above it, or to 0)
    const uint32_t run = m & ~t;
    m &= t;
    return run;
}

// One warp per packed word, one lane per pixel: coalesced label rows.
__global__ void __launch_bounds__(256) k2c_init_kernel(const uint32_t* __restrict__ bits, int n_masks, int H, int W, int Wp,
                                                       int* __restrict__ labels) {
    const long long n = (long long)n_masks * H * Wp;
    const int lane = threadIdx.x & 31;
    for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += ((long long)gridDim.x * blockDim.x) >> 5) {
        const int iw = (int)(i % Wp);
        const long long row = i / Wp;                        // mask * H + y
        const int y = (int)(row % H);
        const uint32_t m = bits[i];
        if (32 * iw + lane >= W) continue;
        int lab = -1;
        if ((m >> lane) & 1u) {
            const uint32_t below = ~m & ((1u << lane) - 1u);  // the closest clear bit below this pixel ends its run
            lab = y * W + 32 * iw + (below ? 32 - __clz(below) : 0);
        }
        labels[row * W + 32 * iw + lane] = lab;
    }
}

__global__ void __launch_bounds__(256) k2c_merge_kernel(const uint32_t* __restrict__ bits, int n_masks, int H, int W, int Wp,
                                                        int* __restrict__ labels) {
    const long long n = (long long)n_masks * H * Wp;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int iw = (int)(i % Wp);
        const long long row = i / Wp;
        const int y = (int)(row % H);
        const long long mk = row / H;
        int* const L = labels + mk * (long long)H * W;
        uint32_t m = bits[i];
        if (!m) continue;
        const int base = y * W + 32 * iw;
        if ((m & 1u) && iw > 0 && (bits[i - 1] >> 31)) unite(L, base, base - 1);
        if (y == 0) continue;
        const uint32_t up = bits[i - Wp];
        while (m) {
            const uint32_t run = pop_run(m);
            uint32_t o = run & up;                           // every maximal group of touching pixels needs one hook
            while (o) {
                const uint32_t sub = pop_run(o);
                const int b = __ffs(sub) - 1;
                unite(L, base + b, base - W + b);
            }
        }
    }
}

__global__ void __launch_bounds__(256) k2c_roots_kernel(const uint32_t* __restrict__ bits, int n_masks, int H, int W, int Wp,
                                                        const int* __restrict__ labels, int4* __restrict__ ext) {
    const long long n = (long long)n_masks * H * Wp;
    const int lane = threadIdx.x & 31;
    for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += ((long long)gridDim.x * blockDim.x) >> 5) {
        const uint32_t m = bits[i];
        if (!((m >> lane) & 1u)) continue;
        const int iw = (int)(i % Wp);
        const long long row = i / Wp;
        const int y = (int)(row % H);
        const long long mk = row / H;
        const int p = y * W + 32 * iw + lane;
        if (labels[mk * (long long)H * W + p] == p) ext[mk * (long long)H * W + p] = make_int4(INT_MIN, INT_MIN, INT_MIN, INT_MIN);
    }
}

// label = root for every mask pixel; extents: one atomicMax per (warp, root, extent) at most, and none when the
// extent is already large enough (it only grows, so a stale read errs on the safe side).
__global__ void __launch_bounds__(256) k2c_flatten_kernel(const uint32_t* __restrict__ bits, int n_masks, int H, int W, int Wp,
                                                          int* __restrict__ labels, int4* __restrict__ ext) {
    const long long n = (long long)n_masks * H * Wp;
    const int lane = threadIdx.x & 31;
    for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += ((long long)gridDim.x * blockDim.x) >> 5) {
        const uint32_t m = bits[i];
        if (!m) continue;                                    // warp-uniform
        const int iw = (int)(i % Wp);
        const long long row = i / Wp;
        const int y = (int)(row % H);
        const long long mk = row / H;
        int* const L = labels + mk * (long long)H * W;
        const int x = 32 * iw + lane;
        const int p = y * W + x;
        int r = -1;
        int4 v = make_int4(INT_MIN, INT_MIN, INT_MIN, INT_MIN);
        if ((m >> lane) & 1u) {
            r = find_root(L, L[p]);
            L[p] = r;
            v = make_int4(y + x, y - x, -y + x, -y - x);
        }
        const unsigned grp = __match_any_sync(kFullMask, r);
        v.x = __reduce_max_sync(grp, v.x); v.y = __reduce_max_sync(grp, v.y);
        v.z = __reduce_max_sync(grp, v.z); v.w = __reduce_max_sync(grp, v.w);
        if (r >= 0 && (int)(__ffs(grp) - 1) == lane) {
            int* const e = reinterpret_cast<int*>(ext + mk * (long long)H * W + r);
            if (*(volatile int*)(e + 0) < v.x) atomicMax(e + 0, v.x);
            if (*(volatile int*)(e + 1) < v.y) atomicMax(e + 1, v.y);
            if (*(volatile int*)(e + 2) < v.z) atomicMax(e + 2, v.z);
            if (*(volatile int*)(e + 3) < v.w) atomicMax(e + 3, v.w);
        }
    }
}

}  // namespace

// labels int32 [n_masks][H][W], ext int4 [n_masks][H * W]
void launch_k2_components(const uint32_t* mask_bits, int n_masks, int H, int W, int Wp, int* labels, int4* ext, cudaStream_t stream) {
    const long long n = (long long)n_masks * H * Wp;          // packed words
    const int grid_w = (int)((n + 255) / 256 < 148 * 16 ? (n + 255) / 256 : 148 * 16);       // a thread per word
    const int grid_p = (int)((n + 7) / 8 < 148 * 32 ? (n + 7) / 8 : 148 * 32);               // a warp per word
    k2c_init_kernel<<<grid_p, 256, 0, stream>>>(mask_bits, n_masks, H, W, Wp, labels);
    k2c_merge_kernel<<<grid_w, 256, 0, stream>>>(mask_bits, n_masks, H, W, Wp, labels);
    k2c_roots_kernel<<<grid_p, 256, 0, stream>>>(mask_bits, n_masks, H, W, Wp, labels, ext);
    k2c_flatten_kernel<<<grid_p, 256, 0, stream>>>(mask_bits, n_masks, H, W, Wp, labels, ext);
}

}  // namespace axb
