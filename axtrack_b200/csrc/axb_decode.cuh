// Shared device helpers of the K1 kernels (axb_decode_nms.cu: one CTA per frame; axb_decode_wide.cu: a frame
// spread over many CTAs of a cooperative launch): per-frame workspace layout, key order, compaction, spatial hash.
#pragma once
#include "axb_common.cuh"

namespace axb {

namespace {

constexpr int kK1Threads = 256;
constexpr int kSortSmem = 4096;     // keys sorted in shared memory up to this many

__device__ __forceinline__ uint32_t ordered_bits(float f) {
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

struct K1Work {
    float* c_conf; int32_t* c_x; int32_t* c_y; int32_t* c_g;      // compaction order
    float* p_conf; int32_t* p_x; int32_t* p_y; int32_t* p_tile;   // pre-NMS order
    unsigned long long* key;                                       // [npad]
    int32_t* s_x; int32_t* s_y; int32_t* s_p;                      // rank order
    int32_t* head;                                                 // [2*npad]
    int32_t* next;
    int32_t* state;
    int32_t* rank;                                                 // wide kernel: rank of each pre-NMS candidate
    int32_t* tile_cnt;                                             // wide kernel: candidates per tile
    int32_t* slice_keep;                                           // wide kernel: survivors per CTA slice [kK1MaxSlices]
};
constexpr int kK1MaxSlices = 1024;

__device__ __forceinline__ K1Work carve(uint8_t* base, int ncell, int npad, int ntiles = 0) {
    K1Work w;
    size_t o = 0;
    auto take = [&](size_t bytes) { uint8_t* p = base + o; o += (bytes + 15) & ~(size_t)15; return p; };
    w.key = (unsigned long long*)take((size_t)npad * 8);
    w.c_conf = (float*)take((size_t)ncell * 4);  w.c_x = (int32_t*)take((size_t)ncell * 4);
    w.c_y = (int32_t*)take((size_t)ncell * 4);   w.c_g = (int32_t*)take((size_t)ncell * 4);
    w.p_conf = (float*)take((size_t)ncell * 4);  w.p_x = (int32_t*)take((size_t)ncell * 4);
    w.p_y = (int32_t*)take((size_t)ncell * 4);   w.p_tile = (int32_t*)take((size_t)ncell * 4);
    w.s_x = (int32_t*)take((size_t)ncell * 4);   w.s_y = (int32_t*)take((size_t)ncell * 4);
    w.s_p = (int32_t*)take((size_t)ncell * 4);
    w.head = (int32_t*)take((size_t)npad * 2 * 4);
    w.next = (int32_t*)take((size_t)ncell * 4);
    w.state = (int32_t*)take((size_t)ncell * 4);
    w.rank = (int32_t*)take((size_t)ncell * 4);
    w.tile_cnt = (int32_t*)take((size_t)ntiles * 4);
    w.slice_keep = (int32_t*)take((size_t)kK1MaxSlices * 4);
    return w;
}

// order-preserving block compaction step: returns this thread's output slot (or -1) and
// advances *s_base.  3 barriers.  s_wc must hold NT/32 ints.
template <int NT>
__device__ __forceinline__ int block_compact_slot(bool pred, int* s_base, int* s_wc) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t bal = __ballot_sync(0xFFFFFFFFu, pred);
    if (lane == 0) s_wc[warp] = __popc(bal);
    __syncthreads();
    int off = *s_base, tot = 0;
    #pragma unroll
    for (int w = 0; w < NT / 32; ++w) { const int c = s_wc[w]; if (w < warp) off += c; tot += c; }
    const int slot = pred ? off + __popc(bal & ((1u << lane) - 1)) : -1;
    __syncthreads();
    if (threadIdx.x == 0) *s_base += tot;
    __syncthreads();
    return slot;
}

__device__ __forceinline__ uint32_t bin_hash(int bx, int by, uint32_t mask) {
    return (((uint32_t)bx * 73856093u) ^ ((uint32_t)by * 19349663u)) & mask;
}

}  // namespace

}  // namespace axb
