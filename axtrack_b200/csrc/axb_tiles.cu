// K6 -- tiling front end of the detector (sm_100a), SURVEY 8f rank 4.
//
// Replaces, for inference, the image half of Timelapse.construct_tiles (axtrack/Timelapse.py:508-525,
// 539-558: torch.split into tilesize x tilesize tiles, ZeroPad2d at the right / bottom border, the
// "non-empty tile" flags (X[:, :, :, 0] > 0).any(-1).any(-1)) and Timelapse.get_frametiles_stack
// (Timelapse.py:150-157 over __getitem__ :111-125: per tile the temporal stack
// X_tiled[t - tc : t + tc + 1, tile, channels].flatten(end_dim=1)), which the reference runs as a Python
// loop over tiles per frame and one model forward per frame.  Here the timelapse X [T][C][H][W] stays
// where it is in HBM (no X_tiled copy of the whole timelapse) and
//   k6_nonempty_kernel  CTAs per (tile, frame, 64-row band): any(X[t][0][tile] > 0), early exit per 8 rows;
//   k6_gather_kernel    writes the model input of a BATCH of frames [B][ntiles][(2tc+1)*nch][ts][ts]
//                       straight from X, zero filled outside the image: a pure copy, 4 B read + 4 B
//                       written per element, float4 when the rows allow it.  Both are HBM bound.
#include "axb_common.cuh"
#include <cuda.h>          // CUtensorMap + the prototype of cuTensorMapEncodeTiled (resolved at run time, no -lcuda)
#include <algorithm>
#include <mutex>

namespace axb {

// flags are zeroed by the launcher; blockIdx.z = a band of kBandRows tile rows, so an empty 512x512 tile is
// read by 8 CTAs instead of one.  A CTA that finds a positive pixel stores 1 (every writer stores the same
// word, no atomic needed) and leaves; a CTA whose flag is already set leaves at once.  The vote flags sit
// one per 32-byte sector (kK6FlagStride words apart): packed bytes made the stores of all bands of 32
// (tile, frame) neighbours serialise on one L2 sector; k6_flags_to_bytes_kernel packs them afterwards.
constexpr int kBandRows = 64;
constexpr int kRoundRows = 8;      // rows between two early-exit checks
template <bool VEC>
__global__ void __launch_bounds__(256) k6_nonempty_kernel(const float* __restrict__ X, int T, int C, int H, int W, int ts,
                                                          int xtiles, uint32_t* __restrict__ flags) {
    const int tile = blockIdx.x, t = blockIdx.y;
    const int ty = tile / xtiles, tx = tile - ty * xtiles;
    const int band = (int)blockIdx.z;                              // signed: H - y0 below may be negative
    const int y0 = ty * ts + band * kBandRows, x0 = tx * ts;
    const int h = min(min(kBandRows, ts - band * kBandRows), H - y0), w = min(ts, W - x0);
    if (h <= 0) return;
    const float* const img = X + ((size_t)t * C) * H * W;          // channel 0 of frame t
    volatile uint32_t* const flag = flags + ((size_t)tile * T + t) * kK6FlagStride;
    for (int r0 = 0; r0 < h; r0 += kRoundRows) {
        const int rows = min(kRoundRows, h - r0);
        int mine = (int)*flag;
        if (VEC) {                                                 // W % 4 == 0 and ts % 4 == 0: w % 4 == 0
            const int wq = w >> 2;
            for (int i = threadIdx.x; i < rows * wq; i += blockDim.x) {
                const int r = i / wq, c = (i - r * wq) << 2;
                const float4 v = __ldg(reinterpret_cast<const float4*>(img + (size_t)(y0 + r0 + r) * W + x0 + c));
                mine |= (v.x > 0.0f) | (v.y > 0.0f) | (v.z > 0.0f) | (v.w > 0.0f);
            }
        } else {
            for (int i = threadIdx.x; i < rows * w; i += blockDim.x) {
                const int r = i / w, c = i - r * w;
                mine |= __ldg(img + (size_t)(y0 + r0 + r) * W + x0 + c) > 0.0f;
            }
        }
        if (__syncthreads_or(mine)) {
            if (threadIdx.x == 0 && *flag == 0) *flag = 1;
            return;
        }
    }
}

__global__ void __launch_bounds__(256) k6_flags_to_bytes_kernel(const uint32_t* __restrict__ flags, size_t n, uint8_t* __restrict__ nonempty) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) nonempty[i] = (uint8_t)(flags[i * kK6FlagStride] != 0);
}

template <bool VEC>
__global__ void __launch_bounds__(256) k6_gather_kernel(const float* __restrict__ X, int T, int C, int H, int W, int ts,
                                                        const int32_t* __restrict__ tile_yx, int ntiles,
                                                        const int32_t* __restrict__ frames, int B, int tc, int chan_lo,
                                                        int nch, int chunks, float* __restrict__ out,
                                                        unsigned long long* err_flags) {
    constexpr int ROWS = 16;
    const int nst = 2 * tc + 1;
    const long long n_planes = (long long)B * ntiles * nst * nch;
    for (long long job = blockIdx.x; job < n_planes * chunks; job += gridDim.x) {
        // job order: (tile, chunk, ch) major, (b, dt) minor -- the 2tc+1 readers of one source row block
        // (neighbouring timepoints) run back to back, so HBM delivers every input row once and L2 the rest
        long long q = job;
        const int dt = (int)(q % nst); q /= nst;
        const int b = (int)(q % B); q /= B;
        const int ch = (int)(q % nch); q /= nch;
        const int chunk = (int)(q % chunks);
        const int tile = (int)(q / chunks);
        const long long plane = (((long long)b * ntiles + tile) * nst + dt) * nch + ch;
        const int t = frames[b] + dt - tc;
        const bool bad = t < 0 || t >= T;                         // CTA-uniform
        if (bad && threadIdx.x == 0) atomicOr(err_flags, 16ull);
        const int y0 = tile_yx[2 * tile] * ts, x0 = tile_yx[2 * tile + 1] * ts;
        const float* const img = X + ((size_t)(bad ? 0 : t) * C + chan_lo + ch) * H * W;
        float* const dst = out + (size_t)plane * ts * ts;
        const int r_lo = chunk * ROWS, r_hi = min(ts, r_lo + ROWS);
        if (VEC) {
            const int wq = ts >> 2;
            for (int i = threadIdx.x; i < (r_hi - r_lo) * wq; i += blockDim.x) {
                const int r = r_lo + i / wq, c = (i % wq) << 2;
                const int y = y0 + r, x = x0 + c;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (!bad && y < H && x < W) v = __ldg(reinterpret_cast<const float4*>(img + (size_t)y * W + x));   // W % 4 == 0: all four in range
                __stcs(reinterpret_cast<float4*>(dst + (size_t)r * ts + c), v);
            }
        } else {
            for (int i = threadIdx.x; i < (r_hi - r_lo) * ts; i += blockDim.x) {
                const int r = r_lo + i / ts, c = i % ts;
                const int y = y0 + r, x = x0 + c;
                dst[(size_t)r * ts + c] = (!bad && y < H && x < W) ? __ldg(img + (size_t)y * W + x) : 0.0f;
            }
        }
    }
}

// The frame a tracker draws its crops from (AxonDetections.get_frame_and_truedets, AxonDetections.py:355-376 ->
// Timelapse.stitch_tiles image half, Timelapse.py:173-188): the reference cuts the frame into tiles, drops the
// tiles that are empty at every timepoint and stitches the rest back -- i.e. the dense frame with the dropped
// tiles zeroed.  Here: one pass over the dense X, out[ch][y][x] = keep[tile(y, x)] ? X[frame][chan_lo + ch][y][x] : 0.
__global__ void __launch_bounds__(256) k6_frame_kernel(const float* __restrict__ X, int C, int H, int W, int ts, int xtiles,
                                                       const uint8_t* __restrict__ keep, int frame, int chan_lo, int nch,
                                                       float* __restrict__ out) {
    const size_t n = (size_t)nch * H * W;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int x = (int)(i % W);
        const size_t r = i / W;
        const int y = (int)(r % H), ch = (int)(r / H);
        const bool k = keep[(y / ts) * xtiles + x / ts] != 0;
        out[i] = k ? __ldg(X + (((size_t)frame * C + chan_lo + ch) * H + y) * W + x) : 0.0f;
    }
}

void launch_k6_frame(const float* X, int C, int H, int W, int ts, const uint8_t* keep, int frame, int chan_lo, int nch,
                     float* out, int n_sms, cudaStream_t stream) {
    const size_t n = (size_t)nch * H * W;
    const int grid = (int)std::min<size_t>((n + 255) / 256, (size_t)n_sms * 16);
    k6_frame_kernel<<<grid > 0 ? grid : 1, 256, 0, stream>>>(X, C, H, W, ts, (W + ts - 1) / ts, keep, frame, chan_lo, nch, out);
}


// ---- k6_gather_tma_kernel: the same copy through the Tensor Memory Accelerator ------------------------------------
// The gather is a strided 2-D copy with zero fill at the image border -- what TMA does in hardware.  The dense
// timelapse is described by ONE 3-D tensor map (W, H, T*C; out-of-bound elements read as 0, which is both the
// ZeroPad2d of Timelapse.py:521-525 and the zero plane of a frame outside the timelapse) and the output stack by
// one 2-D map ([n_planes * ts] rows of ts floats).  A CTA is one warp whose elected lane drives a ring of kTmaStages
// shared-memory boxes (32 KB each): cp.async.bulk.tensor global -> shared (arrives on the stage's mbarrier), then
// cp.async.bulk.tensor shared -> global of the same box, kTmaAhead loads in flight ahead of the stores.  No
// register ever holds a pixel; addresses, bounds and the zero fill are the copy engine's.
#ifndef AXB_TMA_BOXW
// measured on B200 (profiles/r02i_k6_tma_variants.log, 16 frames of 2048^2 / 4096^2 per launch): box 256 x 32, 4 stages,
// 2 loads ahead, ONE CTA per SM 6402 / 6502 GB/s; 128 x 64 6049 / 6079; 128 x 32 with 8 stages x 2 CTAs 5611 / 5630;
// 256 x 16 5088 / 4840; 128 x 16 x 4 CTAs 4876 / 4668 -- against 5152 / 5311 for the LDG/STG kernel
#define AXB_TMA_BOXW 256
#define AXB_TMA_BOXH 32
#define AXB_TMA_STAGES 4
#define AXB_TMA_AHEAD 2
#define AXB_TMA_CTAS 1
#endif
constexpr int kTmaBoxW = AXB_TMA_BOXW, kTmaBoxH = AXB_TMA_BOXH;   // one box: 256 x 32 floats = 32 KB
constexpr int kTmaStages = AXB_TMA_STAGES, kTmaAhead = AXB_TMA_AHEAD;
constexpr int kTmaBoxBytes = kTmaBoxW * kTmaBoxH * 4;

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(32) k6_gather_tma_kernel(const __grid_constant__ CUtensorMap in_map,
                                                           const __grid_constant__ CUtensorMap out_map,
                                                           int T, int C, int ts, const int32_t* __restrict__ tile_yx, int ntiles,
                                                           const int32_t* __restrict__ frames, int B, int tc, int chan_lo, int nch,
                                                           unsigned long long* err_flags) {
    extern __shared__ __align__(128) uint8_t tma_smem[];
    __shared__ __align__(8) unsigned long long mbar[kTmaStages];
    __shared__ int2 out_xy[kTmaStages];                      // destination of the box a stage holds
    if (threadIdx.x != 0) return;                            // one elected thread issues every copy
    const unsigned nst = 2 * tc + 1;
    const unsigned bx = ts / kTmaBoxW, by = ts / kTmaBoxH;   // boxes per tile row / column (ts is a multiple of both)
    const unsigned n_jobs = (unsigned)B * ntiles * nst * nch * bx * by;      // < 2^31 (checked by the launcher)
    for (int s = 0; s < kTmaStages; ++s)
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_addr(&mbar[s])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const unsigned n_mine = blockIdx.x < n_jobs ? (n_jobs - blockIdx.x + gridDim.x - 1) / gridDim.x : 0u;
    // job -> (source plane / coordinates, destination row): order (tile, box row, box column, ch) major, (b, dt) minor,
    // so the 2tc+1 readers of one source box run back to back (HBM delivers it once, L2 the rest)
    auto decode = [&](unsigned job, int& x, int& y, int& plane_in, int& col_out, int& row_out) {
        unsigned q = job;
        const int dt = (int)(q % nst); q /= nst;
        const int b = (int)(q % (unsigned)B); q /= (unsigned)B;
        const int ch = (int)(q % (unsigned)nch); q /= (unsigned)nch;
        const int cb = (int)(q % bx); q /= bx;
        const int rb = (int)(q % by);
        const int tile = (int)(q / by);
        const int t = frames[b] + dt - tc;
        if (t < 0 || t >= T) atomicOr(err_flags, 16ull);
        x = tile_yx[2 * tile + 1] * ts + cb * kTmaBoxW;
        y = tile_yx[2 * tile] * ts + rb * kTmaBoxH;
        plane_in = (t < 0 || t >= T) ? -1 : t * C + chan_lo + ch;           // -1: out of bounds -> a box of zeros
        const long long plane = (((long long)b * ntiles + tile) * nst + dt) * nch + ch;
        col_out = cb * kTmaBoxW;
        row_out = (int)(plane * ts) + rb * kTmaBoxH;
    };
    for (unsigned it = 0; it < n_mine + kTmaAhead; ++it) {
        if (it < n_mine) {                                   // load job `it` into stage it % S
            const int s = (int)(it % kTmaStages);
            // the store that last read this stage (job it - S) has finished reading shared memory
            asm volatile("cp.async.bulk.wait_group.read %0;" :: "n"(kTmaStages - kTmaAhead - 1) : "memory");
            int x, y, pl, co, ro;
            decode(blockIdx.x + it * gridDim.x, x, y, pl, co, ro);
            out_xy[s] = make_int2(co, ro);
            const uint32_t bar = smem_addr(&mbar[s]), dst = smem_addr(tma_smem + (size_t)s * kTmaBoxBytes);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(kTmaBoxBytes) : "memory");
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         :: "r"(dst), "l"(&in_map), "r"(x), "r"(y), "r"(pl), "r"(bar) : "memory");
        }
        if (it >= kTmaAhead) {                               // store job it - A from its stage
            const unsigned k = it - kTmaAhead;
            const int s = (int)(k % kTmaStages);
            const uint32_t parity = (uint32_t)((k / kTmaStages) & 1);
            const uint32_t bar = smem_addr(&mbar[s]), src = smem_addr(tma_smem + (size_t)s * kTmaBoxBytes);
            asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@!p bra WAIT_%=;\n\t}"
                         :: "r"(bar), "r"(parity) : "memory");
            const int2 o = out_xy[s];
            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];"
                         :: "l"(&out_map), "r"(o.x), "r"(o.y), "r"(src) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

namespace {
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess && qr == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}
}  // namespace

// true when the launch went through TMA; false: the shapes do not fit its descriptors, use k6_gather_kernel
bool launch_k6_gather_tma(const float* X, int T, int C, int H, int W, int ts, const int32_t* tile_yx, int ntiles,
                          const int32_t* frames, int B, int tc, int chan_lo, int nch, float* out,
                          unsigned long long* err_flags, int n_sms, cudaStream_t stream) {
    static const bool off = getenv("AXB_K6_NO_TMA") != nullptr;
    const long long n_planes = (long long)B * ntiles * (2 * tc + 1) * nch;
    if (off || ts % kTmaBoxW || ts % kTmaBoxH || W % 4 || (reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(out)) % 16 ||
        n_planes * ts > 0x7fffffffLL || (long long)T * C > 0x7fffffffLL ||
        n_planes * (ts / kTmaBoxW) * (ts / kTmaBoxH) > 0x7fffffffLL)
        return false;
    EncodeTiledFn enc = encode_tiled();
    if (!enc) return false;
    CUtensorMap in_map, out_map;
    {
        const cuuint64_t dim[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)T * C};
        const cuuint64_t stride[2] = {(cuuint64_t)W * 4, (cuuint64_t)W * H * 4};
        const cuuint32_t box[3] = {kTmaBoxW, kTmaBoxH, 1}, es[3] = {1, 1, 1};
        if (enc(&in_map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)X, dim, stride, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
            return false;
    }
    {
        const cuuint64_t dim[2] = {(cuuint64_t)ts, (cuuint64_t)(n_planes * ts)};
        const cuuint64_t stride[1] = {(cuuint64_t)ts * 4};
        const cuuint32_t box[2] = {kTmaBoxW, kTmaBoxH}, es[2] = {1, 1};
        if (enc(&out_map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)out, dim, stride, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
            return false;
    }
    static unsigned long long attr_set = 0ull;
    const size_t smem = (size_t)kTmaStages * kTmaBoxBytes;
    if (first_use_on_device(attr_set))
        cudaFuncSetAttribute(k6_gather_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const long long jobs = n_planes * (ts / kTmaBoxW) * (ts / kTmaBoxH);
    const int grid = (int)std::min<long long>(jobs, (long long)n_sms * AXB_TMA_CTAS);
    k6_gather_tma_kernel<<<grid, 32, smem, stream>>>(in_map, out_map, T, C, ts, tile_yx, ntiles, frames, B, tc, chan_lo, nch, err_flags);
    return true;
}

void launch_k6_nonempty(const float* X, int T, int C, int H, int W, int ts, uint32_t* flags, uint8_t* nonempty, cudaStream_t stream) {
    const int ytiles = (H + ts - 1) / ts, xtiles = (W + ts - 1) / ts;
    const size_t n = (size_t)ytiles * xtiles * T;
    cudaMemsetAsync(flags, 0, n * kK6FlagStride * sizeof(uint32_t), stream);
    const dim3 grid((unsigned)(ytiles * xtiles), (unsigned)T, (unsigned)((ts + kBandRows - 1) / kBandRows));
    const bool vec = (ts % 4 == 0) && (W % 4 == 0) && (reinterpret_cast<uintptr_t>(X) % 16 == 0);
    if (vec) k6_nonempty_kernel<true><<<grid, 256, 0, stream>>>(X, T, C, H, W, ts, xtiles, flags);
    else k6_nonempty_kernel<false><<<grid, 256, 0, stream>>>(X, T, C, H, W, ts, xtiles, flags);
    k6_flags_to_bytes_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(flags, n, nonempty);
}

// ---- axb_tile_nonempty_noalloc: the same two kernels with the vote flags in a file-scope device array ----------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, serving loops, harnesses that forbid it).  Every GPU that loads the
// library holds the array (kK6StaticMaxFlags * 32 B = 2 MiB) until the process exits.  All calls on a device share
// it: launch_k6_nonempty zeroes the flags of the call first, and calls that overlap on the GPU race.
static __device__ __align__(256) uint32_t g_k6_static_flags[(size_t)AXB_NOALLOC_MAX_TILE_FRAMES * kK6FlagStride];

// Host code must not pass g_k6_static_flags itself to a kernel (there the name is a host-side stand-in, not a
// device address): the device address is looked up once per device, with that device current.
cudaError_t launch_k6_nonempty_static(const float* X, int T, int C, int H, int W, int ts, uint8_t* nonempty, cudaStream_t stream) {
    constexpr int kMaxDev = 64;
    static std::mutex mu;
    static uint32_t* flags_of_device[kMaxDev] = {};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    uint32_t* flags = nullptr;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (dev >= 0 && dev < kMaxDev) flags = flags_of_device[dev];
        if (!flags) {
            e = cudaGetSymbolAddress((void**)&flags, g_k6_static_flags);
            if (e != cudaSuccess) return e;
            if (dev >= 0 && dev < kMaxDev) flags_of_device[dev] = flags;
        }
    }
    launch_k6_nonempty(X, T, C, H, W, ts, flags, nonempty, stream);     // cudaMemsetAsync of this call's flags + 2 kernels
    return cudaGetLastError();
}

void launch_k6_gather(const float* X, int T, int C, int H, int W, int ts, const int32_t* tile_yx, int ntiles,
                      const int32_t* frames, int B, int tc, int chan_lo, int nch, float* out,
                      unsigned long long* err_flags, int n_sms, cudaStream_t stream) {
    if (launch_k6_gather_tma(X, T, C, H, W, ts, tile_yx, ntiles, frames, B, tc, chan_lo, nch, out, err_flags, n_sms, stream)) return;
    const int chunks = (ts + 15) / 16;
    const long long jobs = (long long)B * ntiles * (2 * tc + 1) * nch * chunks;
    const int grid = (int)std::min<long long>(jobs, (long long)n_sms * 8 * 16);
    const bool vec = (ts % 4 == 0) && (W % 4 == 0) && ((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(out)) % 16 == 0);
    if (vec) k6_gather_kernel<true><<<grid, 256, 0, stream>>>(X, T, C, H, W, ts, tile_yx, ntiles, frames, B, tc, chan_lo, nch, chunks, out, err_flags);
    else k6_gather_kernel<false><<<grid, 256, 0, stream>>>(X, T, C, H, W, ts, tile_yx, ntiles, frames, B, tc, chan_lo, nch, chunks, out, err_flags);
}

}  // namespace axb
