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

void launch_k6_gather(const float* X, int T, int C, int H, int W, int ts, const int32_t* tile_yx, int ntiles,
                      const int32_t* frames, int B, int tc, int chan_lo, int nch, float* out,
                      unsigned long long* err_flags, int n_sms, cudaStream_t stream) {
    const int chunks = (ts + 15) / 16;
    const long long jobs = (long long)B * ntiles * (2 * tc + 1) * nch * chunks;
    const int grid = (int)std::min<long long>(jobs, (long long)n_sms * 8 * 16);
    const bool vec = (ts % 4 == 0) && (W % 4 == 0) && ((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(out)) % 16 == 0);
    if (vec) k6_gather_kernel<true><<<grid, 256, 0, stream>>>(X, T, C, H, W, ts, tile_yx, ntiles, frames, B, tc, chan_lo, nch, chunks, out, err_flags);
    else k6_gather_kernel<false><<<grid, 256, 0, stream>>>(X, T, C, H, W, ts, tile_yx, ntiles, frames, B, tc, chan_lo, nch, chunks, out, err_flags);
}

}  // namespace axb
