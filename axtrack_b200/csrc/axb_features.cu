// K5 -- appearance features of the detections of a frame (sm_100a), SURVEY 8f rank 3.
//
// Replaces axtrack/mincostflow_models.py:30-64 (feature_model): for every box (x, y, w, h) the
// histogram of the image crop rows [max(y,0), max(y0+1, y0+h)) x columns [max(x,0), max(x0+1, x0+w))
// (numpy slicing: clipped at the image border) with cv2.calcHist([crop], [0], None, [bins], [lo, hi])
// and cv2.normalize(hist, hist, 0, 1, cv2.NORM_MINMAX).  The arithmetic follows OpenCV 4.x exactly
// (pinned against cv2 4.13 through the reference's own function, tests/golden/feature_hist_*.npz):
//   bin      = floor(double(v) * (bins / (hi - lo)) - lo * (bins / (hi - lo))), counted if 0 <= bin < bins
//              (a NaN pixel is counted in bin 0: observed cv2 behaviour, kept for fidelity)
//   scale    = float(1 / (max - min)) if max - min > DBL_EPSILON else 0      (min, max over the bins)
//   shift    = 0.f - float(double(min) * double(scale))
//   feature  = fmaf(count, scale, shift)
// One CTA per box: shared-memory histogram with integer atomics (order independent, bit-exact),
// coalesced row reads of the crop (the frame stays in L2 across the boxes of a launch).
// Algorithmic bytes per box: w*h*4 B read + bins*4 B written; HBM/L2 bound, launch-latency sized.
#include "axb_common.cuh"
#include <cfloat>

namespace axb {

constexpr int kK5Threads = 128;

__global__ void __launch_bounds__(kK5Threads) k5_feature_hist_kernel(const float* __restrict__ images, int n_frames, int H, int W,
                                                                      const int32_t* __restrict__ boxes,
                                                                      const int32_t* __restrict__ box_frame, int n_boxes,
                                                                      int bins, double a, double b, float* __restrict__ out) {
    __shared__ int hist[AXB_MAX_BINS];
    __shared__ int s_min[kK5Threads / 32], s_max[kK5Threads / 32];
    const int tid = threadIdx.x;
    for (int box = blockIdx.x; box < n_boxes; box += gridDim.x) {
        for (int i = tid; i < bins; i += kK5Threads) hist[i] = 0;
        __syncthreads();
        const int bx = boxes[4 * box], by = boxes[4 * box + 1], bw = boxes[4 * box + 2], bh = boxes[4 * box + 3];
        // a block of frames in one launch: box_frame[box] = which frame of `images` the box lies on (out of range: empty crop)
        const int fr = box_frame ? box_frame[box] : 0;
        const float* const image = images + (size_t)min(max(fr, 0), n_frames - 1) * H * W;
        const bool fr_ok = fr >= 0 && fr < n_frames;
        const int r0 = max(by, 0), c0 = max(bx, 0);                      // mincostflow_models.py:55-58
        const int r1 = min(max(r0 + 1, r0 + bh), H), c1 = min(max(c0 + 1, c0 + bw), W);
        const int cw = c1 - c0, chh = r1 - r0;
        if (cw > 0 && chh > 0 && fr_ok) {
            const int npx = cw * chh;
            for (int i = tid; i < npx; i += kK5Threads) {
                const int r = i / cw, c = i - r * cw;
                const float v = __ldg(image + (size_t)(r0 + r) * W + (c0 + c));
                const double f = v != v ? 0.0 : floor((double)v * a + b);   // cvFloor(*p0 * a + b); cv2 counts NaN in bin 0
                if (f >= 0.0 && f < (double)bins) atomicAdd(&hist[(int)f], 1);
            }
        }
        __syncthreads();
        int mn = 0x7fffffff, mx = -1;
        for (int i = tid; i < bins; i += kK5Threads) { mn = min(mn, hist[i]); mx = max(mx, hist[i]); }
        mn = __reduce_min_sync(0xFFFFFFFFu, mn);
        mx = __reduce_max_sync(0xFFFFFFFFu, mx);
        if ((tid & 31) == 0) { s_min[tid >> 5] = mn; s_max[tid >> 5] = mx; }
        __syncthreads();
        mn = s_min[0]; mx = s_max[0];
        #pragma unroll
        for (int wv = 1; wv < kK5Threads / 32; ++wv) { mn = min(mn, s_min[wv]); mx = max(mx, s_max[wv]); }
        const double smin = (double)mn, smax = (double)mx;
        const float scale = (float)((smax - smin > DBL_EPSILON) ? 1.0 / (smax - smin) : 0.0);
        const float shift = 0.f - (float)(smin * (double)scale);
        for (int i = tid; i < bins; i += kK5Threads) out[(size_t)box * bins + i] = __fmaf_rn((float)hist[i], scale, shift);
        __syncthreads();
    }
}

void launch_k5(const float* images, int n_frames, int H, int W, const int32_t* boxes, const int32_t* box_frame, int n_boxes,
               int bins, float lo, float hi, float* out, cudaStream_t stream) {
    if (n_boxes <= 0) return;
    const double a = (double)bins / ((double)hi - (double)lo), b = -(double)lo * a;      // OpenCV's uniform range
    const int grid = n_boxes < 148 * 16 ? n_boxes : 148 * 16;
    k5_feature_hist_kernel<<<grid, kK5Threads, 0, stream>>>(images, n_frames, H, W, boxes, box_frame, n_boxes, bins, a, b, out);
}

// K7 -- visual similarity of the detections of two frames, SURVEY 8f rank 1 (the vis_sim term of transition_model).
//
// Replaces the double loop axtrack/mincostflow_models.py:106-113:
//     vis_sim[i][j] = 1 - cv2.compareHist(f1_i, f2_j, cv2.HISTCMP_BHATTACHARYYA),  then np.nan_to_num
// with f1 / f2 the float32 feature histograms (K5) of the predecessor / current detections.  OpenCV 4.x computes,
// in double, s1 = sum h1, s2 = sum h2, r = sum sqrt(h1 * h2) and returns sqrt(max(1 - r / sqrt(s1 * s2), 0))
// (1 / sqrt(s1 * s2) replaced by 1 when |s1 * s2| <= FLT_EPSILON).  The sums run in the order of its baseline SSE
// build -- four floats per step, the low pair then the high pair converted to double and added to a TWO-lane
// accumulator (lane 0 = even elements, lane 1 = odd elements, each in index order), lanes added 0 + lane0 + lane1,
// then the len % 4 tail one by one -- which is what makes the result bit-identical to cv2 4.13 (pinned by
// tests/golden/vis_sim_*.npz, generated through the reference's own transition_model).
// The value only enters the cost where the A* distance is below max_px (costs[inf_dist] = inf,
// mincostflow_models.py:118): entries with path_len == max_px are skipped and written as 0.
// One thread per (i, j): the histograms of 4 predecessors and 32 current detections staged in shared memory
// (rows padded to an odd word count: conflict-free), 2 * ceil(bins / 2) dependent double adds per entry.
constexpr int kK7Src = 4, kK7Dst = 32;

__device__ __forceinline__ double lane_sum(const float* __restrict__ h, int bins) {   // s = sum h in cv2's order
    double l0 = 0.0, l1 = 0.0;
    const int body = bins & ~3;
    for (int k = 0; k < body; k += 2) { l0 = __dadd_rn(l0, (double)h[k]); l1 = __dadd_rn(l1, (double)h[k + 1]); }
    double s = __dadd_rn(__dadd_rn(0.0, l0), l1);
    for (int k = body; k < bins; ++k) s = __dadd_rn(s, (double)h[k]);
    return s;
}

__global__ void __launch_bounds__(kK7Src * kK7Dst) k7_vis_sim_kernel(const float* __restrict__ f_src, int n_src,
                                                                     const float* __restrict__ f_dst, int n_dst, int bins,
                                                                     const int32_t* __restrict__ path_len, int max_px,
                                                                     double* __restrict__ vis) {
    extern __shared__ float sh[];                  // [kK7Src][pitch] | [kK7Dst][pitch]
    const int pitch = bins | 1;
    float* const hs = sh;
    float* const hd = sh + kK7Src * pitch;
    const int i0 = blockIdx.y * kK7Src, j0 = blockIdx.x * kK7Dst;
    for (int t = threadIdx.x; t < kK7Src * bins; t += blockDim.x) {
        const int r = t / bins, k = t - r * bins;
        hs[r * pitch + k] = i0 + r < n_src ? f_src[(size_t)(i0 + r) * bins + k] : 0.f;
    }
    for (int t = threadIdx.x; t < kK7Dst * bins; t += blockDim.x) {
        const int r = t / bins, k = t - r * bins;
        hd[r * pitch + k] = j0 + r < n_dst ? f_dst[(size_t)(j0 + r) * bins + k] : 0.f;
    }
    __syncthreads();
    const int i = i0 + (int)(threadIdx.x / kK7Dst), j = j0 + (int)(threadIdx.x % kK7Dst);
    if (i >= n_src || j >= n_dst) return;
    const size_t o = (size_t)i * n_dst + j;
    if (path_len && path_len[o] == max_px) { vis[o] = 0.0; return; }
    const float* const h1 = hs + (threadIdx.x / kK7Dst) * pitch;
    const float* const h2 = hd + (threadIdx.x % kK7Dst) * pitch;
    double l0 = 0.0, l1 = 0.0;
    const int body = bins & ~3;
    for (int k = 0; k < body; k += 2) {
        l0 = __dadd_rn(l0, __dsqrt_rn(__dmul_rn((double)h1[k], (double)h2[k])));
        l1 = __dadd_rn(l1, __dsqrt_rn(__dmul_rn((double)h1[k + 1], (double)h2[k + 1])));
    }
    double r = __dadd_rn(__dadd_rn(0.0, l0), l1);
    for (int k = body; k < bins; ++k) r = __dadd_rn(r, __dsqrt_rn(__dmul_rn((double)h1[k], (double)h2[k])));
    double s = __dmul_rn(lane_sum(h1, bins), lane_sum(h2, bins));
    s = fabs(s) > (double)FLT_EPSILON ? __ddiv_rn(1.0, __dsqrt_rn(s)) : 1.0;
    // std::max(1. - result * s1, 0.) keeps a NaN (a slightly negative normalised bin makes sqrt(h1 * h2) NaN in cv2 too);
    // fmax() would drop it
    const double x = __dadd_rn(1.0, -__dmul_rn(r, s));
    const double d = __dsqrt_rn(x < 0.0 ? 0.0 : x);
    double v = __dadd_rn(1.0, -d);
    if (v != v) v = 0.0;                                              // np.nan_to_num
    else if (isinf(v)) v = v > 0 ? DBL_MAX : -DBL_MAX;
    vis[o] = v;
}

void launch_k7(const float* f_src, int n_src, const float* f_dst, int n_dst, int bins, const int32_t* path_len, int max_px,
               double* vis, cudaStream_t stream) {
    if (n_src <= 0 || n_dst <= 0) return;
    const dim3 grid((n_dst + kK7Dst - 1) / kK7Dst, (n_src + kK7Src - 1) / kK7Src);
    const size_t smem = (size_t)(kK7Src + kK7Dst) * (bins | 1) * sizeof(float);
    k7_vis_sim_kernel<<<grid, kK7Src * kK7Dst, smem, stream>>>(f_src, n_src, f_dst, n_dst, bins, path_len, max_px, vis);
}

}  // namespace axb
