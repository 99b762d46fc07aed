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

__global__ void __launch_bounds__(kK5Threads) k5_feature_hist_kernel(const float* __restrict__ image, int H, int W,
                                                                      const int32_t* __restrict__ boxes, int n_boxes,
                                                                      int bins, double a, double b, float* __restrict__ out) {
    __shared__ int hist[AXB_MAX_BINS];
    __shared__ int s_min[kK5Threads / 32], s_max[kK5Threads / 32];
    const int tid = threadIdx.x;
    for (int box = blockIdx.x; box < n_boxes; box += gridDim.x) {
        for (int i = tid; i < bins; i += kK5Threads) hist[i] = 0;
        __syncthreads();
        const int bx = boxes[4 * box], by = boxes[4 * box + 1], bw = boxes[4 * box + 2], bh = boxes[4 * box + 3];
        const int r0 = max(by, 0), c0 = max(bx, 0);                      // mincostflow_models.py:55-58
        const int r1 = min(max(r0 + 1, r0 + bh), H), c1 = min(max(c0 + 1, c0 + bw), W);
        const int cw = c1 - c0, chh = r1 - r0;
        if (cw > 0 && chh > 0) {
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

void launch_k5(const float* image, int H, int W, const int32_t* boxes, int n_boxes, int bins, float lo, float hi,
               float* out, cudaStream_t stream) {
    if (n_boxes <= 0) return;
    const double a = (double)bins / ((double)hi - (double)lo), b = -(double)lo * a;      // OpenCV's uniform range
    const int grid = n_boxes < 148 * 8 ? n_boxes : 148 * 8;
    k5_feature_hist_kernel<<<grid, kK5Threads, 0, stream>>>(image, H, W, boxes, n_boxes, bins, a, b, out);
}

}  // namespace axb
