// K3 -- transition cost of every (track, detection) pair from its A* path length (sm_100a).
//
// Replaces the arithmetic of mincostflow_models.transition_model
// (axtrack/mincostflow_models.py:100-119) and the per-arc cost of libmot's graph builder
// (libmot/libmot/data_association/offline_mincostflow.py:192-198) for the deployed
// MCF_VIS_SIM_WEIGHT = 0.  A path length is an integer in [0, max_px] and the time gap is one of
// max_num_misses + 1 values, so the float64 cost
//     -log((1 - w) * (1 - d / max_px) * miss_rate^(gap - 1) + w * vis_sim + 1e-6),  inf where d == max_px
// and the arc cost int(1e5 * cost) take at most (max_px + 1) x n_gaps different values.  The caller
// evaluates the reference's own numpy expression once per (gap, d) on the host -- bit-identical by
// construction, no device libm in the result -- and this kernel gathers: 4 B read, 8 (+8) B
// written per matrix entry, HBM bound.  One CTA column per frame pair (blockIdx.y), the pair's
// entries strided over blockIdx.x; the two tables of the pair's gap are staged in shared memory.
#include "axb_common.cuh"

namespace axb {

__global__ void __launch_bounds__(256) k3_transition_kernel(const int32_t* __restrict__ path_len,
                                                            const axb_pair_desc* __restrict__ pairs, int n_pairs,
                                                            const double* __restrict__ cost_tab, const long long* __restrict__ arc_tab,
                                                            int n_gaps, int max_px, double* __restrict__ cost,
                                                            long long* __restrict__ arc, unsigned long long* err_flags) {
    extern __shared__ __align__(16) unsigned char k3_smem[];
    double* const s_cost = reinterpret_cast<double*>(k3_smem);
    long long* const s_arc = reinterpret_cast<long long*>(s_cost + (max_px + 1));
    for (int pi = blockIdx.y; pi < n_pairs; pi += gridDim.y) {
        const axb_pair_desc pd = pairs[pi];
        const long long n = (long long)max(pd.n_src, 0) * max(pd.n_dst, 0);
        const int g = pd.time_gap - 1;
        __syncthreads();
        if (n == 0) continue;
        if (g < 0 || g >= n_gaps) {                       // CTA-uniform
            if (threadIdx.x == 0 && blockIdx.x == 0) atomicOr(err_flags, 8ull);
            continue;
        }
        for (int i = threadIdx.x; i <= max_px; i += blockDim.x) {
            s_cost[i] = cost_tab[(size_t)g * (max_px + 1) + i];
            if (arc) s_arc[i] = arc_tab[(size_t)g * (max_px + 1) + i];
        }
        __syncthreads();
        const int32_t* const in = path_len + pd.out_off;
        double* const oc = cost + pd.out_off;
        long long* const oa = arc ? arc + pd.out_off : nullptr;
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
            int d = in[i];
            if (d < 0 || d > max_px) { atomicOr(err_flags, 8ull); d = max_px; }
            oc[i] = s_cost[d];
            if (oa) oa[i] = s_arc[d];
        }
    }
}

void launch_k3(const int32_t* path_len, const axb_pair_desc* pairs, int n_pairs, const double* cost_tab,
               const long long* arc_tab, int n_gaps, int max_px, double* cost, long long* arc,
               unsigned long long* err_flags, int max_elems_per_pair, cudaStream_t stream) {
    const int gx = max(1, min((max_elems_per_pair + 256 * 8 - 1) / (256 * 8), 64));
    const dim3 grid(gx, (unsigned)min(n_pairs, 65535));
    k3_transition_kernel<<<grid, 256, (size_t)(max_px + 1) * 16, stream>>>(path_len, pairs, n_pairs, cost_tab, arc_tab, n_gaps,
                                                                          max_px, cost, arc, err_flags);
}

}  // namespace axb
