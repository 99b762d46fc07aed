// TEST INFRASTRUCTURE ONLY -- CPU oracle for the AxTrack association hot path.
//
// Nothing in the product path (axtrack_b200/, the C-ABI library) may link, import or call
// this file.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference legs use it, and only as the checker / the timed CPU baseline.
//
// It restates, on the CPU, what the reference computes for the A* cost matrix:
//
//   * axo_astar_weights / axo_astar_mask      -- the forked pyastar2d search
//         (/root/reference/pyastar2d/src/cpp/astar.cpp:46-172): 4-connected A*, L1
//         heuristic, float32 costs, `path_length` carried in the heap entry, and the fork's
//         cut-off "stop with no result as soon as the entry on top of the heap has
//         path_length > max_path_length" (astar.cpp:88-92) evaluated BEFORE the goal test
//         (astar.cpp:99-102).  The heap is std::priority_queue with the same ordering
//         (smallest priority on top, astar.cpp:23-25) so that ties pop in the same order as
//         in the reference.
//   * axo_cost_matrix_astar                   -- the pair loop of
//         AxonDetections._compute_detections_astar_paths / _get_path_between_dets
//         (/root/reference/axtrack/AxonDetections.py:553-579, 600-629) followed by
//         _get_astar_path_distances (AxonDetections.py:717-752): Euclid < max prefilter,
//         one A* per in-range pair on weights where(mask==1, 1, 65536) (AxonDetections.py:598),
//         leaf -> `max if None else path_length`.
//   * axo_cost_matrix_rule                    -- the closed form the CUDA kernel implements
//         (one bounded lexicographic search per SOURCE, DESIGN.md "K2 semantics"); used to
//         cross-check at sizes where the per-pair search is too slow, and itself validated
//         against axo_cost_matrix_astar and against the compiled reference (oracle/_ref).
//
// Parity status: pinned -- see tests/test_oracle_golden.py (the reference's own
// known-answer vectors, pyastar2d/tests/test_astar.py:22-26,29-42,63-78) and
// tests/golden/*.npz (outputs of the reference's Python + compiled pyastar2d run in the
// build container by tests/golden/make_golden.py).
//
// Build: oracle/Makefile -> oracle/liboracle.so
#include <cstdint>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <queue>
#include <vector>
#include <limits>
#include <algorithm>

namespace {

struct HeapEntry {
    float prio;    // g + h  (float32, as in the reference)
    int cell;      // flat index y*W + x
    int plen;      // pixels on the path that produced this entry (start counts as 1)
};
// std::priority_queue keeps the LARGEST element on top; the reference flips the comparison
// on its float priority (astar.cpp:23-25).  Same here, so equal priorities order identically.
struct HeapOrder {
    bool operator()(const HeapEntry& a, const HeapEntry& b) const { return a.prio > b.prio; }
};

// Scratch that survives across pairs: best-known g per cell and the predecessor, reset
// lazily through a touched list (the reference re-allocates and re-fills H*W floats per
// pair, astar.cpp:64,69-72 -- same values, without the O(H*W) cost).
struct Scratch {
    int H = 0, W = 0;
    std::vector<float> g;
    std::vector<int> parent;
    std::vector<int> touched;
    void fit(int h, int w) {
        if (h == H && w == W) return;
        H = h; W = w;
        g.assign((size_t)h * w, std::numeric_limits<float>::infinity());
        parent.assign((size_t)h * w, -1);
        touched.clear();
    }
    void reset() {
        for (int c : touched) { g[c] = std::numeric_limits<float>::infinity(); parent[c] = -1; }
        touched.clear();
    }
};

template <class WeightFn>
int astar_search(Scratch& S, int H, int W, WeightFn weight_of, int start, int goal,
                 int max_len, float* out_cost, int32_t* out_path /* [max_len*2] or null */) {
    S.fit(H, W);
    S.reset();
    const int gy = goal / W, gx = goal % W;

    std::priority_queue<HeapEntry, std::vector<HeapEntry>, HeapOrder> heap;
    S.g[start] = 0.f;
    S.touched.push_back(start);
    heap.push(HeapEntry{0.f, start, 1});

    int found_len = -1;
    while (!heap.empty()) {
        const HeapEntry top = heap.top();
        if (top.plen > max_len) break;                 // fork cut-off, astar.cpp:88-92
        if (top.cell == goal) { found_len = top.plen; break; }   // astar.cpp:99-102
        heap.pop();

        const int y = top.cell / W, x = top.cell % W;
        // neighbour order as in the reference's nbrs[] slots 1,3,4,6: up, left, right, down
        // (astar.cpp:109-116 with diag_ok == false)
        const int nb[4] = {
            y > 0     ? top.cell - W : -1,
            x > 0     ? top.cell - 1 : -1,
            x + 1 < W ? top.cell + 1 : -1,
            y + 1 < H ? top.cell + W : -1,
        };
        for (int k = 0; k < 4; ++k) {
            const int n = nb[k];
            if (n < 0) continue;
            const float cand = S.g[top.cell] + weight_of(n);       // astar.cpp:122
            if (cand < S.g[n]) {                                   // strict, astar.cpp:123
                const float h = (float)(std::abs(n / W - gy) + std::abs(n % W - gx));
                if (S.parent[n] < 0 && S.g[n] == std::numeric_limits<float>::infinity())
                    S.touched.push_back(n);
                S.g[n] = cand;
                S.parent[n] = top.cell;
                heap.push(HeapEntry{cand + h, n, top.plen + 1});
            }
        }
    }
    if (found_len < 0) return -1;
    if (out_cost) *out_cost = S.g[goal];
    if (out_path) {
        int c = goal;
        for (int i = found_len - 1; i >= 0; --i) {      // astar.cpp:146-159
            out_path[2 * i] = c / W;
            out_path[2 * i + 1] = c % W;
            c = S.parent[c];
        }
    }
    return found_len;
}

thread_local Scratch tls_scratch;

inline float mask_weight(uint8_t m) { return m == 1 ? 1.0f : 65536.0f; }   // AxonDetections.py:598

}  // namespace

extern "C" {

// Forked pyastar2d search on an arbitrary float32 weight grid.  Returns the path length
// (pixels, both ends included) or -1 for "None".  out_path (optional) receives (row, col)
// pairs start->goal and must hold 2*max_len (or 2*H*W if smaller) int32.
int axo_astar_weights(const float* weights, int H, int W, int sy, int sx, int gy, int gx,
                      int max_len, float* out_cost, int32_t* out_path) {
    auto wf = [weights](int c) { return weights[c]; };
    return astar_search(tls_scratch, H, W, wf, sy * W + sx, gy * W + gx, max_len, out_cost, out_path);
}

// Same search on the association weights where(mask==1, 1, 65536).
int axo_astar_mask(const uint8_t* mask, int H, int W, int sy, int sx, int gy, int gx,
                   int max_len, float* out_cost) {
    auto wf = [mask](int c) { return mask_weight(mask[c]); };
    return astar_search(tls_scratch, H, W, wf, sy * W + sx, gy * W + gx, max_len, out_cost, nullptr);
}

// Pair loop of the reference.  src_yx / dst_yx are int32 [n,2] as (y, x).
//   path_len[i*n_dst+j]    = L if a path was returned else max_px      (AxonDetections.py:736)
//   unreachable[i*n_dst+j] = 1 if the reference leaf is None (out of Euclid range or A* -> None)
//   wcost[i*n_dst+j]       = float32 g(goal) for returned paths, +inf otherwise
// Returns 0, or the error codes of include/axtrack_b200.h when an in-range pair has an
// out-of-grid end point (astar_wrapper.py:39-45); err_pair receives i*n_dst+j of the first one.
int axo_cost_matrix_astar(const uint8_t* mask, int H, int W,
                          const int32_t* src_yx, int n_src, const int32_t* dst_yx, int n_dst,
                          int max_px, int32_t* path_len, uint8_t* unreachable, float* wcost,
                          int64_t* err_pair) {
    const int64_t r2 = (int64_t)max_px * max_px;
    for (int i = 0; i < n_src; ++i) {
        const int sy = src_yx[2 * i], sx = src_yx[2 * i + 1];
        for (int j = 0; j < n_dst; ++j) {
            const int ty = dst_yx[2 * j], tx = dst_yx[2 * j + 1];
            const size_t o = (size_t)i * n_dst + j;
            path_len[o] = max_px; unreachable[o] = 1;
            if (wcost) wcost[o] = std::numeric_limits<float>::infinity();
            const int64_t dy = (int64_t)sy - ty, dx = (int64_t)sx - tx;
            if (dy * dy + dx * dx >= r2) continue;               // AxonDetections.py:620-628
            if (sy < 0 || sy >= H || sx < 0 || sx >= W) { if (err_pair) *err_pair = (int64_t)o; return -2; }
            if (ty < 0 || ty >= H || tx < 0 || tx >= W) { if (err_pair) *err_pair = (int64_t)o; return -3; }
            float c = 0.f;
            const int L = axo_astar_mask(mask, H, W, sy, sx, ty, tx, max_px, &c);
            if (L >= 0) { path_len[o] = L; unreachable[o] = 0; if (wcost) wcost[o] = c; }
        }
    }
    return 0;
}

// Closed form (DESIGN.md "K2 semantics").  With weights {1, 65536} the exact cost of a path
// is n_fg + 65536*n_bg over the entered pixels, so a path's hop count is a function of its
// cost.  Per source: settle cells in lexicographic (n_bg, n_fg) order ("level" k = n_bg),
// never expanding a cell that sits at hop == max_px (it would be pushed with
// path_length max_px+1, astar.cpp:136).  Let k* be the first level in which some cell is
// reached at hop max_px.  Then for target T with settled (k_T, j_T):
//     L = k_T + j_T + 1 if  L <= max_px  and  k_T <= k*   else None
// because an entry at hop max_px of a lower level always has a smaller A* priority than the
// goal (65536 >> 2*max_px) and is therefore popped -- and triggers the cut-off -- first.
//
// tie_regime (optional, [n_src*n_dst]): 1 where the reference's answer is decided by the pop order of
// equal-priority heap entries (libstdc++ std::priority_queue), which no closed form reproduces: the
// target sits in a level k_T >= 1 and some lower level k expanded cells at hop max_px-1 without
// reaching hop max_px.  Such a cell can push an already-better-known neighbour with path_length
// max_px+1 when it is expanded before that neighbour's own parent -- possible only on an exact
// f-tie (DESIGN.md "Heap-tie corner") -- and that stale entry then triggers the cut-off
// (astar.cpp:88-92) ahead of any goal of a higher level.  Everywhere else the rule is exact.
static int cost_matrix_rule_impl(const uint8_t* mask, int H, int W,
                         const int32_t* src_yx, int n_src, const int32_t* dst_yx, int n_dst,
                         int max_px, int32_t* path_len, uint8_t* unreachable, float* wcost,
                         int64_t* err_pair, uint8_t* tie_regime) {
    const int64_t r2 = (int64_t)max_px * max_px;
    const int R = max_px;                 // cells at hop <= R are looked at
    const int side = 2 * R + 1;
    // window-local state: level (k) and fg-count (j) per cell, -1 = unsettled
    std::vector<int32_t> lev((size_t)side * side), fgc((size_t)side * side);
    std::vector<int> cur, nxt;                       // frontier of the running (k, j)
    std::vector<std::pair<int, int>> seeds, seeds_next;   // (j, cell) for the level / next level

    for (int i = 0; i < n_src; ++i) {
        const int sy = src_yx[2 * i], sx = src_yx[2 * i + 1];
        bool any = false;
        for (int j = 0; j < n_dst; ++j) {
            const size_t o = (size_t)i * n_dst + j;
            path_len[o] = max_px; unreachable[o] = 1;
            if (tie_regime) tie_regime[o] = 0;
            if (wcost) wcost[o] = std::numeric_limits<float>::infinity();
            const int ty = dst_yx[2 * j], tx = dst_yx[2 * j + 1];
            const int64_t dy = (int64_t)sy - ty, dx = (int64_t)sx - tx;
            if (dy * dy + dx * dx >= r2) continue;
            if (sy < 0 || sy >= H || sx < 0 || sx >= W) { if (err_pair) *err_pair = (int64_t)o; return -2; }
            if (ty < 0 || ty >= H || tx < 0 || tx >= W) { if (err_pair) *err_pair = (int64_t)o; return -3; }
            any = true;
        }
        if (!any) continue;

        std::fill(lev.begin(), lev.end(), -1);
        auto wl = [&](int y, int x) { return (y - sy + R) * side + (x - sx + R); };
        auto in_img = [&](int y, int x) { return y >= 0 && y < H && x >= 0 && x < W; };

        seeds.clear(); seeds_next.clear();
        seeds.push_back({0, wl(sy, sx)});
        lev[wl(sy, sx)] = 0; fgc[wl(sy, sx)] = 0;
        int kstar = -1;
        int ktie = -1;                     // first level that expanded hop max_px-1 without reaching hop max_px
        for (int k = 0; k <= R && kstar < 0 && !seeds.empty(); ++k) {
            // seeds are already settled (level k, their j); sorted by j by construction
            size_t sp = 0;
            cur.clear();
            int j = seeds[0].first;
            bool hit_limit = false, at_rm1 = false;
            while (true) {
                while (sp < seeds.size() && seeds[sp].first == j) cur.push_back(seeds[sp++].second);
                if (cur.empty()) {
                    if (sp >= seeds.size()) break;
                    j = seeds[sp].first; continue;
                }
                const int hop = k + j;
                if (hop >= R) { hit_limit = true; /* cells at hop == max_px: never expanded */
                    cur.clear();
                    if (sp >= seeds.size()) break;
                    j = seeds[sp].first; continue; }
                if (hop == R - 1) at_rm1 = true;
                nxt.clear();
                for (int c : cur) {
                    const int y = c / side - R + sy, x = c % side - R + sx;
                    const int ny[4] = {y - 1, y, y, y + 1}, nx[4] = {x, x - 1, x + 1, x};
                    for (int d = 0; d < 4; ++d) {
                        if (!in_img(ny[d], nx[d])) continue;
                        const int n = wl(ny[d], nx[d]);
                        if (lev[n] >= 0) continue;
                        if (mask[(size_t)ny[d] * W + nx[d]] == 1) {
                            lev[n] = k; fgc[n] = j + 1; nxt.push_back(n);
                        } else {
                            lev[n] = k + 1; fgc[n] = j; seeds_next.push_back({j, n});
                        }
                    }
                }
                cur.swap(nxt);
                ++j;
            }
            if (hit_limit) kstar = k;
            else if (at_rm1 && ktie < 0) ktie = k;
            seeds.swap(seeds_next); seeds_next.clear();
        }
        for (int j = 0; j < n_dst; ++j) {
            const size_t o = (size_t)i * n_dst + j;
            const int ty = dst_yx[2 * j], tx = dst_yx[2 * j + 1];
            const int64_t dy = (int64_t)sy - ty, dx = (int64_t)sx - tx;
            if (dy * dy + dx * dx >= r2) continue;
            if (std::abs(ty - sy) + std::abs(tx - sx) > R) continue;
            const int n = wl(ty, tx);
            if (lev[n] < 0) continue;
            const int k = lev[n], jj = fgc[n];
            if (kstar >= 0 && k > kstar) continue;
            if (k + jj + 1 > max_px) continue;
            path_len[o] = k + jj + 1; unreachable[o] = 0;
            if (wcost) wcost[o] = (float)((double)k * 65536.0 + (double)jj);
            if (tie_regime && ktie >= 0 && k > ktie) tie_regime[o] = 1;
        }
    }
    return 0;
}

int axo_cost_matrix_rule(const uint8_t* mask, int H, int W,
                         const int32_t* src_yx, int n_src, const int32_t* dst_yx, int n_dst,
                         int max_px, int32_t* path_len, uint8_t* unreachable, float* wcost,
                         int64_t* err_pair) {
    return cost_matrix_rule_impl(mask, H, W, src_yx, n_src, dst_yx, n_dst, max_px, path_len, unreachable, wcost, err_pair, nullptr);
}

int axo_cost_matrix_rule_ex(const uint8_t* mask, int H, int W,
                            const int32_t* src_yx, int n_src, const int32_t* dst_yx, int n_dst,
                            int max_px, int32_t* path_len, uint8_t* unreachable, float* wcost,
                            int64_t* err_pair, uint8_t* tie_regime) {
    return cost_matrix_rule_impl(mask, H, W, src_yx, n_src, dst_yx, n_dst, max_px, path_len, unreachable, wcost, err_pair, tie_regime);
}

}  // extern "C"
