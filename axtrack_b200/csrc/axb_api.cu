// C ABI of libaxtrack_b200.so (include/axtrack_b200.h): context, scratch management, launches.
// No CPU fallback anywhere: every compute entry point needs a CUDA device.
#include "axb_common.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

struct axb_ctx {
    int device = 0;
    int num_sms = 0;
    std::string err;
    int64_t launches = 0;
    cudaStream_t own_stream = nullptr;
    // chunked host entry (axb_cost_matrix_batch_host): a second scratch set + stream so that the searches of chunk
    // k+1 fill the SMs while chunk k drains, and a copy stream that takes finished chunks to the host meanwhile
    axb_ctx* alt = nullptr;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_in = nullptr, ev_chunk[2] = {nullptr, nullptr};

    // device-side sticky status: [0] first input error key (atomicMin), [1] flag bits (atomicOr)
    unsigned long long* status_dev = nullptr;
    unsigned long long* status_host = nullptr;   // pinned

    // K2
    uint32_t* k2_fast_scratch = nullptr;   int k2_fast_ctas = 0;
    uint32_t* k2_gen_scratch = nullptr;    int k2_gen_ctas = 0;
    uint32_t* k2_wide_scratch = nullptr;   // max_px > 511: per general CTA kGenScratchWordsWide (allocated on first use)
    uint8_t* trace_par_wide = nullptr;
    int* counters = nullptr;
    int* prefix_dev = nullptr;        int prefix_cap = 0;   int last_n_pairs = 0;   // [work | all | next] x prefix_cap
    int* gen_list = nullptr;          size_t gen_list_bytes = 0;
    int* dense_list = nullptr;        size_t dense_list_bytes = 0;      // [2][total_src]: windows the first / the directed pass could not hold
    int* dir_list = nullptr;          size_t dir_list_bytes = 0;        // work items that need a directed search after the cover pass
    uint32_t* item_bits = nullptr;    size_t item_bits_bytes = 0;
    int* comp_label = nullptr;        size_t comp_label_bytes = 0;      // mask components (axb_components.cu)
    int4* comp_ext = nullptr;         size_t comp_ext_bytes = 0;
    int dense_stride = 0;
    // axb_prepare_masks: tiles / components of these mask bits are ready for the next axb_cost_matrix_batch (one shot)
    const uint32_t* prep_bits = nullptr;   int prep_n_masks = 0, prep_H = 0, prep_W = 0;   bool prep_components = false;
    int* early_list = nullptr;        size_t early_list_bytes = 0;      // in-grid off-mask detections: the all-levels kernel runs them beside the level-0 kernels
    cudaStream_t side_stream = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    uint8_t* trace_par = nullptr;                                      // K4: per general CTA entry directions
    int* trace_iota = nullptr;        size_t trace_iota_bytes = 0;
    uint32_t* k2_tiles = nullptr;     size_t k2_tiles_bytes = 0;       // 8x4-pixel tile words of the masks of the current call
    // K1
    uint8_t* k1_scratch = nullptr;    size_t k1_scratch_bytes = 0;
    // staging for the single-pair and host entry points
    uint8_t* st_mask = nullptr;       size_t st_mask_bytes = 0;
    uint32_t* st_bits = nullptr;      size_t st_bits_bytes = 0;
    int32_t* st_dets = nullptr;       size_t st_dets_bytes = 0;
    axb_pair_desc* st_pairs = nullptr; size_t st_pairs_bytes = 0;
    int32_t* st_plen = nullptr;       size_t st_plen_bytes = 0;
    uint8_t* st_unr = nullptr;        size_t st_unr_bytes = 0;
    float* st_wc = nullptr;           size_t st_wc_bytes = 0;
    // staging for axb_decode_nms_host: one arena
    uint8_t* st_k1 = nullptr;         size_t st_k1_bytes = 0;
    // K6: one flag word per 32-byte sector while the bands of a tile vote
    uint32_t* k6_flags = nullptr;     size_t k6_flags_bytes = 0;
};

namespace {

bool cuda_ok(axb_ctx* c, cudaError_t e, const char* what) {
    if (e == cudaSuccess) return true;
    char buf[512];
    snprintf(buf, sizeof buf, "%s: %s", what, cudaGetErrorString(e));
    if (c) c->err = buf;
    return false;
}
#define AXB_CUDA(c, call) do { if (!cuda_ok((c), (call), #call)) return AXB_ERR_CUDA; } while (0)

int bad_arg(axb_ctx* c, const char* msg) { if (c) c->err = msg; return AXB_ERR_BAD_ARG; }

// Every entry point runs on the ctx's device and leaves the calling thread's current device as it
// found it (the caller may be a framework that tracks its own current device).
struct DeviceGuard {
    int prev = -1;
    cudaError_t err = cudaSuccess;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != dev) err = cudaSetDevice(dev); else prev = -1;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
#define AXB_ON_DEVICE(c) DeviceGuard axb_guard_((c)->device); AXB_CUDA((c), axb_guard_.err)

template <class T>
int grow(axb_ctx* c, T** ptr, size_t* have, size_t need) {
    if (need <= *have && *ptr) return AXB_OK;
    if (*ptr) { AXB_CUDA(c, cudaFree(*ptr)); *ptr = nullptr; *have = 0; }
    size_t bytes = std::max<size_t>(need + need / 4, 256);
    AXB_CUDA(c, cudaMalloc((void**)ptr, bytes));
    *have = bytes;
    return AXB_OK;
}

int ensure_k2(axb_ctx* c, int n_pairs, int total_src) {
    if (!c->k2_fast_scratch) {
        c->k2_fast_ctas = c->num_sms * std::max(axb::kFastCtasPerSm, axb::kSparseCtasPerSm);
        AXB_CUDA(c, cudaMalloc((void**)&c->k2_fast_scratch, (size_t)c->k2_fast_ctas * axb::kFastScratchWords * sizeof(uint32_t)));
    }
    if (!c->k2_gen_scratch) {
        c->k2_gen_ctas = c->num_sms;
        AXB_CUDA(c, cudaMalloc((void**)&c->k2_gen_scratch, (size_t)c->k2_gen_ctas * axb::kGenScratchWords * sizeof(uint32_t)));
    }
    if (!c->counters) AXB_CUDA(c, cudaMalloc((void**)&c->counters, 256));
    if (n_pairs + 1 > c->prefix_cap) {
        if (c->prefix_dev) AXB_CUDA(c, cudaFree(c->prefix_dev));
        c->prefix_dev = nullptr;
        c->prefix_cap = std::max(n_pairs + 1, 1024) * 2;
        AXB_CUDA(c, cudaMalloc((void**)&c->prefix_dev, axb::kPrepArrays * (size_t)c->prefix_cap * sizeof(int)));
    }
    const size_t n_items = (size_t)std::max(total_src, 1);
    int rc = grow(c, &c->dense_list, &c->dense_list_bytes, 2 * n_items * sizeof(int));
    if (rc != AXB_OK) return rc;
    c->dense_stride = (int)(c->dense_list_bytes / (2 * sizeof(int)));
    if ((rc = grow(c, &c->dir_list, &c->dir_list_bytes, n_items * sizeof(int))) != AXB_OK) return rc;
    if ((rc = grow(c, &c->early_list, &c->early_list_bytes, n_items * sizeof(int))) != AXB_OK) return rc;
    if (!c->side_stream) {
        AXB_CUDA(c, cudaStreamCreateWithFlags(&c->side_stream, cudaStreamNonBlocking));
        AXB_CUDA(c, cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
        AXB_CUDA(c, cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
    }
    if ((rc = grow(c, &c->item_bits, &c->item_bits_bytes, (n_items + 31) / 32 * sizeof(uint32_t))) != AXB_OK) return rc;
    return grow(c, &c->gen_list, &c->gen_list_bytes, n_items * sizeof(int));
}

axb::K2Prep prep_arrays(axb_ctx* c) {
    axb::K2Prep o;
    int* const b = c->prefix_dev;
    const size_t n = (size_t)c->prefix_cap;
    o.work_prefix = b; o.all_prefix = b + n; o.next_same = b + 2 * n; o.p1_prefix = b + 3 * n; o.rev_head = b + 4 * n;
    o.rev_next = b + 5 * n; o.lead_of = b + 6 * n; o.dst_head = b + 7 * n; o.state = b + 8 * n;
    o.item_bits = c->item_bits;
    return o;
}

void set_prep(axb::K2Params& p, const axb::K2Prep& o, axb_ctx* c) {
    p.work_prefix = o.work_prefix; p.all_prefix = o.all_prefix; p.next_same = o.next_same; p.p1_prefix = o.p1_prefix;
    p.rev_head = o.rev_head; p.rev_next = o.rev_next; p.lead_of = o.lead_of; p.item_bits = o.item_bits;
    p.dir_list = c->dir_list; p.gen_list = c->gen_list; p.dense_list = c->dense_list; p.counters = c->counters;
    p.gen_src = c->gen_list; p.gen_len_ctr = axb::kCtrGenLen; p.gen_ticket_ctr = axb::kCtrGenTicket;
}

// where a fast-kernel launch takes its items from / puts the windows it cannot hold
void set_source(axb::K2Params& p, const int* list, int len_ctr, int ticket_ctr, int* dense_out, int dense_out_ctr, int reverse) {
    p.src_list = list; p.src_len_ctr = len_ctr; p.src_ticket_ctr = ticket_ctr;
    p.dense_out = dense_out; p.dense_out_ctr = dense_out_ctr; p.reverse = reverse;
}

}  // namespace

extern "C" {

const char* axb_version(void) { return "axtrack_b200 0.1 (sm_100a)"; }

int axb_create(int device, axb_ctx** out) {
    if (!out) return AXB_ERR_BAD_ARG;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) return AXB_ERR_CUDA;
    axb_ctx* c = new axb_ctx();
    c->device = device;
    DeviceGuard guard(device);          // the caller's current device is restored on every return path
    if (guard.err != cudaSuccess) { delete c; return AXB_ERR_CUDA; }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { delete c; return AXB_ERR_CUDA; }
    c->num_sms = prop.multiProcessorCount;
    if (cudaMalloc((void**)&c->status_dev, 2 * sizeof(unsigned long long)) != cudaSuccess ||
        cudaMallocHost((void**)&c->status_host, 2 * sizeof(unsigned long long)) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete c; return AXB_ERR_CUDA;
    }
    c->status_host[0] = ~0ull; c->status_host[1] = 0ull;
    if (cudaMemcpy(c->status_dev, c->status_host, 16, cudaMemcpyHostToDevice) != cudaSuccess) { delete c; return AXB_ERR_CUDA; }
    *out = c;
    return AXB_OK;
}

void axb_destroy(axb_ctx* c) {
    if (!c) return;
    DeviceGuard guard(c->device);
    cudaDeviceSynchronize();
    if (c->alt) { axb_destroy(c->alt); c->alt = nullptr; }
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->side_stream) cudaStreamDestroy(c->side_stream);
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    if (c->ev_in) cudaEventDestroy(c->ev_in);
    for (int i = 0; i < 2; ++i) if (c->ev_chunk[i]) cudaEventDestroy(c->ev_chunk[i]);
    cudaFree(c->status_dev); cudaFreeHost(c->status_host);
    cudaFree(c->k2_fast_scratch); cudaFree(c->k2_gen_scratch); cudaFree(c->counters); cudaFree(c->prefix_dev);
    cudaFree(c->trace_par); cudaFree(c->trace_iota); cudaFree(c->k2_wide_scratch); cudaFree(c->trace_par_wide);
    cudaFree(c->gen_list); cudaFree(c->dense_list); cudaFree(c->dir_list); cudaFree(c->early_list); cudaFree(c->item_bits); cudaFree(c->comp_label); cudaFree(c->comp_ext); cudaFree(c->k2_tiles); cudaFree(c->k1_scratch);
    cudaFree(c->st_mask); cudaFree(c->st_bits); cudaFree(c->st_dets); cudaFree(c->st_pairs);
    cudaFree(c->st_plen); cudaFree(c->st_unr); cudaFree(c->st_wc); cudaFree(c->st_k1); cudaFree(c->k6_flags);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
}

const char* axb_last_error(const axb_ctx* c) { return c ? c->err.c_str() : "no context"; }
int64_t axb_launch_count(const axb_ctx* c) { return c ? c->launches : 0; }
int axb_mask_words(int W) { return (W + 31) / 32; }

int axb_pack_mask(axb_ctx* c, const uint8_t* mask_dev, int n_masks, int H, int W,
                  uint32_t* mask_bits_dev, void* stream) {
    if (!c) return AXB_ERR_BAD_ARG;
    if (!mask_dev || !mask_bits_dev || n_masks < 0 || H <= 0 || W <= 0) return bad_arg(c, "axb_pack_mask: bad argument");
    AXB_ON_DEVICE(c);
    axb::launch_pack_mask_u8(mask_dev, n_masks, H, W, mask_bits_dev, (cudaStream_t)stream);
    c->launches += 1;
    AXB_CUDA(c, cudaGetLastError());
    return AXB_OK;
}

int axb_pack_weights(axb_ctx* c, const float* weights_dev, int n_masks, int H, int W,
                     uint32_t* mask_bits_dev, void* stream) {
    if (!c) return AXB_ERR_BAD_ARG;
    if (!weights_dev || !mask_bits_dev || n_masks < 0 || H <= 0 || W <= 0) return bad_arg(c, "axb_pack_weights: bad argument");
    AXB_ON_DEVICE(c);
    axb::launch_pack_mask_f32(weights_dev, n_masks, H, W, mask_bits_dev, c->status_dev + 1, (cudaStream_t)stream);
    c->launches += 1;
    AXB_CUDA(c, cudaGetLastError());
    return AXB_OK;
}

}  // extern "C"

namespace {
// Device-side sticky status -> host, cleared.  Synchronises `s`.
int status_fetch(axb_ctx* c, cudaStream_t s, unsigned long long* key, unsigned long long* flags) {
    AXB_CUDA(c, cudaMemcpyAsync(c->status_host, c->status_dev, 16, cudaMemcpyDeviceToHost, s));
    AXB_CUDA(c, cudaStreamSynchronize(s));
    *key = c->status_host[0]; *flags = c->status_host[1];
    if (*key != ~0ull || *flags) {
        c->status_host[0] = ~0ull; c->status_host[1] = 0ull;
        AXB_CUDA(c, cudaMemcpyAsync(c->status_dev, c->status_host, 16, cudaMemcpyHostToDevice, s));
        AXB_CUDA(c, cudaStreamSynchronize(s));
    }
    return AXB_OK;
}

// (key, flags) -> return code; for an out-of-grid end point info = (pair, source, target), the work item of the key
// mapped to its pair by `locate(item, &pair, &source)`.
template <class Locate>
int status_decode(axb_ctx* c, unsigned long long key, unsigned long long flags, int64_t info[3], Locate locate) {
    if (info) { info[0] = info[1] = info[2] = -1; }
    int rc = AXB_OK;
    if (flags & 1ull) rc = AXB_ERR_BAD_WEIGHT;                       // checked first by the reference (astar_wrapper.py:35)
    else if (key != ~0ull) {
        const int code = (int)(key & 3ull);
        const long long j = (long long)((key >> 2) & 0xFFFFFFull);
        const long long item = (long long)(key >> 26);
        rc = code == 1 ? AXB_ERR_OOB_START : AXB_ERR_OOB_GOAL;
        if (info) { info[2] = j; info[0] = -1; info[1] = item; locate(item, &info[0], &info[1]); }
    }
    else if (flags & 2ull) rc = AXB_ERR_UNSUPPORTED_WEIGHT;
    else if (flags & 4ull) rc = AXB_ERR_CAPACITY;
    else if (flags & 8ull) { rc = AXB_ERR_BAD_ARG; c->err = "axb_transition_cost_batch: path length or time gap outside its table"; }
    else if (flags & 16ull) { rc = AXB_ERR_BAD_ARG; c->err = "axb_gather_frametiles: a frame's temporal context leaves the timelapse"; }
    else if (flags & 32ull) { rc = AXB_ERR_BAD_ARG; c->err = "axb_cost_matrix_batch: a pair descriptor points outside the detections / masks / outputs, or total_src is smaller than the pairs' sources"; }
    return rc;
}

// What K2 derives from the masks alone: the chunks of one host call share them (computed once on the first ctx).
struct MaskProducts {
    const uint32_t* tiles;
    const int* comp_label;      // NULL: no cover
    const int4* comp_ext;
};

bool cover_possible(int n_masks, int H, int W) {
    static const bool no_reverse = getenv("AXB_K2_NO_REVERSE") != nullptr;
    const size_t px = (size_t)n_masks * H * W;
    return !no_reverse && (size_t)H * W < (1ull << 31) && px * 20 <= (4ull << 30);
}

// tiled masks + (when the cover can be used) their components, on stream s into the ctx's own buffers
int mask_products(axb_ctx* c, const uint32_t* mask_bits_dev, int n_masks, int H, int W, bool want_components, cudaStream_t s,
                  MaskProducts* out) {
    const int Wp = (W + 31) / 32, TH = (H + 3) / 4;
    int rc = grow(c, &c->k2_tiles, &c->k2_tiles_bytes, (size_t)n_masks * TH * Wp * 4 * sizeof(uint32_t));
    if (rc != AXB_OK) return rc;
    axb::launch_k2_tiles(mask_bits_dev, n_masks, H, Wp, c->k2_tiles, s);
    c->launches += 1;
    out->tiles = c->k2_tiles; out->comp_label = nullptr; out->comp_ext = nullptr;
    if (want_components) {
        const size_t px = (size_t)n_masks * H * W;
        if ((rc = grow(c, &c->comp_label, &c->comp_label_bytes, px * sizeof(int))) != AXB_OK) return rc;
        if ((rc = grow(c, &c->comp_ext, &c->comp_ext_bytes, px * sizeof(int4))) != AXB_OK) return rc;
        axb::launch_k2_components(mask_bits_dev, n_masks, H, W, Wp, c->comp_label, c->comp_ext, s);
        c->launches += 4;
        out->comp_label = c->comp_label; out->comp_ext = c->comp_ext;
    }
    return AXB_OK;
}

int cost_matrix_batch_impl(axb_ctx* c, const uint32_t* mask_bits_dev, int n_masks, int H, int W,
                           const int32_t* dets_yx_dev, int n_dets, const axb_pair_desc* pairs_dev, int n_pairs,
                           int total_src, int max_px,
                           int32_t* path_len_dev, uint8_t* unreachable_dev, float* wcost_dev, int64_t out_capacity,
                           long long orig_base, void* stream, const MaskProducts* shared = nullptr) {
    if (!c) return AXB_ERR_BAD_ARG;
    if (n_pairs < 0 || total_src < 0 || H <= 0 || W <= 0 || n_masks <= 0 || n_dets < 0 || out_capacity < 0)
        return bad_arg(c, "axb_cost_matrix_batch: bad size");
    if (max_px < 1 || max_px > AXB_MAX_PX) return bad_arg(c, "axb_cost_matrix_batch: max_px outside [1, AXB_MAX_PX]");
    if (n_pairs == 0 || total_src == 0) return AXB_OK;
    if (!mask_bits_dev || !dets_yx_dev || !pairs_dev || !path_len_dev || !unreachable_dev)
        return bad_arg(c, "axb_cost_matrix_batch: NULL pointer");
    AXB_ON_DEVICE(c);
    cudaStream_t s = (cudaStream_t)stream;
    int rc = ensure_k2(c, n_pairs, total_src);
    if (rc != AXB_OK) return rc;
    const int Wp = (W + 31) / 32, TH = (H + 3) / 4;
    const bool wide = max_px > axb::kFastMaxPx;
    if (wide && !c->k2_wide_scratch)
        AXB_CUDA(c, cudaMalloc((void**)&c->k2_wide_scratch, (size_t)c->k2_gen_ctas * axb::kWideCtasPerSm * axb::kGenScratchWordsWide * sizeof(uint32_t)));
    // Cover (axb_search.cu k2_prep_kernel): pairs whose source frame is not searched are answered from their target
    // side; that needs the components of the masks.  Off for a single pair (nothing to share), for label arrays
    // beyond 4 GiB and with AXB_K2_NO_REVERSE (A/B timing, tests).
    MaskProducts own;
    if (!shared && c->prep_bits == mask_bits_dev && c->prep_n_masks == n_masks && c->prep_H == H && c->prep_W == W) {
        own.tiles = c->k2_tiles;                 // axb_prepare_masks ran on this stream for exactly these bits
        own.comp_label = c->prep_components ? c->comp_label : nullptr;
        own.comp_ext = c->prep_components ? c->comp_ext : nullptr;
        shared = &own;
    }
    c->prep_bits = nullptr;                      // one shot
    if (!shared) {
        if ((rc = mask_products(c, mask_bits_dev, n_masks, H, W, n_pairs >= 2 && cover_possible(n_masks, H, W), s, &own)) != AXB_OK) return rc;
        shared = &own;
    }
    const bool cover = n_pairs >= 2 && shared->comp_label != nullptr;
    const axb::K2Prep prep = prep_arrays(c);
    axb::launch_k2_prep(pairs_dev, n_pairs, prep, c->counters, n_masks, n_dets, (long long)out_capacity, total_src, cover ? 1 : 0,
                        c->status_dev + 1, s);
    c->last_n_pairs = n_pairs;
    axb::K2Params p;
    memset(&p, 0, sizeof p);
    p.mask_bits = mask_bits_dev; p.H = H; p.W = W; p.Wp = Wp;
    p.mask_tiles = shared->tiles; p.TH = TH; p.TWp = 4 * Wp;
    p.dets_yx = dets_yx_dev; p.pairs = pairs_dev; p.n_pairs = n_pairs;
    set_prep(p, prep, c);
    p.max_px = max_px; p.hash_smem = 0; p.sparse_words = axb::kSparseWords; p.path_len = path_len_dev; p.unreachable = unreachable_dev; p.wcost = wcost_dev;
    p.fast_scratch = c->k2_fast_scratch; p.gen_scratch = wide ? c->k2_wide_scratch : c->k2_gen_scratch;
    p.err_key = c->status_dev;
    p.trace_par = nullptr; p.trace_yx = nullptr; p.trace_len = nullptr;
    p.orig_base = orig_base;
    p.comp_label = cover ? shared->comp_label : nullptr; p.comp_ext = cover ? shared->comp_ext : nullptr;
    // In-grid off-mask detections (a few per cent on real data) need the all-levels kernel, one search latency each:
    // it takes them on a side stream while the level-0 kernels run (k2_classify_kernel).  AXB_K2_NO_EARLY: A/B timing.
    static const bool no_early = getenv("AXB_K2_NO_EARLY") != nullptr;
    const bool early = !no_early && n_pairs >= 2;
    p.early = early ? 1 : 0;
    const int gen_grid = std::max(1, std::min(c->k2_gen_ctas * (wide ? axb::kWideCtasPerSm : 1), total_src));
    if (cover) { axb::launch_k2_fill_reverse(p, prep.dst_head, prep.state, s); c->launches += 1; }     // before the fork: the side stream overwrites rows of it
    if (early) {
        axb::launch_k2_classify(p, c->early_list, total_src, s);
        AXB_CUDA(c, cudaEventRecord(c->ev_fork, s));
        AXB_CUDA(c, cudaStreamWaitEvent(c->side_stream, c->ev_fork, 0));
        axb::K2Params pe = p;
        pe.gen_src = c->early_list; pe.gen_len_ctr = axb::kCtrEarlyLen; pe.gen_ticket_ctr = axb::kCtrEarlyTicket;
        axb::launch_k2_general(pe, gen_grid, c->side_stream);
        AXB_CUDA(c, cudaEventRecord(c->ev_join, c->side_stream));
        c->launches += 2;
    }
    int* const dense_a = c->dense_list;
    int* const dense_b = c->dense_list + c->dense_stride;
    // The launches of a call: (1) level-0 search from every detection of the searched sets, with the reverse members of
    // their chains; (2) the windows (1) could not hold; (3) directed searches for the detections of the other sets that
    // (1) and (2) could not answer from the target side (none on connected masks); (4) the windows (3) could not hold;
    // (5) all levels for what level 0 cannot answer.
    if (wide) {
        // max_px 512 .. 1023: sparse tile-word windows on the 256-column circle (level 0, 2 CTAs per SM, then 1 CTA per SM
        // for the windows that did not fit); what they cannot hold or answer goes to the all-levels kernel with its bitmaps
        // in global memory.  AXB_K2_WIDE_GENERAL_ONLY (round-1 behaviour, for A/B timing): every source on that kernel.
        static const bool general_only = getenv("AXB_K2_WIDE_GENERAL_ONLY") != nullptr;
        const int g1 = std::max(1, std::min(c->num_sms * axb::kWideSparseCtasPerSm, total_src));
        const int g2 = std::max(1, std::min(c->num_sms, total_src));
        if (general_only) axb::launch_k2_all_to_general(prep.work_prefix, n_pairs, c->gen_list, c->counters, s);
        else {
            set_source(p, nullptr, 0, axb::kCtrP1Ticket, dense_a, axb::kCtrDenseALen, cover ? 1 : 0);
            axb::launch_k2_sparse_wide(p, g1, false, s);
            set_source(p, dense_a, axb::kCtrDenseALen, axb::kCtrDenseATicket, nullptr, 0, cover ? 1 : 0);
            axb::launch_k2_sparse_wide(p, g2, true, s);
            if (cover) {
                set_source(p, c->dir_list, axb::kCtrDirLen, axb::kCtrDirTicket, dense_b, axb::kCtrDenseBLen, 0);
                axb::launch_k2_sparse_wide(p, g1, false, s);
                set_source(p, dense_b, axb::kCtrDenseBLen, axb::kCtrDenseBTicket, nullptr, 0, 0);
                axb::launch_k2_sparse_wide(p, g2, true, s);
                c->launches += 2;
            }
        }
        if (early) AXB_CUDA(c, cudaStreamWaitEvent(s, c->ev_join, 0));      // one scratch set: the late list after the early one
        axb::launch_k2_general(p, gen_grid, s);
        c->launches += 4;
        AXB_CUDA(c, cudaGetLastError());
        static const bool dbg = getenv("AXB_K2_DEBUG") != nullptr;     // profiling aid: where the work items went
        if (dbg) {
            int h[16];
            cudaStreamSynchronize(s);
            cudaMemcpy(h, c->counters, sizeof h, cudaMemcpyDeviceToHost);
            fprintf(stderr, "[axb] wide K2: %d sources, %d first-pass tickets, %d to the second kind, %d directed, all-levels kernel: %d early + %d late\n",
                    total_src, h[0], h[3], h[5], h[9], h[1]);
        }
        return AXB_OK;
    }
    // one search per CTA.  Measured alternatives (profiles/r02a_*, r02p_*, bit-identical results): one WARP per search
    // (13.3 ms; a rewrite with two entries per lane 14.3 .. 14.9 ms) and two searches POOLED per CTA (13.6 ms) against
    // 12.1 ms on the config-2 block -- fewer instructions, but too few warps per SM to hide the latency between
    // dependent instructions; all were removed again.
    const int gs = std::max(1, std::min(c->num_sms * axb::kSparseCtasPerSm, total_src));
    const int gd = std::max(1, std::min(c->num_sms * axb::kFastCtasPerSm, total_src));
    set_source(p, nullptr, 0, axb::kCtrP1Ticket, dense_a, axb::kCtrDenseALen, cover ? 1 : 0);
    axb::launch_k2_sparse(p, gs, s);
    set_source(p, dense_a, axb::kCtrDenseALen, axb::kCtrDenseATicket, nullptr, 0, cover ? 1 : 0);
    axb::launch_k2_fast(p, gd, s);
    if (cover) {
        set_source(p, c->dir_list, axb::kCtrDirLen, axb::kCtrDirTicket, dense_b, axb::kCtrDenseBLen, 0);
        axb::launch_k2_sparse(p, gs, s);
        set_source(p, dense_b, axb::kCtrDenseBLen, axb::kCtrDenseBTicket, nullptr, 0, 0);
        axb::launch_k2_fast(p, gd, s);
        c->launches += 2;
    }
    if (early) AXB_CUDA(c, cudaStreamWaitEvent(s, c->ev_join, 0));          // one scratch set: the late list after the early one
    axb::launch_k2_general(p, gen_grid, s);
    c->launches += 4;
    AXB_CUDA(c, cudaGetLastError());
    static const bool dbg = getenv("AXB_K2_DEBUG") != nullptr;
    if (dbg) {
        int h[16];
        cudaStreamSynchronize(s);
        cudaMemcpy(h, c->counters, sizeof h, cudaMemcpyDeviceToHost);
        fprintf(stderr, "[axb] K2: %d sources, %d first-pass tickets, %d dense windows, %d directed (+%d dense), all-levels kernel: %d early + %d late\n",
                total_src, h[0], h[3], h[5], h[7], h[9], h[1]);
    }
    return AXB_OK;
}

}  // namespace

extern "C" {

int axb_prepare_masks(axb_ctx* c, const uint32_t* mask_bits_dev, int n_masks, int H, int W, void* stream) {
    if (!c) return AXB_ERR_BAD_ARG;
    if (!mask_bits_dev) { c->prep_bits = nullptr; return AXB_OK; }      // discard what an earlier call prepared
    if (H <= 0 || W <= 0 || n_masks <= 0) return bad_arg(c, "axb_prepare_masks: bad size");
    AXB_ON_DEVICE(c);
    c->prep_bits = nullptr;
    MaskProducts mp;
    const bool comps = cover_possible(n_masks, H, W);
    const int rc = mask_products(c, mask_bits_dev, n_masks, H, W, comps, (cudaStream_t)stream, &mp);
    if (rc != AXB_OK) return rc;
    AXB_CUDA(c, cudaGetLastError());
    c->prep_bits = mask_bits_dev; c->prep_n_masks = n_masks; c->prep_H = H; c->prep_W = W; c->prep_components = comps;
    return AXB_OK;
}

int axb_cost_matrix_batch(axb_ctx* c, const uint32_t* mask_bits_dev, int n_masks, int H, int W,
                          const int32_t* dets_yx_dev, int n_dets, const axb_pair_desc* pairs_dev, int n_pairs,
                          int total_src, int max_px,
                          int32_t* path_len_dev, uint8_t* unreachable_dev, float* wcost_dev, int64_t out_capacity,
                          void* stream) {
    return cost_matrix_batch_impl(c, mask_bits_dev, n_masks, H, W, dets_yx_dev, n_dets, pairs_dev, n_pairs, total_src, max_px,
                                  path_len_dev, unreachable_dev, wcost_dev, out_capacity, 0, stream);
}

int axb_cost_matrix(axb_ctx* c, const uint32_t* mask_bits_dev, int H, int W,
                    const int32_t* src_yx_dev, int n_src, const int32_t* dst_yx_dev, int n_dst,
                    int max_px, int32_t* path_len_dev, uint8_t* unreachable_dev, float* wcost_dev,
                    void* stream) {
    if (!c) return AXB_ERR_BAD_ARG;
    if (n_src < 0 || n_dst < 0) return bad_arg(c, "axb_cost_matrix: negative count");
    if (n_src == 0 || n_dst == 0) return AXB_OK;
    if (!src_yx_dev || !dst_yx_dev) return bad_arg(c, "axb_cost_matrix: NULL detections");
    AXB_ON_DEVICE(c);
    cudaStream_t s = (cudaStream_t)stream;
    int rc = grow(c, &c->st_dets, &c->st_dets_bytes, (size_t)(n_src + n_dst) * 8);
    if (rc != AXB_OK) return rc;
    rc = grow(c, &c->st_pairs, &c->st_pairs_bytes, sizeof(axb_pair_desc));
    if (rc != AXB_OK) return rc;
    AXB_CUDA(c, cudaMemcpyAsync(c->st_dets, src_yx_dev, (size_t)n_src * 8, cudaMemcpyDeviceToDevice, s));
    AXB_CUDA(c, cudaMemcpyAsync(c->st_dets + 2 * (size_t)n_src, dst_yx_dev, (size_t)n_dst * 8, cudaMemcpyDeviceToDevice, s));
    axb_pair_desc pd;
    memset(&pd, 0, sizeof pd);
    pd.mask_idx = 0; pd.src_off = 0; pd.n_src = n_src; pd.dst_off = n_src; pd.n_dst = n_dst; pd.out_off = 0;
    AXB_CUDA(c, cudaMemcpyAsync(c->st_pairs, &pd, sizeof pd, cudaMemcpyHostToDevice, s));   // pageable source: staged before return
    return axb_cost_matrix_batch(c, mask_bits_dev, 1, H, W, c->st_dets, n_src + n_dst, c->st_pairs, 1, n_src, max_px,
                                 path_len_dev, unreachable_dev, wcost_dev, (int64_t)n_src * n_dst, stream);
}

int axb_cost_matrix_batch_host(axb_ctx* c, const uint8_t* masks_host, int n_masks, int H, int W,
                               const int32_t* dets_yx_host, int n_dets,
                               const axb_pair_desc* pairs_host, int n_pairs, int max_px,
                               int32_t* path_len_host, uint8_t* unreachable_host, float* wcost_host) {
    if (!c) return AXB_ERR_BAD_ARG;
    if (n_pairs < 0 || n_dets < 0 || n_masks <= 0 || H <= 0 || W <= 0) return bad_arg(c, "axb_cost_matrix_batch_host: bad size");
    if (n_pairs == 0) return AXB_OK;
    if (!masks_host || !pairs_host || !path_len_host || !unreachable_host || (n_dets > 0 && !dets_yx_host))
        return bad_arg(c, "axb_cost_matrix_batch_host: NULL pointer");
    long long total_out = 0; long long total_src = 0;
    std::vector<long long> src_prefix((size_t)n_pairs + 1, 0);
    for (int i = 0; i < n_pairs; ++i) {
        const axb_pair_desc& pd = pairs_host[i];
        if (pd.n_src < 0 || pd.n_dst < 0 || pd.src_off < 0 || pd.dst_off < 0 || pd.src_off + pd.n_src > n_dets ||
            pd.dst_off + pd.n_dst > n_dets || pd.mask_idx < 0 || pd.mask_idx >= n_masks || pd.out_off < 0)
            return bad_arg(c, "axb_cost_matrix_batch_host: pair descriptor out of range");
        total_out = std::max(total_out, (long long)pd.out_off + (long long)pd.n_src * pd.n_dst);
        total_src += pd.n_src;
        src_prefix[(size_t)i + 1] = total_src;
    }
    if (total_src > 0x7fffffffLL) return bad_arg(c, "axb_cost_matrix_batch_host: too many sources");
    AXB_ON_DEVICE(c);
    cudaStream_t s = c->own_stream;
    const size_t mask_bytes = (size_t)n_masks * H * W;
    const size_t bits_bytes = (size_t)n_masks * H * ((W + 31) / 32) * 4;
    int rc;
    if ((rc = grow(c, &c->st_mask, &c->st_mask_bytes, mask_bytes)) != AXB_OK) return rc;
    if ((rc = grow(c, &c->st_bits, &c->st_bits_bytes, bits_bytes)) != AXB_OK) return rc;
    if ((rc = grow(c, &c->st_dets, &c->st_dets_bytes, (size_t)std::max(n_dets, 1) * 8)) != AXB_OK) return rc;
    if ((rc = grow(c, &c->st_pairs, &c->st_pairs_bytes, (size_t)n_pairs * sizeof(axb_pair_desc))) != AXB_OK) return rc;
    if ((rc = grow(c, &c->st_plen, &c->st_plen_bytes, (size_t)std::max(total_out, 1LL) * 4)) != AXB_OK) return rc;
    if ((rc = grow(c, &c->st_unr, &c->st_unr_bytes, (size_t)std::max(total_out, 1LL))) != AXB_OK) return rc;
    if (wcost_host && (rc = grow(c, &c->st_wc, &c->st_wc_bytes, (size_t)std::max(total_out, 1LL) * 4)) != AXB_OK) return rc;
    AXB_CUDA(c, cudaMemcpyAsync(c->st_mask, masks_host, mask_bytes, cudaMemcpyHostToDevice, s));
    if (n_dets > 0) AXB_CUDA(c, cudaMemcpyAsync(c->st_dets, dets_yx_host, (size_t)n_dets * 8, cudaMemcpyHostToDevice, s));
    AXB_CUDA(c, cudaMemcpyAsync(c->st_pairs, pairs_host, (size_t)n_pairs * sizeof(axb_pair_desc), cudaMemcpyHostToDevice, s));
    if ((rc = axb_pack_mask(c, c->st_mask, n_masks, H, W, c->st_bits, s)) != AXB_OK) return rc;

    // Chunks of frame pairs: the searches of chunk k+1 (other scratch set, other stream) fill the SMs as chunk k
    // drains, and the copy stream takes chunk k's matrices to the host meanwhile.  Tiles and components of the masks
    // are computed once for all chunks.  One chunk up to 320 MB of results (configs[1]: 44 MB, measured faster
    // unchunked), up to 8 beyond (configs[2], 1.36 GB: unchunked 97.7 ms, 5 equal chunks 87.7, 8 shrinking chunks
    // 78.9 ms end to end against 72.3 ms on the device).
    constexpr double kChunkRatio = 0.8;
    static const bool no_chunks = getenv("AXB_HOST_NO_CHUNKS") != nullptr;
    const double out_mb = (double)total_out * (wcost_host ? 9.0 : 5.0) / (1 << 20);
    int n_chunks = (int)std::min<long long>(8, (long long)(out_mb / 160.0));
    if (const char* e = getenv("AXB_HOST_CHUNKS")) n_chunks = std::max(1, std::min(8, atoi(e)));      // experiments
    // the smallest (last) chunk still holds two full rounds of searches
    while (n_chunks > 1 && total_src * (1.0 - kChunkRatio) * std::pow(kChunkRatio, n_chunks - 1) / (1.0 - std::pow(kChunkRatio, n_chunks)) < 2.0 * c->num_sms * axb::kSparseCtasPerSm) --n_chunks;
    if (n_chunks < 2 || no_chunks || total_out <= 0) n_chunks = 1;
    axb_ctx* ctxs[2] = {c, c};
    MaskProducts shared;               // tiles and components once for all chunks, ahead of the fork
    if ((rc = mask_products(c, c->st_bits, n_masks, H, W, n_pairs >= 2 && cover_possible(n_masks, H, W), s, &shared)) != AXB_OK) return rc;
    if (n_chunks > 1) {
        if (!c->alt && (rc = axb_create(c->device, &c->alt)) != AXB_OK) return rc;
        if (!c->copy_stream) AXB_CUDA(c, cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
        if (!c->ev_in) AXB_CUDA(c, cudaEventCreateWithFlags(&c->ev_in, cudaEventDisableTiming));
        for (int i = 0; i < 2; ++i)
            if (!c->ev_chunk[i]) AXB_CUDA(c, cudaEventCreateWithFlags(&c->ev_chunk[i], cudaEventDisableTiming));
        ctxs[1] = c->alt;
        AXB_CUDA(c, cudaEventRecord(c->ev_in, s));
        AXB_CUDA(c, cudaStreamWaitEvent(c->alt->own_stream, c->ev_in, 0));
    }
    if (total_out > 0 && total_src > 0) {
        int p0 = 0;
        for (int k = 0; k < n_chunks; ++k) {
            int p1 = n_pairs;
            if (k + 1 < n_chunks) {
                // Shrinking chunks (each 0.8 of the one before): a chunk's copy (~54 GB/s) hides behind the searches of
                // the next one (~19 GB/s of results on configs[2]) as long as that one is at least a third as large,
                // and the copy of the LAST chunk is exposed -- so it should be the smallest (measured on configs[2]:
                // 5 equal chunks 87.7 ms, 8 chunks of ratio 0.8 78.9 ms, unchunked 97.7 ms end to end).
                static const double r = getenv("AXB_HOST_CHUNK_RATIO") ? atof(getenv("AXB_HOST_CHUNK_RATIO")) : kChunkRatio;
                const double frac = r >= 0.999 ? (double)(k + 1) / n_chunks : (1.0 - std::pow(r, k + 1)) / (1.0 - std::pow(r, n_chunks));
                const long long want = (long long)(total_src * frac);
                p1 = (int)(std::lower_bound(src_prefix.begin(), src_prefix.end(), want) - src_prefix.begin());
                p1 = std::min(std::max(p1, p0), n_pairs);
            }
            if (p1 == p0) continue;
            axb_ctx* const cc = ctxs[k & 1];
            cudaStream_t cs = cc->own_stream;
            rc = cost_matrix_batch_impl(cc, c->st_bits, n_masks, H, W, c->st_dets, n_dets, c->st_pairs + p0, p1 - p0,
                                        (int)(src_prefix[p1] - src_prefix[p0]), max_px, c->st_plen, c->st_unr,
                                        wcost_host ? c->st_wc : nullptr, (int64_t)total_out, src_prefix[p0], cs, &shared);
            if (rc != AXB_OK) { if (cc != c) c->err = cc->err; return rc; }
            long long lo = total_out, hi = 0;
            for (int i = p0; i < p1; ++i) {
                lo = std::min(lo, (long long)pairs_host[i].out_off);
                hi = std::max(hi, (long long)pairs_host[i].out_off + (long long)pairs_host[i].n_src * pairs_host[i].n_dst);
            }
            cudaStream_t ds = cs;
            if (n_chunks > 1) {
                AXB_CUDA(c, cudaEventRecord(c->ev_chunk[k & 1], cs));
                AXB_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev_chunk[k & 1], 0));
                ds = c->copy_stream;
            }
            if (hi > lo) {
                AXB_CUDA(c, cudaMemcpyAsync(path_len_host + lo, c->st_plen + lo, (size_t)(hi - lo) * 4, cudaMemcpyDeviceToHost, ds));
                AXB_CUDA(c, cudaMemcpyAsync(unreachable_host + lo, c->st_unr + lo, (size_t)(hi - lo), cudaMemcpyDeviceToHost, ds));
                if (wcost_host) AXB_CUDA(c, cudaMemcpyAsync(wcost_host + lo, c->st_wc + lo, (size_t)(hi - lo) * 4, cudaMemcpyDeviceToHost, ds));
            }
            p0 = p1;
        }
    }
    // status of both scratch sets: the first error in the reference's iteration order (the keys are global)
    unsigned long long key = ~0ull, flags = 0ull;
    for (int i = 0; i < (n_chunks > 1 ? 2 : 1); ++i) {
        unsigned long long k2 = ~0ull, f2 = 0ull;
        if ((rc = status_fetch(ctxs[i], ctxs[i]->own_stream, &k2, &f2)) != AXB_OK) return rc;
        key = std::min(key, k2); flags |= f2;
    }
    if (n_chunks > 1) AXB_CUDA(c, cudaStreamSynchronize(c->copy_stream));
    int64_t info[3];
    rc = status_decode(c, key, flags, info, [&](long long item, int64_t* pair, int64_t* src) {
        const int lo = (int)(std::upper_bound(src_prefix.begin(), src_prefix.begin() + n_pairs, item) - src_prefix.begin()) - 1;
        *pair = std::max(lo, 0); *src = item - src_prefix[std::max(lo, 0)];
    });
    if ((rc == AXB_ERR_OOB_START || rc == AXB_ERR_OOB_GOAL) && info[0] >= 0 && info[0] < n_pairs) {
        // the reference's message (astar_wrapper.py:41,45)
        const axb_pair_desc& pd = pairs_host[info[0]];
        const bool st = rc == AXB_ERR_OOB_START;
        const int64_t row = st ? (int64_t)pd.src_off + info[1] : (int64_t)pd.dst_off + info[2];
        char buf[128];
        snprintf(buf, sizeof buf, "%s of (%d, %d) lies outside grid.", st ? "Start" : "Goal",
                 dets_yx_host[2 * row], dets_yx_host[2 * row + 1]);
        c->err = buf;
    }
    return rc;
}

int axb_cost_matrix_host(axb_ctx* c, const uint8_t* mask_host, int H, int W,
                         const int32_t* src_yx_host, int n_src, const int32_t* dst_yx_host, int n_dst,
                         int max_px, int32_t* path_len_host, uint8_t* unreachable_host, float* wcost_host) {
    if (!c) return AXB_ERR_BAD_ARG;
    if (n_src < 0 || n_dst < 0) return bad_arg(c, "axb_cost_matrix_host: negative count");
    if (n_src == 0 || n_dst == 0) return AXB_OK;
    if (!src_yx_host || !dst_yx_host) return bad_arg(c, "axb_cost_matrix_host: NULL detections");
    std::vector<int32_t> dets((size_t)(n_src + n_dst) * 2);
    memcpy(dets.data(), src_yx_host, (size_t)n_src * 8);
    memcpy(dets.data() + 2 * (size_t)n_src, dst_yx_host, (size_t)n_dst * 8);
    axb_pair_desc pd;
    memset(&pd, 0, sizeof pd);
    pd.n_src = n_src; pd.dst_off = n_src; pd.n_dst = n_dst;
    return axb_cost_matrix_batch_host(c, mask_host, 1, H, W, dets.data(), n_src + n_dst, &pd, 1, max_px,
                                      path_len_host, unreachable_host, wcost_host);
}

int axb_decode_nms(axb_ctx* c, const float* Y_dev, const int32_t* tile_yx_dev,
                   int F, int ntiles, int Sx, int Sy, int tilesize, float conf_thr, int nms_dist, int cap,
                   float* pre_conf_dev, int32_t* pre_yx_dev, int32_t* pre_cell_dev, int32_t* n_pre_dev,
                   float* conf_dev, int32_t* yx_dev, int32_t* keep_idx_dev, int32_t* n_out_dev, void* stream) {
    if (!c) return AXB_ERR_BAD_ARG;
    if (F < 0 || ntiles <= 0 || Sx <= 0 || Sy <= 0 || tilesize <= 0 || cap <= 0 || nms_dist < 0 || nms_dist > 30000)
        return bad_arg(c, "axb_decode_nms: bad size");
    if (F == 0) return AXB_OK;
    if (!Y_dev || !tile_yx_dev || !n_pre_dev || !conf_dev || !yx_dev || !keep_idx_dev || !n_out_dev)
        return bad_arg(c, "axb_decode_nms: NULL pointer");
    const long long ncell_ll = (long long)ntiles * Sx * Sy;
    if (ncell_ll > (1 << 24)) return bad_arg(c, "axb_decode_nms: more than 2^24 grid cells per frame");
    AXB_ON_DEVICE(c);
    const int ncell = (int)ncell_ll;
    int npad = 1;
    while (npad < ncell) npad <<= 1;
    const size_t per_frame = axb::k1_scratch_per_frame(ncell, npad, ntiles);
    int rc = grow(c, &c->k1_scratch, &c->k1_scratch_bytes, per_frame * (size_t)F);
    if (rc != AXB_OK) return rc;
    axb::K1Params p;
    p.Y = Y_dev; p.tile_yx = tile_yx_dev; p.F = F; p.ntiles = ntiles; p.Sx = Sx; p.Sy = Sy; p.tilesize = tilesize;
    p.conf_thr = conf_thr; p.nms_d2 = nms_dist * nms_dist; p.cap = cap;
    p.pre_conf = pre_conf_dev; p.pre_yx = pre_yx_dev; p.pre_cell = pre_cell_dev; p.n_pre = n_pre_dev;
    p.conf = conf_dev; p.yx = yx_dev; p.keep_idx = keep_idx_dev; p.n_out = n_out_dev;
    p.scratch = c->k1_scratch; p.scratch_per_frame = per_frame; p.npad = npad; p.err_flags = c->status_dev + 1;
    if (!c->counters) AXB_CUDA(c, cudaMalloc((void**)&c->counters, 256));
    p.wide_ctl = c->counters + 32; p.num_sms = c->num_sms;       // counters[0..7] belong to K2, [32..39] to the wide K1
    axb::launch_k1(p, (cudaStream_t)stream);
    c->launches += 1;
    AXB_CUDA(c, cudaGetLastError());
    return AXB_OK;
}

int axb_decode_nms_host(axb_ctx* c, const float* Y_host, const int32_t* tile_yx_host,
                        int F, int ntiles, int Sx, int Sy, int tilesize, float conf_thr, int nms_dist, int cap,
                        float* pre_conf_host, int32_t* pre_yx_host, int32_t* pre_cell_host, int32_t* n_pre_host,
                        float* conf_host, int32_t* yx_host, int32_t* keep_idx_host, int32_t* n_out_host) {
    if (!c) return AXB_ERR_BAD_ARG;
    if (F < 0 || ntiles <= 0 || Sx <= 0 || Sy <= 0 || cap <= 0) return bad_arg(c, "axb_decode_nms_host: bad size");
    if (F == 0) return AXB_OK;
    if (!Y_host || !tile_yx_host || !n_pre_host || !conf_host || !yx_host || !keep_idx_host || !n_out_host)
        return bad_arg(c, "axb_decode_nms_host: NULL pointer");
    AXB_ON_DEVICE(c);
    cudaStream_t s = c->own_stream;
    auto al = [](size_t b) { return (b + 255) & ~(size_t)255; };
    const size_t nY = (size_t)F * ntiles * Sx * Sy * 3 * 4, nT = (size_t)ntiles * 8, nFC = (size_t)F * cap * 4, nF = (size_t)F * 4;
    const size_t total = al(nY) + al(nT) + 2 * al(nF) + 3 * al(nFC) + 2 * al(2 * nFC) + al(nFC);
    int rc = grow(c, &c->st_k1, &c->st_k1_bytes, total);
    if (rc != AXB_OK) return rc;
    uint8_t* b = c->st_k1;
    auto take = [&](size_t bytes) { uint8_t* p = b; b += al(bytes); return p; };
    float* dY = (float*)take(nY);            int32_t* dT = (int32_t*)take(nT);
    int32_t* d_npre = (int32_t*)take(nF);    int32_t* d_nout = (int32_t*)take(nF);
    float* d_pconf = (float*)take(nFC);      int32_t* d_ptile = (int32_t*)take(nFC);    int32_t* d_pyx = (int32_t*)take(2 * nFC);
    float* d_conf = (float*)take(nFC);       int32_t* d_yx = (int32_t*)take(2 * nFC);   int32_t* d_keep = (int32_t*)take(nFC);
    AXB_CUDA(c, cudaMemcpyAsync(dY, Y_host, nY, cudaMemcpyHostToDevice, s));
    AXB_CUDA(c, cudaMemcpyAsync(dT, tile_yx_host, nT, cudaMemcpyHostToDevice, s));
    rc = axb_decode_nms(c, dY, dT, F, ntiles, Sx, Sy, tilesize, conf_thr, nms_dist, cap,
                        pre_conf_host ? d_pconf : nullptr, pre_yx_host ? d_pyx : nullptr, pre_cell_host ? d_ptile : nullptr,
                        d_npre, d_conf, d_yx, d_keep, d_nout, s);
    if (rc != AXB_OK) return rc;
    AXB_CUDA(c, cudaMemcpyAsync(n_pre_host, d_npre, nF, cudaMemcpyDeviceToHost, s));
    AXB_CUDA(c, cudaMemcpyAsync(n_out_host, d_nout, nF, cudaMemcpyDeviceToHost, s));
    AXB_CUDA(c, cudaMemcpyAsync(conf_host, d_conf, nFC, cudaMemcpyDeviceToHost, s));
    AXB_CUDA(c, cudaMemcpyAsync(yx_host, d_yx, 2 * nFC, cudaMemcpyDeviceToHost, s));
    AXB_CUDA(c, cudaMemcpyAsync(keep_idx_host, d_keep, nFC, cudaMemcpyDeviceToHost, s));
    if (pre_conf_host) AXB_CUDA(c, cudaMemcpyAsync(pre_conf_host, d_pconf, nFC, cudaMemcpyDeviceToHost, s));
    if (pre_yx_host) AXB_CUDA(c, cudaMemcpyAsync(pre_yx_host, d_pyx, 2 * nFC, cudaMemcpyDeviceToHost, s));
    if (pre_cell_host) AXB_CUDA(c, cudaMemcpyAsync(pre_cell_host, d_ptile, nFC, cudaMemcpyDeviceToHost, s));
    return axb_status(c, s, nullptr);
}

int axb_trace_paths(axb_ctx* c, const uint32_t* mask_bits_dev, int n_masks, int H, int W,
                    const int32_t* req_host, int n_req, int max_px,
                    int32_t* path_yx_dev, int32_t* path_len_dev, void* stream) {
    if (!c) return AXB_ERR_BAD_ARG;
    if (n_req < 0 || H <= 0 || W <= 0 || n_masks <= 0) return bad_arg(c, "axb_trace_paths: bad size");
    if (max_px < 1 || max_px > AXB_MAX_PX) return bad_arg(c, "axb_trace_paths: max_px outside [1, AXB_MAX_PX]");
    if (n_req == 0) return AXB_OK;
    if (!mask_bits_dev || !req_host || !path_yx_dev || !path_len_dev) return bad_arg(c, "axb_trace_paths: NULL pointer");
    AXB_ON_DEVICE(c);
    cudaStream_t s = (cudaStream_t)stream;
    int rc = ensure_k2(c, n_req, n_req);
    if (rc != AXB_OK) return rc;
    const bool wide = max_px > axb::kFastMaxPx;
    if (!wide && !c->trace_par) AXB_CUDA(c, cudaMalloc((void**)&c->trace_par, (size_t)c->k2_gen_ctas * axb::kTraceParBytes));
    if (wide && !c->trace_par_wide) AXB_CUDA(c, cudaMalloc((void**)&c->trace_par_wide, (size_t)c->k2_gen_ctas * axb::kWideCtasPerSm * axb::kTraceParBytesWide));
    if (wide && !c->k2_wide_scratch)
        AXB_CUDA(c, cudaMalloc((void**)&c->k2_wide_scratch, (size_t)c->k2_gen_ctas * axb::kWideCtasPerSm * axb::kGenScratchWordsWide * sizeof(uint32_t)));
    // one 1 x 1 frame pair per request: detections rows 2i (source), 2i+1 (target)
    std::vector<int32_t> dets((size_t)n_req * 4);
    std::vector<axb_pair_desc> pairs((size_t)n_req);
    std::vector<int> iota((size_t)n_req + 1);
    for (int i = 0; i < n_req; ++i) {
        const int32_t* r = req_host + 5 * (size_t)i;
        if (r[0] < 0 || r[0] >= n_masks) return bad_arg(c, "axb_trace_paths: mask index out of range");
        dets[4 * (size_t)i] = r[1]; dets[4 * (size_t)i + 1] = r[2]; dets[4 * (size_t)i + 2] = r[3]; dets[4 * (size_t)i + 3] = r[4];
        axb_pair_desc& pd = pairs[i];
        memset(&pd, 0, sizeof pd);
        pd.mask_idx = r[0]; pd.src_off = 2 * i; pd.n_src = 1; pd.dst_off = 2 * i + 1; pd.n_dst = 1; pd.out_off = i;
        iota[i] = i;
    }
    iota[n_req] = n_req;
    if ((rc = grow(c, &c->st_dets, &c->st_dets_bytes, dets.size() * 4)) != AXB_OK) return rc;
    if ((rc = grow(c, &c->st_pairs, &c->st_pairs_bytes, pairs.size() * sizeof(axb_pair_desc))) != AXB_OK) return rc;
    if ((rc = grow(c, &c->st_plen, &c->st_plen_bytes, (size_t)n_req * 4)) != AXB_OK) return rc;
    if ((rc = grow(c, &c->st_unr, &c->st_unr_bytes, (size_t)n_req)) != AXB_OK) return rc;
    AXB_CUDA(c, cudaMemcpyAsync(c->st_dets, dets.data(), dets.size() * 4, cudaMemcpyHostToDevice, s));      // pageable: staged before return
    AXB_CUDA(c, cudaMemcpyAsync(c->st_pairs, pairs.data(), pairs.size() * sizeof(axb_pair_desc), cudaMemcpyHostToDevice, s));
    const axb::K2Prep prep = prep_arrays(c);
    axb::launch_k2_prep(c->st_pairs, n_req, prep, c->counters, n_masks, 2 * n_req, n_req, n_req, 0, c->status_dev + 1, s);
    c->last_n_pairs = n_req;
    // every request goes straight to the all-levels kernel: gen_list = 0 .. n-1, counters[1] = n
    AXB_CUDA(c, cudaMemcpyAsync(c->gen_list, iota.data(), (size_t)n_req * sizeof(int), cudaMemcpyHostToDevice, s));
    AXB_CUDA(c, cudaMemcpyAsync(c->counters + 1, &iota[n_req], sizeof(int), cudaMemcpyHostToDevice, s));
    axb::K2Params p;
    memset(&p, 0, sizeof p);
    p.mask_bits = mask_bits_dev; p.H = H; p.W = W; p.Wp = (W + 31) / 32;
    p.dets_yx = c->st_dets; p.pairs = c->st_pairs; p.n_pairs = n_req;
    set_prep(p, prep, c);
    p.max_px = max_px; p.path_len = c->st_plen; p.unreachable = c->st_unr; p.wcost = nullptr;
    p.fast_scratch = c->k2_fast_scratch; p.gen_scratch = wide ? c->k2_wide_scratch : c->k2_gen_scratch;
    p.err_key = c->status_dev;
    p.trace_par = wide ? c->trace_par_wide : c->trace_par; p.trace_yx = path_yx_dev; p.trace_len = path_len_dev;
    axb::launch_k2_general(p, std::max(1, std::min(c->k2_gen_ctas * (wide ? axb::kWideCtasPerSm : 1), n_req)), s);
    c->launches += 2;
    AXB_CUDA(c, cudaGetLastError());
    return AXB_OK;
}

int axb_feature_hist_frames(axb_ctx* c, const float* images_dev, int n_frames, int H, int W, const int32_t* boxes_dev,
                            const int32_t* box_frame_dev, int n_boxes, int bins, float lo, float hi, float* out_dev, void* stream) {
    if (!c) return AXB_ERR_BAD_ARG;
    if (H <= 0 || W <= 0 || n_frames < 1 || n_boxes < 0 || bins < 1 || bins > AXB_MAX_BINS || !(hi > lo))
        return bad_arg(c, "axb_feature_hist: bad size, bins outside [1, AXB_MAX_BINS] or empty range");
    if (n_boxes == 0) return AXB_OK;
    if (!images_dev || !boxes_dev || !out_dev) return bad_arg(c, "axb_feature_hist: NULL pointer");
    AXB_ON_DEVICE(c);
    axb::launch_k5(images_dev, n_frames, H, W, boxes_dev, box_frame_dev, n_boxes, bins, lo, hi, out_dev, (cudaStream_t)stream);
    c->launches += 1;
    AXB_CUDA(c, cudaGetLastError());
    return AXB_OK;
}

int axb_feature_hist(axb_ctx* c, const float* image_dev, int H, int W, const int32_t* boxes_dev, int n_boxes,
                     int bins, float lo, float hi, float* out_dev, void* stream) {
    return axb_feature_hist_frames(c, image_dev, 1, H, W, boxes_dev, nullptr, n_boxes, bins, lo, hi, out_dev, stream);
}

int axb_vis_sim(axb_ctx* c, const float* feats_src_dev, int n_src, const float* feats_dst_dev, int n_dst, int bins,
                const int32_t* path_len_dev, int max_px, double* vis_dev, void* stream) {
    if (!c) return AXB_ERR_BAD_ARG;
    if (n_src < 0 || n_dst < 0 || bins < 1 || bins > AXB_MAX_BINS) return bad_arg(c, "axb_vis_sim: bad size or bins outside [1, AXB_MAX_BINS]");
    if (n_src == 0 || n_dst == 0) return AXB_OK;
    if (!feats_src_dev || !feats_dst_dev || !vis_dev) return bad_arg(c, "axb_vis_sim: NULL pointer");
    AXB_ON_DEVICE(c);
    axb::launch_k7(feats_src_dev, n_src, feats_dst_dev, n_dst, bins, path_len_dev, max_px, vis_dev, (cudaStream_t)stream);
    c->launches += 1;
    AXB_CUDA(c, cudaGetLastError());
    return AXB_OK;
}

int axb_transition_cost_batch(axb_ctx* c, const int32_t* path_len_dev, const axb_pair_desc* pairs_dev, int n_pairs,
                              int max_elems_per_pair, const double* cost_tab_dev, const int64_t* arc_tab_dev,
                              int n_gaps, int max_px, double* cost_dev, int64_t* arc_dev, void* stream) {
    if (!c) return AXB_ERR_BAD_ARG;
    if (n_pairs < 0 || n_gaps <= 0 || max_px < 1 || max_px > AXB_MAX_PX || max_elems_per_pair < 0)
        return bad_arg(c, "axb_transition_cost_batch: bad size or max_px outside [1, AXB_MAX_PX]");
    if (n_pairs == 0 || max_elems_per_pair == 0) return AXB_OK;
    if (!path_len_dev || !pairs_dev || !cost_tab_dev || !cost_dev || (arc_tab_dev && !arc_dev))
        return bad_arg(c, "axb_transition_cost_batch: NULL pointer");
    AXB_ON_DEVICE(c);
    axb::launch_k3(path_len_dev, pairs_dev, n_pairs, cost_tab_dev, (const long long*)arc_tab_dev, n_gaps, max_px,
                   cost_dev, arc_tab_dev ? (long long*)arc_dev : nullptr, c->status_dev + 1, max_elems_per_pair,
                   (cudaStream_t)stream);
    c->launches += 1;
    AXB_CUDA(c, cudaGetLastError());
    return AXB_OK;
}

int axb_tile_nonempty(axb_ctx* c, const float* X_dev, int T, int C, int H, int W, int tilesize, uint8_t* nonempty_dev,
                      void* stream) {
    if (!c) return AXB_ERR_BAD_ARG;
    if (T < 0 || C < 1 || H <= 0 || W <= 0 || tilesize < 1 || T > 65535) return bad_arg(c, "axb_tile_nonempty: bad size");
    if (T == 0) return AXB_OK;
    if (!X_dev || !nonempty_dev) return bad_arg(c, "axb_tile_nonempty: NULL pointer");
    AXB_ON_DEVICE(c);
    const size_t n_flags = (size_t)((H + tilesize - 1) / tilesize) * ((W + tilesize - 1) / tilesize) * T;
    int rc = grow(c, &c->k6_flags, &c->k6_flags_bytes, n_flags * axb::kK6FlagStride * sizeof(uint32_t));
    if (rc != AXB_OK) return rc;
    axb::launch_k6_nonempty(X_dev, T, C, H, W, tilesize, c->k6_flags, nonempty_dev, (cudaStream_t)stream);
    c->launches += 2;
    AXB_CUDA(c, cudaGetLastError());
    return AXB_OK;
}

// No cudaMalloc / cudaFree on this path: the vote flags live in a static device array of axb_tiles.cu.
int axb_tile_nonempty_noalloc(axb_ctx* c, const float* X_dev, int T, int C, int H, int W, int tilesize, uint8_t* nonempty_dev,
                              void* stream) {
    if (!c) return AXB_ERR_BAD_ARG;
    if (T < 0 || C < 1 || H <= 0 || W <= 0 || tilesize < 1 || T > 65535) return bad_arg(c, "axb_tile_nonempty_noalloc: bad size");
    if (T == 0) return AXB_OK;
    if (!X_dev || !nonempty_dev) return bad_arg(c, "axb_tile_nonempty_noalloc: NULL pointer");
    const unsigned long long n_flags = (unsigned long long)((H + tilesize - 1) / tilesize) * ((W + tilesize - 1) / tilesize) * T;
    if (n_flags > (unsigned long long)AXB_NOALLOC_MAX_TILE_FRAMES) {
        char buf[200];
        snprintf(buf, sizeof buf, "axb_tile_nonempty_noalloc: ytiles * xtiles * T = %llu exceeds AXB_NOALLOC_MAX_TILE_FRAMES = %d "
                 "(use axb_tile_nonempty, or fewer frames per call)", n_flags, AXB_NOALLOC_MAX_TILE_FRAMES);
        return bad_arg(c, buf);
    }
    AXB_ON_DEVICE(c);
    AXB_CUDA(c, axb::launch_k6_nonempty_static(X_dev, T, C, H, W, tilesize, nonempty_dev, (cudaStream_t)stream));
    c->launches += 2;
    return AXB_OK;
}

int axb_gather_frametiles(axb_ctx* c, const float* X_dev, int T, int C, int H, int W, int tilesize,
                          const int32_t* tile_yx_dev, int ntiles, const int32_t* frames_dev, int n_frames,
                          int temporal_context, int chan_lo, int chan_n, float* out_dev, void* stream) {
    if (!c) return AXB_ERR_BAD_ARG;
    if (T <= 0 || C < 1 || H <= 0 || W <= 0 || tilesize < 1 || ntiles < 0 || n_frames < 0 || temporal_context < 0)
        return bad_arg(c, "axb_gather_frametiles: bad size");
    if (chan_lo < 0 || chan_n < 1 || chan_lo + chan_n > C) return bad_arg(c, "axb_gather_frametiles: channel range outside [0, C)");
    if (ntiles == 0 || n_frames == 0) return AXB_OK;
    if (!X_dev || !tile_yx_dev || !frames_dev || !out_dev) return bad_arg(c, "axb_gather_frametiles: NULL pointer");
    AXB_ON_DEVICE(c);
    axb::launch_k6_gather(X_dev, T, C, H, W, tilesize, tile_yx_dev, ntiles, frames_dev, n_frames, temporal_context, chan_lo,
                          chan_n, out_dev, c->status_dev + 1, c->num_sms, (cudaStream_t)stream);
    c->launches += 1;
    AXB_CUDA(c, cudaGetLastError());
    return AXB_OK;
}

int axb_stitched_frame(axb_ctx* c, const float* X_dev, int T, int C, int H, int W, int tilesize, const uint8_t* tile_keep_dev,
                       int frame, int chan_lo, int chan_n, float* out_dev, void* stream) {
    if (!c) return AXB_ERR_BAD_ARG;
    if (T <= 0 || C < 1 || H <= 0 || W <= 0 || tilesize < 1 || frame < 0 || frame >= T) return bad_arg(c, "axb_stitched_frame: bad size or frame outside [0, T)");
    if (chan_lo < 0 || chan_n < 1 || chan_lo + chan_n > C) return bad_arg(c, "axb_stitched_frame: channel range outside [0, C)");
    if (!X_dev || !tile_keep_dev || !out_dev) return bad_arg(c, "axb_stitched_frame: NULL pointer");
    AXB_ON_DEVICE(c);
    axb::launch_k6_frame(X_dev, C, H, W, tilesize, tile_keep_dev, frame, chan_lo, chan_n, out_dev, c->num_sms, (cudaStream_t)stream);
    c->launches += 1;
    AXB_CUDA(c, cudaGetLastError());
    return AXB_OK;
}

int axb_status(axb_ctx* c, void* stream, int64_t info[3]) {
    if (!c) return AXB_ERR_BAD_ARG;
    AXB_ON_DEVICE(c);
    unsigned long long key = ~0ull, flags = 0ull;
    int rc = status_fetch(c, (cudaStream_t)stream, &key, &flags);
    if (rc != AXB_OK) return rc;
    std::vector<int> pre;
    return status_decode(c, key, flags, info, [&](long long item, int64_t* pair, int64_t* src) {
        const int n = c->last_n_pairs;
        if (n <= 0 || !c->prefix_dev) return;
        pre.resize((size_t)n + 1);
        if (cudaMemcpy(pre.data(), c->prefix_dev + c->prefix_cap, ((size_t)n + 1) * sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess) return;
        int lo = (int)(std::upper_bound(pre.begin(), pre.begin() + n, (int)item) - pre.begin()) - 1;
        lo = std::max(lo, 0);
        *pair = lo; *src = item - pre[lo];
    });
}

}  // extern "C"
