// bp_handle.h — the opaque handle behind the C ABI of include/bp_b200.h and the helpers shared by the host translation
// units of libbp_b200.so (bp_b200.cu: ABI entry points, read-back; bp_process.cu: the mask ingest / launch pipeline).
#ifndef BP_HANDLE_H
#define BP_HANDLE_H
#include <cuda_runtime.h>

#include <utility>
#include <vector>

#include "bp_b200.h"
#include "bp_internal.h"
#include "bp_types.h"

// ------------------------------------------------------------------------------------------
// handle
// ------------------------------------------------------------------------------------------
struct bp_handle {
    int device = 0;
    int mode = 0;
    int shape[3] = {0, 0, 0};
    float default_value = 0.f;
    GridInfo g{};
    cudaStream_t stream = nullptr;
    void *grid = nullptr;  // int8 or float32, [ni, ny, nzp]
    size_t grid_bytes = 0;
    void *pool = nullptr;  // tiled maps of the current batch
    size_t pool_bytes = 0;
    cudaStream_t cstream = nullptr;  // H2D copies of host masks, overlapped with the main stream
    void *stage[2] = {nullptr, nullptr};  // double-buffered device staging for raw host masks
    size_t stage_bytes[2] = {0, 0};
    cudaEvent_t ev_copied[2] = {nullptr, nullptr}, ev_stage_free[2] = {nullptr, nullptr};
    void *outbuf = nullptr;  // unpadded int32 / float32 readback staging
    size_t out_bytes = 0;
    float *lut_d = nullptr;  // 256 floats
    void *views_d[2] = {nullptr, nullptr};  // global-memory copy of a launch's ViewParams (carve sparse phase)
    float lut_h[256];        // the table as uploaded (inversion applied)
    dim3 launch_grid{1, 1, 1};  // carve / average launches (super-block order, see GridInfo)
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev_kernel, ev_ingest;
    std::vector<cudaEvent_t> ev_free;
    float kernel_ms = 0.f, ingest_ms = 0.f;
    int64_t launches = 0;
    int64_t h2d_bytes = 0;  // mask bytes copied host -> device since creation (bp_ingest_bytes)
    // block cull of the carve path (bp_kernels.cu: cull_blocks_kernel / carve_items_kernel); NULL when the handle is not
    // a carving handle or its block grid does not fit the 16-bit block coordinates of the work list
    CullGrid cg{};
    uint8_t *blk_dead = nullptr;        // one byte per cull block, persistent until bp_clear
    bool lazy_unseen = false;           // carving: bp_clear has NOT written the state yet ("all unseen"): the cull's first
                                        // pass writes every block, anything else that touches the state calls settle_seed
    uint2 *blk_live = nullptr;          // work list of one launch
    uint2 *blk_cand = nullptr;          // survivors of the cull's first pass
    uint8_t *patch_cls = nullptr;       // one byte per work item: how alive the patch is after the first views
    uint32_t *seg_items = nullptr;      // one word per work item: the live items carve_segments_kernel draws from
    uint32_t *cull_counters = nullptr;  // {live blocks, next work item, blocks killed, -}
    CullView *cull_views_d = nullptr;   // kCullViews entries
    void *sat_pool = nullptr;           // summed-area tables (+ scratch) of one launch
    size_t sat_bytes = 0;
    int items_ctas = 0;                 // SM count (persistent grids of carve_items / carve_segments = this x occupancy)
    bool count_work = false;            // BP_COUNT_WORK=1: count the projections the carve kernels evaluate
    unsigned long long *work_d = nullptr;  // carving: projections evaluated; averaging: voxel-views skipped
    // view skipping of the averaging path (bp_kernels.cu: avg_skip_kernel); NULL for carving handles
    uint32_t *avg_skip_d = nullptr;     // kSkipWords words per skip block
    int avg_nbj = 0, avg_nbk = 0, avg_lgk = 4;  // skip blocks along j / k, log2 of their extent along k
    uint32_t avg_nblocks = 0;
    int64_t avg_nominal_vv = 0;         // voxel-views of the averaging launches since the last bp_work_stats
    // speculative read-back (bp_set_readback_target; carving, host masks in several chunks)
    void *spec_target = nullptr;      // pinned host volume the next bp_get_values will be given
    void *spec_target_dev = nullptr;  // the same memory as the device addresses it
    cudaStream_t rstream = nullptr;   // the early D2H runs here, under the remaining uploads
    cudaEvent_t ev_snap = nullptr, ev_spec_done = nullptr;
    void *snap = nullptr;             // int8 state as it was when the early D2H was taken
    size_t snap_bytes = 0;
    unsigned long long *spec_count_d = nullptr;
    bool spec_pending = false;        // the target holds (or is being filled with) the expansion of `snap`
    int64_t spec_rows = 0, spec_rows_total = 0;  // rows patched / rows compared by the last bp_get_values
};

#define CU(call)                                                                             \
    do {                                                                                     \
        cudaError_t e_ = (call);                                                             \
        if (e_ != cudaSuccess)                                                               \
            return bp_fail(BP_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                           __FILE__, __LINE__);                                              \
    } while (0)

namespace bph {

// grow a device buffer (frees the old one after syncing `s`); *have is updated
int ensure(void **p, size_t *have, size_t want, cudaStream_t s);
// timing events are pooled per handle
int get_event(bp_handle *h, cudaEvent_t *ev);
// fold the recorded (start, stop) pairs into kernel_ms / ingest_ms; syncs the main stream
int fold_events(bp_handle *h);
size_t elem_size(int dtype);
// int32 values the int8 carve states expand to (kernels/backprojection.c:79-83 on the device coding)
void carve_values(const bp_handle *h, int *v_carved, int *v_seen, int *v_unseen);
// global index of a handle's local plane (see GridInfo)
int global_plane_host(const GridInfo &G, int li);
// speculative read-back: first half (bp_process.cu calls it after the chosen chunk), and the wait that frees outbuf
int start_spec_readback(bp_handle *h);
int spec_settle(bp_handle *h);
// lazy clear: write the pending "all unseen" seed now (no-op when nothing is pending)
int settle_seed(bp_handle *h);

// carving: 32x32-pixel bit tiles of one view (32 words each) in the pool
struct TiledDims {
    int tpr, tiles_y;
    size_t elems;  // uint32 words
};
TiledDims tiled_dims(int H, int W);

// bp_process.cu
int process_impl(bp_handle *h, int32_t n, const float *cams, const void *const *masks, const int32_t *hw, int mask_dtype,
                 uint32_t flags, const float *lut, void *pack_dst, int64_t pack_stride);

}  // namespace bph
#endif  // BP_HANDLE_H
