// bp_process.cu — the host pipeline behind bp_process_views / bp_pack_views: view ordering, staging of host masks
// (double-buffered H2D on a copy stream), mask ingest launches and the fused carve / average launches.
// Replaces the per-view upload + launch + finish() loop of plant3dvision/cl.py:190-227, :282-303 of the reference.
#include "bp_handle.h"
#include "bp_geom.h"
#include "bp_kernels.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <vector>

namespace {
using namespace bph;

// View rectangles: copy (host masks) and pack only the rectangle of every view that the handle's planes project into
// (bp_box_rect: pixels outside it are never sampled by any voxel of the handle).  On by default since round 2 (C5: the
// 1024^3 box fills 61.5 % of a 6000 x 4000 view; e2e 185 -> 146 ms, profiles/r02_bench_view_rect.json); BP_VIEW_RECT=0
// switches it off.  Read at every call so that a test can switch it.
bool view_rect_enabled() {
    const char *e = std::getenv("BP_VIEW_RECT");
    return !e || std::atoi(e) != 0;
}

// BP_CARVE_CULL=0 switches the block cull off: every launch then sweeps every patch of the grid with carve_kernel
// (the round-1 path; kept for A/B measurements and as the path for non-finite / huge parameters).  Default on.
bool carve_cull_enabled() {
    const char *e = std::getenv("BP_CARVE_CULL");
    return !e || std::atoi(e) != 0;
}

int env_int(const char *name, int dflt) {
    const char *e = std::getenv(name);
    return e ? std::atoi(e) : dflt;
}

// BP_ITEM_GRAB = 1 | 2 | 4 | 8 | 16: consecutive patches of one block a warp of carve_items_kernel takes per atomic (A/B)
int item_grab() {
    const int g = env_int("BP_ITEM_GRAB", 1);
    return (g == 1 || g == 2 || g == 4 || g == 8 || g == 16) ? g : 1;
}

// BP_LIVE_ORDER=0: work list of the carve in row-major block order instead of Morton order (A/B; see compact_live_kernel).
bool live_order_morton() {
    const char *e = std::getenv("BP_LIVE_ORDER");
    return !e || std::atoi(e) != 0;
}

// Two-tier masks (device-resident uint8 masks only): the first BP_PACK_HEAD views of a launch, in visiting order, are
// bit-packed; the later ones are read in place by the few voxels that survive the first (see bp_kernels.cu).
// Negative = pack every view.  Default: see below.
int pack_head_views(const bp_handle *h) {
    const char *e = std::getenv("BP_PACK_HEAD");
    if (e) {
        const int v = std::atoi(e);
        return v < 0 ? kMaxViews : v;
    }
    // 16 views on a large handle, whatever its share of the grid: since the later views run as view segments
    // (carve_segments_kernel) the raw reads cost the kernels nothing measurable (C5, one GPU: kernels 7.07 ms with 16
    // views packed against 7.12 with all 240, and 0.06 instead of 0.73 ms of packing; with 32 / 64 packed views both
    // the pack and the tile tables of the cull grow: profiles/r02_packhead_v2.txt).  A handle with few blocks packs
    // only the views its cull looks at (launch_cull_carve): a C5 rank of 8 goes from 1.205 to 1.162 ms per step.
    return (h->blk_dead && h->cg.n_blocks < 100000u) ? kCullFirstViews : 16;
}

// bit-pack ONE view whose raw mask already sits in device memory (general path: any dtype / width)
int pack_view(bp_handle *h, uint32_t *pool, const void *src, int H, int W, int dtype, uint32_t flags, size_t off,
              const TiledDims &d) {
    bpk::pack_bits(h->stream, dtype, src, pool + off, H, W, d.tpr, d.tiles_y * 32, flags);
    CU(cudaGetLastError());
    h->launches++;
    return BP_OK;
}

// Carving is order-independent (kernels/backprojection.c:65-83: carved is absorbing, seen is
// set once), so the fused kernel is free to visit the views of a batch in any order.  Visiting
// them in farthest-point order of their optical axes (r6 r7 r8; a view and its antipode count
// as the same direction, their silhouettes nearly coincide) carves most of the empty space in
// the first handful of views, after which whole warps leave the view loop.  Averaging keeps the
// caller's order: its float32 sum is order-dependent (cl.py:223-226 accumulates view by view).
void spread_order(const float *cams, int n, std::vector<int> &order) {
    order.resize(n);
    if (n == 0) return;
    std::vector<float> far(n, 2.0f);  // min over chosen of (1 - |cos|); 2 = nothing chosen yet
    std::vector<char> used(n, 0);
    int cur = 0;
    for (int q = 0; q < n; ++q) {
        order[q] = cur;
        used[cur] = 1;
        const float *a = cams + 16 * (size_t)cur + 10;
        const float na = std::sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
        int best = -1;
        float best_d = -1.0f;
        for (int v = 0; v < n; ++v) {
            if (used[v]) continue;
            const float *b = cams + 16 * (size_t)v + 10;
            const float nb = std::sqrt(b[0] * b[0] + b[1] * b[1] + b[2] * b[2]);
            float c = (a[0] * b[0] + a[1] * b[1] + a[2] * b[2]) / (na * nb);
            if (!(c == c)) c = 1.0f;
            const float d = 1.0f - std::fabs(c);
            if (d < far[v]) far[v] = d;
            if (far[v] > best_d) {
                best_d = far[v];
                best = v;
            }
        }
        cur = best;
    }
}

int check_views_args(bp_handle *h, int dtype, uint32_t flags, const float *lut) {
    if (elem_size(dtype) == 0) return bp_fail(BP_EINVAL, "unknown mask_dtype %d", dtype);
    if (h->mode == BP_MODE_AVERAGING) {
        if (dtype != BP_MASK_U8 && dtype != BP_MASK_F32)
            return bp_fail(BP_EINVAL, "averaging takes U8 (+lut) or F32 maps; convert other types on the host");
        if (dtype == BP_MASK_U8 && !lut) return bp_fail(BP_EINVAL, "averaging with a U8 map needs a 256-entry lut");
        if (dtype == BP_MASK_F32 && (flags & (BP_FLAG_INVERT | BP_FLAG_BOOL_INVERT)))
            return bp_fail(BP_EINVAL, "invert is not defined for float maps (np.invert raises)");
    } else if (dtype == BP_MASK_F32 && (flags & (BP_FLAG_INVERT | BP_FLAG_BOOL_INVERT))) {
        return bp_fail(BP_EINVAL, "invert is not defined for float masks (np.invert raises)");
    }
    return BP_OK;
}


// One carving launch through the block cull (bp_kernels.cu): summed-area tables of the first bit-packed views, the cull
// (persistent dead flags, zero fill of the blocks it kills, work list of the survivors) and the persistent kernel that
// carves the patches of the surviving blocks.  `order` = this launch's views as indices into hw / cams.
int launch_cull_carve(bp_handle *h, const Batch &batch, const int32_t *hw, const int *order, const float *cams,
                      unsigned long long *work) {
    static thread_local SatJobs sat;
    static thread_local CullView cv[kCullViews];
    int n_cull = 0;
    // How many packed views the cull looks at (BP_CULL_VIEWS).  The first three write off 9 blocks in 10; the second
    // pass over the survivors saves the kernels 0.2 ms on the whole C5 grid and costs 0.05 ms of latency whatever the
    // grid: a handle with few blocks (a plane-sharded rank, a small scan) stops after the first
    // (profiles/r02_cull_views_ab.txt: C5 rank of 8, 65 k blocks, 1.269 -> 1.220 ms; C2, 26 k blocks, 0.410 -> 0.369;
    // but C1, 154 k blocks, 0.508 -> 0.677 without the second pass).
    const int cull_dflt = h->cg.n_blocks >= 100000u ? kCullViews : kCullFirstViews;
    const int cull_cap = std::min(kCullViews, std::max(0, env_int("BP_CULL_VIEWS", cull_dflt)));
    while (n_cull < batch.n && n_cull < cull_cap && batch.v[n_cull].kind == 0) ++n_cull;  // the packed head
    size_t need = 0;
    int max_tiles = 0;
    for (int q = 0; q < n_cull; ++q) {
        const int v = order[q];
        const TiledDims d = tiled_dims(hw[2 * v], hw[2 * v + 1]);
        need += (((size_t)(d.tiles_y + 1) * (d.tpr + 1) * 4 + 255) & ~(size_t)255) + (((size_t)d.tiles_y * d.tpr + 255) & ~(size_t)255);
        if (d.tpr * d.tiles_y > max_tiles) max_tiles = d.tpr * d.tiles_y;
    }
    int rc = ensure(&h->sat_pool, &h->sat_bytes, need, h->stream);
    if (rc) return rc;
    char *pool = (char *)h->sat_pool;
    sat.n = n_cull;
    for (int q = 0; q < n_cull; ++q) {
        const int v = order[q];
        const TiledDims d = tiled_dims(hw[2 * v], hw[2 * v + 1]);
        SatJob &J = sat.j[q];
        J.tiles = (const uint32_t *)batch.v[q].map;
        J.sat = (uint32_t *)pool;
        pool += ((size_t)(d.tiles_y + 1) * (d.tpr + 1) * 4 + 255) & ~(size_t)255;
        J.any = (uint8_t *)pool;
        pool += ((size_t)d.tiles_y * d.tpr + 255) & ~(size_t)255;
        J.tpr = d.tpr;
        J.tiles_y = d.tiles_y;
        CullView &V = cv[q];
        for (int a = 0; a < 16; ++a) V.cam[a] = cams[16 * (size_t)v + a];
        V.sat = J.sat;
        V.tpr1 = d.tpr + 1;
        V.W = hw[2 * v + 1];
        V.H = hw[2 * v];
        V.pad[0] = V.pad[1] = V.pad[2] = 0;
    }
    if (n_cull > 0) {
        bpk::tile_sat(h->stream, sat, max_tiles);
        CU(cudaGetLastError());
        h->launches += 2;
        CU(cudaMemcpyAsync(h->cull_views_d, cv, sizeof(CullView) * (size_t)n_cull, cudaMemcpyHostToDevice, h->stream));
    }
    CU(cudaMemsetAsync(h->cull_counters, 0, 8 * sizeof(uint32_t), h->stream));
    bpk::cull_blocks(h->stream, (int8_t *)h->grid, h->g, h->cg, h->cull_views_d, n_cull, kCullFirstViews, live_order_morton() ? 1 : 0,
                     env_int("BP_LPT", 1), h->blk_dead, h->blk_cand, h->blk_live, h->cull_counters, h->lazy_unseen);
    h->lazy_unseen = false;  // the first pass has written every block of the grid
    CU(cudaGetLastError());
    h->launches += (n_cull > kCullFirstViews) ? 2 : 1;
    // the first views over every surviving patch; the remaining views of the patches still alive, cut into segments
    // that any warp can take (BP_VIEW_SPLIT: where to cut; >= the view count: no second kernel)
    int v_split = std::max(1, env_int("BP_VIEW_SPLIT", 16));
    // a short remainder is not worth three more launches (C1: 18 views, 0.77 ms with segments for the last two, 0.58 without)
    if (batch.n - v_split < env_int("BP_VIEW_SPLIT_MIN_REST", 16)) v_split = batch.n;
    bpk::carve_items(h->items_ctas, h->stream, (int8_t *)h->grid, batch, (const ViewParams *)h->views_d[0], h->g, h->blk_live,
                     h->cg.n_blocks, item_grab(), v_split, h->cull_counters, work);
    h->launches += 1;  // (the caller counts the last carve launch itself)
    if (batch.n > v_split) {
        CU(cudaGetLastError());
        h->launches += 2;  // classify_items_kernel, decode_items_kernel
        bpk::carve_segments(h->items_ctas, h->stream, (int8_t *)h->grid, batch, (const ViewParams *)h->views_d[0], h->g, h->blk_live,
                            h->cg.n_blocks, v_split, h->patch_cls, h->seg_items, h->cull_counters, work);
    } else {
        h->launches -= 1;
    }
    return BP_OK;
}

// BP_AVG_SKIP=0 switches the view skipping of the averaging path off (A/B).  Default on.
bool avg_skip_enabled() {
    const char *e = std::getenv("BP_AVG_SKIP");
    return !e || std::atoi(e) != 0;
}

// View skipping for one averaging launch over uint8 maps (bp_kernels.cu): per view the summed-area table of "tile holds a
// byte != B", then one skip bit per (block, view).  B = the raw byte whose table entry is the value of a zero pixel
// (255 under BP_FLAG_INVERT, else 0): the commonest byte of a probability map.  Returns skip == NULL (nothing skipped)
// when a map does not have 16-byte rows.
int launch_avg_skip(bp_handle *h, const Batch &batch, const int32_t *hw, const int *order, const float *cams, uint32_t flags,
                    const uint32_t **skip_out, float *c0_out) {
    static thread_local SatJobs sat;
    static thread_local MapAnyJobs any;
    static thread_local CullView cv[kMaxViews];
    const int n = batch.n;
    size_t need = 0;
    int tiles_y_max = 0;
    any.tpr_max = 0;
    for (int q = 0; q < n; ++q) {
        const int v = order[q], H = hw[2 * v], W = hw[2 * v + 1];
        if ((W % 16) != 0 || ((uintptr_t)batch.v[q].map % 16) != 0) return BP_OK;
        const TiledDims d = tiled_dims(H, W);
        need += (((size_t)(d.tiles_y + 1) * (d.tpr + 1) * 4 + 255) & ~(size_t)255) + (((size_t)d.tiles_y * d.tpr + 255) & ~(size_t)255);
        if (d.tiles_y > tiles_y_max) tiles_y_max = d.tiles_y;
        if (d.tpr > any.tpr_max) any.tpr_max = d.tpr;
    }
    int rc = ensure(&h->sat_pool, &h->sat_bytes, need, h->stream);
    if (rc) return rc;
    const uint32_t B = (flags & BP_FLAG_INVERT) ? 255u : 0u;
    any.n = sat.n = n;
    any.b4 = B * 0x01010101u;
    any.pad = 0;
    char *pool = (char *)h->sat_pool;
    for (int q = 0; q < n; ++q) {
        const int v = order[q], H = hw[2 * v], W = hw[2 * v + 1];
        const TiledDims d = tiled_dims(H, W);
        SatJob &J = sat.j[q];
        J.tiles = nullptr;
        J.sat = (uint32_t *)pool;
        pool += ((size_t)(d.tiles_y + 1) * (d.tpr + 1) * 4 + 255) & ~(size_t)255;
        J.any = (uint8_t *)pool;
        pool += ((size_t)d.tiles_y * d.tpr + 255) & ~(size_t)255;
        J.tpr = d.tpr;
        J.tiles_y = d.tiles_y;
        MapAnyJob &A = any.j[q];
        A.src = (const uint8_t *)batch.v[q].map;
        A.any = J.any;
        A.H = H;
        A.W = W;
        CullView &V = cv[q];
        for (int a = 0; a < 16; ++a) V.cam[a] = cams[16 * (size_t)v + a];
        V.sat = J.sat;
        V.tpr1 = d.tpr + 1;
        V.W = W;
        V.H = H;
        V.pad[0] = V.pad[1] = V.pad[2] = 0;
    }
    bpk::map_any(h->stream, any, tiles_y_max);
    bpk::tile_sat_only(h->stream, sat);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(h->cull_views_d, cv, sizeof(CullView) * (size_t)n, cudaMemcpyHostToDevice, h->stream));
    CU(cudaMemsetAsync(h->avg_skip_d, 0, (size_t)h->avg_nblocks * kSkipWords * sizeof(uint32_t), h->stream));
    bpk::avg_skip(h->stream, h->g, h->avg_nbj, h->avg_nbk, h->avg_lgk, h->avg_nblocks, h->cull_views_d, n, h->avg_skip_d, h->work_d);
    CU(cudaGetLastError());
    h->launches += 3;
    *skip_out = h->avg_skip_d;
    *c0_out = h->lut_h[B];
    return BP_OK;
}

}  // namespace

namespace bph {

// Shared body of bp_process_views and bp_pack_views.  pack_dst != NULL: ingest only (upload + bit-pack,
// view q at pack_dst + q * pack_stride, caller's order, no carve launch).
int process_impl(bp_handle *h, int32_t n, const float *cams, const void *const *masks, const int32_t *hw, int mask_dtype,
                 uint32_t flags, const float *lut, void *pack_dst, int64_t pack_stride) {
    if (!h) return bp_fail(BP_EINVAL, "null handle");
    if (n < 0) return bp_fail(BP_EINVAL, "negative view count");
    if (n == 0) return BP_OK;
    const bool pack_only = pack_dst != nullptr;
    if ((!cams && !pack_only) || !masks || !hw) return bp_fail(BP_EINVAL, "null argument");
    int rc = check_views_args(h, mask_dtype, flags, lut);
    if (rc) return rc;
    const bool carving = h->mode == BP_MODE_CARVING;
    const bool packed_in = (flags & BP_FLAG_PACKED_MASKS) != 0;
    if ((pack_only || packed_in) && !carving) return bp_fail(BP_ESTATE, "bit-packed masks belong to carving handles");
    if (pack_only && packed_in) return bp_fail(BP_EINVAL, "masks are already packed");
    if (packed_in) flags |= BP_FLAG_DEVICE_MASKS;
    size_t max_raw = 0;
    for (int v = 0; v < n; ++v) {
        if (!masks[v]) return bp_fail(BP_EINVAL, "mask %d is null", v);
        if (hw[2 * v] <= 0 || hw[2 * v + 1] <= 0 || hw[2 * v] >= (1 << 24) || hw[2 * v + 1] >= (1 << 24))
            return bp_fail(BP_EINVAL, "mask %d has bad size %d x %d", v, hw[2 * v], hw[2 * v + 1]);
        if (carving && (size_t)hw[2 * v] * (size_t)hw[2 * v + 1] >= ((size_t)1 << 35))
            return bp_fail(BP_EINVAL, "mask %d is too large (%d x %d): tile word offsets are 32-bit", v, hw[2 * v], hw[2 * v + 1]);
        if (!carving && (size_t)hw[2 * v] * (size_t)hw[2 * v + 1] >= ((size_t)1 << 31))
            return bp_fail(BP_EINVAL, "map %d is too large (%d x %d): element offsets are 32-bit", v, hw[2 * v], hw[2 * v + 1]);
        const size_t raw = ((size_t)hw[2 * v] * hw[2 * v + 1] * elem_size(mask_dtype) + 255) & ~(size_t)255;
        if (raw > max_raw) max_raw = raw;
    }
    CU(cudaSetDevice(h->device));
    if (!pack_only && (h->g.ni == 0 || h->g.ny == 0 || h->g.nz == 0)) return BP_OK;  // empty grid: nothing to do
    if (!carving && mask_dtype == BP_MASK_U8) {
        // np.invert on the map (cl.py:300-301) == the same bytes read through a permuted table
        for (int b = 0; b < 256; ++b) {
            int src = b;
            if (flags & BP_FLAG_BOOL_INVERT) src = (b == 0) ? 1 : 0;
            else if (flags & BP_FLAG_INVERT) src = 255 - b;
            h->lut_h[b] = lut[src];
        }
        CU(cudaMemcpyAsync(h->lut_d, h->lut_h, 256 * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    }
    if (h->ev_kernel.size() > 2048) {
        rc = fold_events(h);
        if (rc) return rc;
    }
    const bool host_masks = !(flags & BP_FLAG_DEVICE_MASKS);

    // visiting order (carving only; see spread_order) and chunking.  Device-resident masks go
    // kMaxViews at a time; host masks in chunks that fit a staging buffer, so that the H2D copy
    // of chunk c+1 (copy stream) overlaps the pack + carve/average of chunk c (main stream).
    std::vector<int> order;
    if (carving && !pack_only) {
        spread_order(cams, n, order);
    } else {
        order.resize(n);
        for (int v = 0; v < n; ++v) order[v] = v;
    }
    int chunk = kMaxViews;
    if (host_masks) {
        size_t stage_bytes = kStageBytes;  // BP_STAGE_BYTES: smaller chunks, for tests of the chunk pipeline
        if (const char *e = std::getenv("BP_STAGE_BYTES")) {
            const long long v = std::atoll(e);
            if (v > 0) stage_bytes = (size_t)v;
        }
        const size_t per = stage_bytes / max_raw;
        chunk = (int)(per < 1 ? 1 : (per > (size_t)kMaxViews ? (size_t)kMaxViews : per));
        const int n_chunks = (n + chunk - 1) / chunk;
        for (int s = 0; s < (n_chunks > 1 ? 2 : 1); ++s) {
            const size_t want = max_raw * (size_t)(chunk < n ? chunk : n);
            if (h->stage_bytes[s] < want) {
                CU(cudaStreamSynchronize(h->stream));
                CU(cudaStreamSynchronize(h->cstream));
                if (h->stage[s]) CU(cudaFree(h->stage[s]));
                h->stage[s] = nullptr;
                h->stage_bytes[s] = 0;
                cudaError_t e = cudaMalloc(&h->stage[s], want);
                if (e != cudaSuccess)
                    return bp_fail(BP_ENOMEM, "cudaMalloc(%zu) for mask staging failed: %s", want, cudaGetErrorString(e));
                h->stage_bytes[s] = want;
            }
        }
    }

    static thread_local Batch batch;     // 22.5 KB, filled per launch
    static thread_local PackJobs jobs;   // 12 KB
    std::vector<TiledDims> dims;
    std::vector<const void *> src;
    std::vector<char> raw_view;
    // BP_VIEW_RECT: per view, the rectangle the handle's planes project into; only that part of a mask is copied
    // (host masks) and packed.  Views that go through the general packer (pack_view) keep the whole image.
    // Packing for other ranks (bp_pack_views_cams): the tiles go to every GPU, so the rectangle is the WHOLE grid's.
    const bool use_rect = view_rect_enabled() && cams && !packed_in && (pack_only || h->g.ni > 0);
    const int rect_i_lo = (use_rect && !pack_only) ? global_plane_host(h->g, 0) : 0;
    const int rect_i_hi = !use_rect ? 0 : pack_only ? h->shape[0] - 1 : global_plane_host(h->g, h->g.ni - 1);
    std::vector<ViewRect> rects;
    auto rect_of = [&](int v) {
        const int H = hw[2 * v], W = hw[2 * v + 1];
        const ViewRect whole = {0, 0, W, H};
        // carving: only the batched uint8 packer knows about rectangles (staged copies are 256-byte aligned)
        const bool can = use_rect && (!carving || (mask_dtype == BP_MASK_U8 && (W % 16) == 0 &&
                                                   (host_masks || ((uintptr_t)masks[v] % 16) == 0)));
        return can ? bp_box_rect(rect_i_lo, rect_i_hi, 0, h->g.ny - 1, 0, h->g.nz - 1, h->g.ox, h->g.oy, h->g.oz, h->g.vs,
                                 cams + 16 * (size_t)v, H, W)
                   : whole;
    };
    // Speculative read-back (bp_set_readback_target): take the early copy of the volume after the chunk that leaves
    // about as much upload time as the read-back needs (both cross PCIe, in opposite directions), but not before
    // the first 16 views (farthest-point order) have removed the bulk of the empty space.
    int ci_snap = -1;
    if (carving && host_masks && !pack_only && h->spec_target && chunk < n) {
        const int n_chunks = (n + chunk - 1) / chunk;
        double h2d = 0.0;
        for (int v = 0; v < n; ++v) {
            const ViewRect R = rect_of(v);
            h2d += (double)(R.x1 - R.x0) * (double)(R.y1 - R.y0) * (double)elem_size(mask_dtype);
        }
        const double d2h = 4.0 * (double)h->g.ni * h->g.ny * h->g.nz;
        double frac = h2d > 0.0 ? 1.0 - d2h / h2d : 0.0;
        if (const char *e = std::getenv("BP_SPEC_AT")) frac = std::atof(e);  // experiments: fraction of the views
        int before = (int)std::ceil(frac * n);
        if (before < 16) before = 16;
        const int ci = (before + chunk - 1) / chunk - 1;  // the chunk whose carve launch sees view number `before`
        if (ci < n_chunks - 1) ci_snap = ci;                  // else nothing would be left to hide the copy behind
    }
    int ci = 0;
    for (int v0 = 0; v0 < n; v0 += chunk, ++ci) {
        const int nb = (n - v0 < chunk) ? (n - v0) : chunk;
        const int s = ci & 1;
        dims.resize(nb);
        src.resize(nb);
        rects.resize(nb);
        for (int q = 0; q < nb; ++q) rects[q] = rect_of(order[v0 + q]);
        // two-tier masks: which views of this launch are read raw (device-resident uint8 masks past the packed head)
        raw_view.assign(nb, 0);
        if (carving && !pack_only && !packed_in && !host_masks && mask_dtype == BP_MASK_U8) {
            const int head = pack_head_views(h);
            for (int q = head; q < nb; ++q) {
                const int v = order[v0 + q];
                raw_view[q] = ((size_t)hw[2 * v] * (size_t)hw[2 * v + 1] < ((size_t)1 << 32)) ? 1 : 0;  // 32-bit byte offsets
            }
        }
        if (carving) {
            size_t total = 0;
            for (int q = 0; q < nb; ++q) {
                const int v = order[v0 + q];
                dims[q] = tiled_dims(hw[2 * v], hw[2 * v + 1]);
                if (!raw_view[q]) total += dims[q].elems;
                if (pack_only && (int64_t)(dims[q].elems * 4) > pack_stride)
                    return bp_fail(BP_EINVAL, "view_stride %lld is smaller than the %zu bytes view %d packs to",
                                (long long)pack_stride, dims[q].elems * 4, v);
            }
            if (!pack_only && !packed_in) {
                rc = ensure(&h->pool, &h->pool_bytes, total * 4, h->stream);
                if (rc) return rc;
            }
        }
        uint32_t *const pool_base = pack_only ? (uint32_t *)pack_dst : (uint32_t *)h->pool;
        if (host_masks) {
            // the staging buffer is free once the work of the chunk that used it last has run
            CU(cudaStreamWaitEvent(h->cstream, h->ev_stage_free[s], 0));
            size_t so = 0;
            for (int q = 0; q < nb; ++q) {
                const int v = order[v0 + q];
                const size_t es = elem_size(mask_dtype);
                const size_t raw = (size_t)hw[2 * v] * hw[2 * v + 1] * es;
                char *dst = (char *)h->stage[s] + so;
                const ViewRect &R = rects[q];
                const size_t Wq = (size_t)hw[2 * v + 1];
                const size_t first = ((size_t)R.y0 * Wq + (size_t)R.x0) * es;  // same layout, rows / columns left out
                if (R.x0 == 0 && (size_t)R.x1 == Wq) {
                    const size_t nbytes = (size_t)(R.y1 - R.y0) * Wq * es;
                    CU(cudaMemcpyAsync(dst + first, (const char *)masks[v] + first, nbytes, cudaMemcpyHostToDevice, h->cstream));
                    h->h2d_bytes += (int64_t)nbytes;
                } else {
                    CU(cudaMemcpy2DAsync(dst + first, Wq * es, (const char *)masks[v] + first, Wq * es,
                                         (size_t)(R.x1 - R.x0) * es, (size_t)(R.y1 - R.y0), cudaMemcpyHostToDevice, h->cstream));
                    h->h2d_bytes += (int64_t)((size_t)(R.x1 - R.x0) * es * (size_t)(R.y1 - R.y0));
                }
                src[q] = dst;
                so += (raw + 255) & ~(size_t)255;
            }
            CU(cudaEventRecord(h->ev_copied[s], h->cstream));
            CU(cudaStreamWaitEvent(h->stream, h->ev_copied[s], 0));
        } else {
            for (int q = 0; q < nb; ++q) src[q] = masks[order[v0 + q]];
        }
        cudaEvent_t i0, i1, k0 = nullptr, k1 = nullptr;
        if ((rc = get_event(h, &i0)) || (rc = get_event(h, &i1))) return rc;
        if (!pack_only && ((rc = get_event(h, &k0)) || (rc = get_event(h, &k1)))) return rc;
        CU(cudaEventRecord(i0, h->stream));
        size_t off = 0;
        batch.n = nb;
        {   // the fast division path assumes no intermediate can overflow (see DivGuard)
            auto tame = [](float f) { return std::fabs(f) < 0x1p40f; };  // false for NaN / Inf too
            const GridInfo &G = h->g;
            const float ext[3] = {(float)h->shape[0] * G.vs, (float)G.ny * G.vs, (float)G.nz * G.vs};
            bool all = tame(G.ox) && tame(G.oy) && tame(G.oz) && tame(G.vs) && tame(ext[0]) && tame(ext[1]) && tame(ext[2]);
            for (int q = 0; q < nb && all && !pack_only; ++q)
                for (int a = 0; a < 16; ++a) all = all && tame(cams[16 * (size_t)order[v0 + q] + a]);
            batch.force_exact = all ? 0 : 1;
        }
        jobs.n = 0;
        int jw = 0, jt = 0;
        for (int q = 0; q < nb; ++q) {
            const int v = order[v0 + q];
            const int H = hw[2 * v], W = hw[2 * v + 1];
            ViewParams &P = batch.v[q];
            if (pack_only) off = (size_t)(v0 + q) * (size_t)(pack_stride / 4);
            P.kind = 0;
            if (packed_in) {
                P.pitch = (uint32_t)dims[q].tpr;
                P.map = src[q];
            } else if (carving && raw_view[q]) {
                P.pitch = (uint32_t)W;
                P.map = src[q];
                P.kind = raw_view_kind(flags);
            } else if (carving) {
                const bool fast = mask_dtype == BP_MASK_U8 && (W % 16) == 0 && ((uintptr_t)src[q] % 16) == 0;
                if (fast) {
                    PackJob &J = jobs.j[jobs.n++];
                    J.src = src[q];
                    J.off = off;
                    J.H = H;
                    J.W = W;
                    J.tpr = dims[q].tpr;
                    const ViewRect &R = rects[q];
                    J.ty0 = R.y0 >> 5;
                    J.ty1 = (R.y1 + 31) >> 5;
                    J.px0 = R.x0;
                    J.px1 = R.x1;
                    J.pad = 0;
                    const int segs = ((R.x1 + 511) >> 9) - (R.x0 >> 9);  // 512-pixel segments = CTAs along x
                    if (segs > jw) jw = segs;
                    if (J.ty1 - J.ty0 > jt) jt = J.ty1 - J.ty0;
                } else {
                    rc = pack_view(h, pool_base, src[q], H, W, mask_dtype, flags, off, dims[q]);
                    if (rc) return rc;
                }
                P.pitch = (uint32_t)dims[q].tpr;
                P.map = pool_base + off;
                off += dims[q].elems;
            } else {
                // averaging reads the caller's row-major map in place (device masks) or its staged copy
                P.pitch = (uint32_t)W;
                P.map = src[q];
            }
            if (pack_only) continue;
            const float *c = cams + 16 * (size_t)v;
            P.fx = c[0]; P.fy = c[1]; P.cx = c[2]; P.cy = c[3];
            for (int a = 0; a < 9; ++a) P.r[a] = c[4 + a];
            for (int a = 0; a < 3; ++a) P.t[a] = c[13 + a];
            P.Wf = (float)W;
            P.Hf = (float)H;
            P.Wi = (uint32_t)W;
            P.Hi = (uint32_t)H;
        }
        if (jobs.n > 0) {
            dim3 grd(jw, jt, jobs.n);
            bpk::pack_u8x16(grd, h->stream, pool_base, jobs, flags);
            CU(cudaGetLastError());
            h->launches++;
        }
        CU(cudaEventRecord(i1, h->stream));
        h->ev_ingest.emplace_back(i0, i1);
        if (host_masks && carving) CU(cudaEventRecord(h->ev_stage_free[s], h->stream));
        if (pack_only) continue;
        CU(cudaEventRecord(k0, h->stream));
        const dim3 grd = h->launch_grid;
        if (carving) {
            // (pageable source: the call returns once the bytes are staged, so `batch` may be reused)
            CU(cudaMemcpyAsync(h->views_d[0], batch.v, sizeof(ViewParams) * (size_t)batch.n, cudaMemcpyHostToDevice, h->stream));
            unsigned long long *work = h->count_work ? h->work_d : nullptr;
            if (h->blk_dead && !batch.force_exact && carve_cull_enabled()) {
                rc = launch_cull_carve(h, batch, hw, order.data() + v0, cams, work);
                if (rc) return rc;
            } else {
                rc = settle_seed(h);  // (no cull pass to seed the state for this launch)
                if (rc) return rc;
                bpk::carve(grd, h->stream, (int8_t *)h->grid, batch, (const ViewParams *)h->views_d[0], h->g, work);
            }
        } else {
            const uint32_t *skip = nullptr;
            float c0 = 0.f;
            if (mask_dtype == BP_MASK_U8 && h->avg_skip_d && !batch.force_exact && avg_skip_enabled()) {
                rc = launch_avg_skip(h, batch, hw, order.data() + v0, cams, flags, &skip, &c0);
                if (rc) return rc;
            }
            h->avg_nominal_vv += (int64_t)h->g.ni * h->g.ny * h->g.nz * (int64_t)batch.n;
            bpk::average(grd, h->stream, mask_dtype == BP_MASK_U8, (float *)h->grid, h->lut_d, batch, h->g, skip, h->avg_nbj,
                         h->avg_nbk, h->avg_lgk, c0);
        }
        CU(cudaGetLastError());
        CU(cudaEventRecord(k1, h->stream));
        h->ev_kernel.emplace_back(k0, k1);
        h->launches++;
        if (host_masks && !carving) CU(cudaEventRecord(h->ev_stage_free[s], h->stream));
        if (ci == ci_snap) {
            rc = start_spec_readback(h);
            if (rc) return rc;
        }
    }
    // host masks are never retained: every H2D copy has completed when this returns
    if (host_masks) CU(cudaStreamSynchronize(h->cstream));
    return BP_OK;
}

}  // namespace bph
