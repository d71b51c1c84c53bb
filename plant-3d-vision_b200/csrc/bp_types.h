// bp_types.h — plain data shared by the kernels (bp_kernels.cu) and the host pipeline (bp_process.cu, bp_b200.cu)
// of libbp_b200.so.  Not part of the C ABI.
#ifndef BP_TYPES_H
#define BP_TYPES_H
#include <cstddef>
#include <cstdint>

constexpr int kMaxViews = 256;  // views per launch (24 KB of the 32 KB parameter space)
constexpr int kMediumMax = 64;   // carve: at most this many live voxels of a warp's 128 leave the dense phase
constexpr int kAvgVpt = 8;       // consecutive-k voxels per thread in average_kernel (two float4)
constexpr size_t kStageBytes = (size_t)256 << 20;  // raw host masks staged per chunk

struct ViewParams {  // 96 bytes
    float r[9];
    float t[3];
    float fx, fy, cx, cy;
    float Wf, Hf;           // image width / height as floats (exact below 2^24)
    uint32_t pitch;         // carving: 32x32-pixel bit tiles per tile row; averaging: elements per row
    uint32_t Wi, Hi;        // the same width / height as integers
    uint32_t kind;          // carving: 0 = bit tiles; else a RAW row-major uint8 mask read in place (kViewRaw...)
    const void *map;        // carving: this view's bit tiles / raw mask; averaging: row-major uint8 / float32 map
};
static_assert(sizeof(ViewParams) == 96, "ViewParams layout");

// ViewParams::kind of a carving view whose uint8 mask is read in place (two-tier masks: only the first views of a launch,
// the ones every voxel meets, are bit-packed; the few voxels that survive them read the later masks raw):
//   foreground = ((byte == ((kind >> 8) & 0xff)) ? 1 : 0) ^ ((kind >> 1) & 1)
// which covers the three readings of nonzero_px<uint8_t>: plain (!= 0), BP_FLAG_INVERT (!= 255), BP_FLAG_BOOL_INVERT (== 0).
constexpr uint32_t kViewRaw = 1u, kViewRawFlip = 2u;
inline uint32_t raw_view_kind(uint32_t bp_flags) {  // bp_flags: BP_FLAG_INVERT = 1, BP_FLAG_BOOL_INVERT = 2
    if (bp_flags & 2u) return kViewRaw | (0u << 8);                  // == 0
    if (bp_flags & 1u) return kViewRaw | kViewRawFlip | (255u << 8); // != 255
    return kViewRaw | kViewRawFlip | (0u << 8);                      // != 0
}

struct Batch {
    int n;
    int force_exact;  // some parameter is non-finite or huge: skip the shared-reciprocal fast path
    int pad[2];
    ViewParams v[kMaxViews];
};

struct GridInfo {
    int ni;     // planes held by this handle
    int ny, nz; // logical row count / row length
    int nzp;    // padded row length (multiple of 16)
    int i_off;  // global index of local plane 0
    // striped sharding (bp_create_striped): local plane li is global plane
    //   i_off + (((li >> il_lg) * il_world + il_rank) << il_lg) + (li & ((1 << il_lg) - 1));
    // a plain handle has il_world = 1, il_rank = 0, i.e. i_off + li.
    int il_lg, il_world, il_rank;
    float ox, oy, oz, vs;
    // CTA scheduling order: the 1-D grid is cut into super-blocks of sb_i planes x sb_j x sb_k CTA
    // tiles (~64^3 voxels) that are launched back to back, so that the CTAs resident at any time
    // cover a compact 3-D block whose footprint in every view stays in L2.
    // Extents are powers of two (lg_* = log2), and the launch is 3-D: blockIdx.x = (super-block along k,
    // CTA inside the super-block), blockIdx.y / .z = super-block along j / i — x runs fastest, so the
    // order is the same as a 1-D walk and the decode needs shifts only (the divisions by run-time
    // extents were 15 % of the carve kernel's instructions, profiles/r01_carve_v6_ncu_full.txt).
    int sb_i, sb_j, sb_k;  // super-block extent in planes / CTA tiles
    int lg_i, lg_j, lg_k;
    int nsb_j, nsb_k;      // super-blocks along j and k
};

// ------------------------------------------------------------------------------------------
// block cull of the carve path (cull_blocks_kernel / carve_items_kernel)
// ------------------------------------------------------------------------------------------
// A cull block is kBlkI planes x kBlkJ rows x kBlkK voxels = 16 warp patches of 4 (j) x 32 (k) voxels in one plane.
constexpr int kBlkI = 8, kBlkJ = 8, kBlkK = 32;
constexpr int kItemsPerBlock = kBlkI * (kBlkJ / 4) * (kBlkK / 32);
constexpr int kCullViews = 32;     // views (in visiting order) the block cull may use
constexpr int kCullFirstViews = 3; // ... of which the first pass (every block) takes these; the rest see survivors only

struct CullView {  // 96 bytes: what the block cull needs of one view
    float cam[16];         // fx fy cx cy | rot[9] | tvec[3]
    const uint32_t *sat;   // (tiles_y + 1) x (tpr + 1) summed-area table of "tile has a foreground bit"
    int32_t tpr1;          // tpr + 1 (row pitch of sat)
    int32_t W, H;
    int32_t pad[3];
};
static_assert(sizeof(CullView) == 96, "CullView layout");

struct SatJob {  // 32 bytes: build the summed-area table of one view's bit tiles
    const uint32_t *tiles;
    uint32_t *sat;
    uint8_t *any;          // tiles_y x tpr scratch
    int32_t tpr, tiles_y;
};
struct SatJobs {
    int n;
    int pad[3];
    SatJob j[kMaxViews];
};

// ------------------------------------------------------------------------------------------
// view skipping of the averaging path (avg_skip_kernel / average_kernel)
// ------------------------------------------------------------------------------------------
// A skip block is kAvgBlkI planes x 32 rows x 2^lgk voxels (lgk = 3..6, bp_handle::avg_lgk): the voxels of 1, 2, 4 or
// 8 warps of an average_kernel CTA tile, in 8 consecutive planes.  One bit per (block, view of the launch): every voxel
// of the block lands inside the picture on a pixel holding the launch's "uniform byte", so the view adds the same
// table entry to all of them.
constexpr int kAvgBlkI = 8, kAvgBlkJ = 32;
constexpr int kAvgTileK = 8 * kAvgVpt;          // voxels along k of an average_kernel CTA tile
constexpr int kSkipWords = kMaxViews / 32;

struct MapAnyJob {  // 24 bytes: "does this 32 x 32-pixel tile of a row-major uint8 map hold a byte != B"
    const uint8_t *src;
    uint8_t *any;          // tiles_y x tpr
    int32_t H, W;
};
struct MapAnyJobs {
    int n;
    int tpr_max;
    uint32_t b4;           // the uniform byte B replicated four times
    int pad;
    MapAnyJob j[kMaxViews];
};

struct CullGrid {
    int nbi, nbj, nbk;       // cull blocks along i (local planes), j, k
    uint32_t n_blocks;
};

// Batched fast path of the mask ingest: one PackJob per view of a chunk (see pack_bits_u8x16_kernel).
struct PackJob {  // 48 bytes
    const void *src;
    unsigned long long off;  // word offset of the view's tiled map in the pool
    int H, W, tpr;
    // the part of the view that is packed: tile rows [ty0, ty1) and pixel columns [px0, px1) (px0 a multiple of
    // 32, px1 a multiple of 32 or W).  The whole view unless BP_VIEW_RECT is on (see view_rect): tiles outside
    // the rectangle the grid projects into are never sampled, so they are neither read nor written.
    int ty0, ty1, px0, px1;
    int pad;
};
struct PackJobs {
    int n;
    int pad[3];
    PackJob j[kMaxViews];
};

#endif  // BP_TYPES_H
