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
    uint32_t pad;
    const void *map;        // carving: this view's bit tiles; averaging: row-major uint8 / float32 map
};
static_assert(sizeof(ViewParams) == 96, "ViewParams layout");

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
