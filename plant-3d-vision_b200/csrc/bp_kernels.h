// bp_kernels.h — launch wrappers of the kernels in bp_kernels.cu (the only translation unit with device code
// for the back-projection path).  Not part of the C ABI.  Every wrapper only enqueues; callers check
// cudaGetLastError() and count the launch.
#ifndef BP_KERNELS_H
#define BP_KERNELS_H
#include <cuda_runtime.h>

#include "bp_types.h"

namespace bpk {

// fused multi-view carve over EVERY patch of the handle's planes (kernels/backprojection.c:57-84); CTA = 64 threads.
// work_total != NULL: the counting variant (adds the projections evaluated; diagnostics only)
void carve(dim3 grd, cudaStream_t st, int8_t *state, const Batch &b, const ViewParams *gviews, const GridInfo &g,
           unsigned long long *work_total);
// block cull of the carve path: summed-area tables of "tile has a foreground bit" for the views of `jobs`
// (two launches), the cull itself (dead flags, zero fill of the blocks it kills, work list of the survivors;
// counters (8 words, zeroed by the caller) = {live blocks listed from the front, next work item, blocks killed,
// candidates after the first pass, live blocks listed from the back};
// three launches: the first n_first views over every block, the rest over the compacted survivors, then the work list
// of the blocks still alive in Morton order of the block coordinates (live_order = 1) or row-major order (0)), and the persistent kernel that
// carves the patches of the surviving blocks (persistent grids: SM count x the occupancy of each kernel).
void tile_sat(cudaStream_t st, const SatJobs &jobs, int max_tiles);
void cull_blocks(cudaStream_t st, int8_t *state, const GridInfo &g, const CullGrid &cg, const CullView *views, int n_views,
                 int n_first, int live_order, int heavy_first, uint8_t *dead, uint2 *cand, uint2 *live, uint32_t *counters,
                 bool fill_unseen);  // fill_unseen: the state is unseeded (lazy clear) - the first pass writes every block
int carve_items_ctas_per_sm();
// carve_items walks every patch of the surviving blocks through the first v_split views; carve_segments (three launches)
// classifies the patches (one byte per work item in cls: 0 dead .. 3 more than 64 voxels live), cuts the remaining views
// of the live ones into segments any warp can take (seg: one uint32 per work item; counters[5..7]), merged by an atomic AND on a transient coding, and
// restores the coding.  n_sm: SM count - the persistent grids are n_sm x the occupancy of each kernel.
void carve_items(int n_sm, cudaStream_t st, int8_t *state, const Batch &b, const ViewParams *gviews, const GridInfo &g,
                 const uint2 *live, uint32_t n_blocks, int grab, int v_split, uint32_t *counters, unsigned long long *work_total);
void carve_segments(int n_sm, cudaStream_t st, int8_t *state, const Batch &b, const ViewParams *gviews, const GridInfo &g,
                    const uint2 *live, uint32_t n_blocks, int v_split, uint8_t *cls, uint32_t *seg, uint32_t *counters,
                    unsigned long long *work_total);
// fused multi-view average (kernels/backprojection.c:36-55); uint8 maps through the 256-entry table, or float32 maps
// skip (u8 maps only, may be NULL): one bit per (skip block, view) from avg_skip - the view adds c0 to every voxel
void average(dim3 grd, cudaStream_t st, bool u8_maps, float *value, const float *lut, const Batch &b, const GridInfo &g,
             const uint32_t *skip, int skip_nbj, int skip_nbk, int skip_lgk, float c0);
// view skipping of the averaging path: per-tile "holds a byte != B" flags of row-major uint8 maps, their summed-area
// tables (tile_sat_only: the second half of tile_sat), and the (block, view) skip bits
void map_any(cudaStream_t st, const MapAnyJobs &jobs, int tiles_y_max);
void tile_sat_only(cudaStream_t st, const SatJobs &jobs);
void avg_skip(cudaStream_t st, const GridInfo &g, int nbj, int nbk, int lgk, uint32_t n_blocks, const CullView *views,
              int n_views, uint32_t *skip, unsigned long long *skipped);
// mask ingest: one view of any dtype / width -> 32x32-pixel bit tiles
void pack_bits(cudaStream_t st, int dtype, const void *src, uint32_t *out, int H, int W, int tpr, int Hpad, uint32_t flags);
// mask ingest, batched fast path (uint8, 16-byte aligned rows): every view of a chunk in one launch
void pack_u8x16(dim3 grd, cudaStream_t st, uint32_t *pool, const PackJobs &jobs, uint32_t flags);
// read-back helpers
void expand_state(cudaStream_t st, const int8_t *state, int32_t *out, size_t n_rows, int nz, int nzp, int v_carved,
                  int v_seen, int v_unseen);
// bit s of *present (zeroed by the caller) is set when carve state s occurs among the unpadded voxels
void state_presence(cudaStream_t st, const int8_t *state, size_t n_rows, int nz, int nzp, uint32_t *present);
void patch_rows(cudaStream_t st, const int8_t *state, const int8_t *snap, int32_t *host_out, size_t n_rows, int nz, int nzp,
                int v_carved, int v_seen, int v_unseen, unsigned long long *count);
void unpad_f32(cudaStream_t st, const float *in, float *out, size_t n, int nz, int nzp);
void tail(cudaStream_t st, int n_blocks, bool carve, bool f64, const void *grid, void *out, size_t n, int nz, int nzp,
          int v_carved, int v_seen, int v_unseen, int exp_clip, double *partial);
void fill_f32(cudaStream_t st, float *p, size_t n, float v);
// device self-test of the shared-reciprocal division (default stream)
void selftest_division(unsigned long long n, unsigned long long seed, unsigned long long *mismatches_d);

}  // namespace bpk
#endif  // BP_KERNELS_H
