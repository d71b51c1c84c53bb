/* vol2pcd_b200.h — C ABI of the volume -> point-cloud handoff in libbp_b200.so.
 *
 * Replaces, on one B200, what the reference's `PointCloud` task does with the `Voxels` volume:
 *   plant3dvision/proc3d.py:490-570   vol2pcd: binarise (> 0.5), two Euclidean distance transforms,
 *                                     signed distance, np.gradient, three Gaussian filters (sigma 1),
 *                                     level-set band selection, per-point normal + projected point
 *   plant3dvision/tasks/proc3d.py:80-115  multiclass argmax + contrast / score gates
 * The volume can stay in HBM: `vol` may be a device pointer (e.g. bp_device_state of a carving
 * handle), so the D2H -> TIFF/NPZ -> read -> H2D round trip of tasks/cl.py:176-184 and
 * tasks/proc3d.py:69-78 disappears.
 *
 * Same conventions as bp_b200.h: 0 on success, negative BP_E* on failure, bp_last_error() for the
 * message, plain pointers and sizes only, no CPU fallback.
 */
#ifndef VOL2PCD_B200_H
#define VOL2PCD_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct v2p_result v2p_result;

/* element type of the volume handed to bp_vol2pcd */
#define V2P_U8 0  /* uint8 / numpy bool                                   */
#define V2P_I8 1  /* int8 (raw carving state of bp_device_state)          */
#define V2P_I32 2 /* int32 (Backprojection.get_values() of a carving run) */
#define V2P_F32 3 /* float32 (averaging)                                  */
#define V2P_F64 4 /* float64 (process_fileset with labels; NPZ volumes)   */

/* how a voxel is binarised (proc3d.py:515 `volume > 0.5`) */
#define V2P_BIN_GT 0 /* foreground iff (double)v >  bin_value (0.5 for the reference's test)       */
#define V2P_BIN_EQ 1 /* foreground iff (double)v == bin_value (int8 carving state: 1 = seen)       */
#define V2P_BIN_GE 2 /* foreground iff (double)v >= bin_value (log-domain sums: v >= cutoff is the
                        same set as exp(v) > 0.5 when the caller computes the cutoff with the same exp) */

#define V2P_FLAG_DEVICE_VOLUME 1u /* `vol` is a device pointer */

/* Runs the whole of proc3d.py:515-566 on the device and keeps the result in device memory.
 *   vol, vol_dtype   volume [nx, ny, nz], C order; row_pitch = elements between consecutive
 *                    (i, j) rows (0 = nz; the padded pitch of bp_device_state otherwise)
 *   bin_op/bin_value see above
 *   origin, voxel_size   proc3d.py:564 index2point: p = voxel_size * idx + origin (float64)
 *   level_set_value      proc3d.py:536; band_hi = -level_set_value + sqrt(3) and
 *   half_diag = sqrt(3)/2 are passed in so that they are the caller's numpy values
 *   weights, radius      the 2*radius+1 Gaussian weights scipy.ndimage.gaussian_filter(., 1) uses
 *                        (radius 4); the device applies them in scipy's summation order
 * Outputs (float64, C order of the voxel index, points whose smoothed gradient is zero dropped,
 * proc3d.py:545-560): *out holds n points and n unit normals until bp_vol2pcd_free.          */
int bp_vol2pcd(v2p_result **out, int device, const void *vol, int vol_dtype, uint32_t flags,
               const int32_t shape[3], int64_t row_pitch, int bin_op, double bin_value,
               const double origin[3], double voxel_size, double level_set_value, double band_hi,
               double half_diag, const double *weights, int radius);

/* Number of points, and device time (ms, CUDA events) of the whole pipeline / of its distance
 * transform passes alone. */
int bp_vol2pcd_info(const v2p_result *r, int64_t *n_points, float *total_ms, float *edt_ms,
                    int64_t *n_band_voxels, int64_t *launches);

/* Copy the n x 3 float64 points and normals to host buffers (either may be NULL). */
int bp_vol2pcd_fetch(const v2p_result *r, double *points, double *normals);

/* Device pointers of the same arrays (stay valid until bp_vol2pcd_free). */
int bp_vol2pcd_device(const v2p_result *r, void **points, void **normals);

/* The signed, half-voxel-shifted distance of proc3d.py:522 for the whole volume (float64 [nx, ny, nz],
 * host buffer).  Exact where |distance| <= the transform's cap (which covers every voxel the point
 * extraction reads); beyond it the value is +/- a lower bound.  *cap receives that bound.  Debug aid. */
int bp_vol2pcd_signed_distance(const v2p_result *r, double *host_dst, double *cap);

int bp_vol2pcd_free(v2p_result *r);

/* tasks/proc3d.py:80-115 on the device: L label volumes [nx, ny, nz] (float32 or float64, host or
 * device pointers, one per label), `background` = index of the 'background' label (-1 if none).
 * Writes for label `label` a uint8 0/1 device volume (*out_dev, caller frees with bp_device_free)
 * that is exactly the boolean `pred_c` the reference hands to vol2pcd.                          */
int bp_multiclass_select(int device, int n_labels, const void *const *vols, int vol_dtype, uint32_t flags,
                         const int32_t shape[3], int background, double background_prior,
                         double min_contrast, double min_score, int label, void **out_dev);
int bp_device_free(int device, void *ptr);

/* SURVEY §8f-4 — the scoring loop of the SegmentedPointCloud task (plant3dvision/tasks/proc3d.py:203-228) with
 * proc3d.backproject_points (plant3dvision/proc3d.py:655-659), which the reference runs as a pure-Python loop
 * over points inside a loop over files: every point is projected into every label map (float64 pinhole, pixel =
 * (int)(u + 0.5), no depth test — as the reference), and scores[label][point] += map[y][x] when the pixel is
 * inside the picture; then labels[point] = argmax over labels (first maximum).
 *   points   n_points x 3 float64 (host)
 *   cams     n_views x 16 float64: fx fy cx cy | rotmat row-major | tvec
 *   maps     n_views host pointers to uint8 H x W label maps, hw n_views x 2 (H, W), view_label n_views
 *            indices into [0, n_labels)
 *   scores   out, n_labels x n_points float64 (exact integers), may be NULL; labels out, n_points int32
 *   ms       out (may be NULL): device time of the scoring kernel                                        */
int bp_score_points(int device, int64_t n_points, const double *points, int32_t n_views, const double *cams,
                    const void *const *maps, const int32_t *hw, const int32_t *view_label, int32_t n_labels,
                    double *scores, int32_t *labels, float *ms);

#ifdef __cplusplus
}
#endif
#endif /* VOL2PCD_B200_H */
