/* masks_b200.h — C ABI of the mask producer in libbp_b200.so.
 *
 * Replaces, on one B200, what the reference's `Masks` task does per image before the `Voxels` task reads it back
 * (plant3dvision/tasks/proc2d.py:213-248):
 *   plant3dvision/proc2d.py:69-115    linear:        rescale_intensity(img, out_range=(0, 1)), c0*R + c1*G + c2*B
 *   plant3dvision/proc2d.py:118-169   excess_green:  same rescale, s = R+G+B + 1e-9, 2*G/s - R/s - B/s
 *   tasks/proc2d.py:232               img > threshold
 *   plant3dvision/proc2d.py:172-220   dilation:      binary_dilation(img, disk(n, decomposition='sequence'))
 *   tasks/proc2d.py:237               np.array(255 * img, dtype=uint8)
 * All arithmetic is float64 with separate multiplies, adds and divides in the reference's order, so the binary
 * mask equals the reference's bit for bit.  The masks can stay in HBM (`out_dev`) and feed bp_process_views /
 * bp_pack_views with BP_FLAG_DEVICE_MASKS: the write-PNG -> read-PNG -> upload round trip between the two tasks
 * (tasks/proc2d.py:239-240 -> cl.py:297) disappears.
 *
 * Same conventions as bp_b200.h: 0 on success, negative BP_E* on failure, bp_last_error() for the message, plain
 * pointers and sizes only, no CPU fallback.
 */
#ifndef MASKS_B200_H
#define MASKS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BP_MASKS_LINEAR 0       /* tasks/proc2d.py:217-218 type == "linear"       */
#define BP_MASKS_EXCESS_GREEN 1 /* tasks/proc2d.py:219-220 type == "excess_green" */

/* n images of the same size in one call.
 *   images       n host pointers to uint8 [H, W, C] arrays, C >= 3 (channels 0..2 are R, G, B; like the reference,
 *                the rescale's min / max run over ALL C channels of the image)
 *   type, coefs  see above; coefs (3 doubles) only for BP_MASKS_LINEAR
 *   threshold    tasks/proc2d.py:232 (strict >, float64)
 *   n_footprints, footprints, repeats
 *                the dilation as skimage hands it to scipy: a sequence of 3x3 footprints (row-major 0/1 bytes,
 *                9 per footprint), footprint f applied repeats[f] times; each application is
 *                scipy.ndimage.binary_dilation(., structure=footprint, border_value=0):
 *                out[y][x] = OR over set (a, b) of in[y - (a-1)][x - (b-1)].  n_footprints = 0: no dilation.
 *   masks_out    n host pointers to uint8 [H, W] (0 / 255); may be NULL when out_dev is given
 *   out_dev      if not NULL, receives ONE device buffer holding the n masks back to back (n * H * W bytes); the
 *                caller frees it with bp_device_free (vol2pcd_b200.h)
 *   ms           out (may be NULL): device time of the kernels                                                    */
int bp_masks_filter(int device, int32_t n, const uint8_t *const *images, int32_t H, int32_t W, int32_t C, int type,
                    const double *coefs, double threshold, int32_t n_footprints, const uint8_t *footprints,
                    const int32_t *repeats, uint8_t *const *masks_out, void **out_dev, float *ms);

#ifdef __cplusplus
}
#endif
#endif /* MASKS_B200_H */
