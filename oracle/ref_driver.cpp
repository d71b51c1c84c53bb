/* TEST INFRASTRUCTURE — not product code.
 *
 * Drives the reference's OWN kernel source, compiled from where it lies under
 * /root/reference (path given as -DREF_KERNEL_FILE=... -I<kernels dir>), through
 * the OpenCL->C++ shim in ref_shim.h.  One call == one `process_view`
 * (plant3dvision/cl.py:190-227): a full pass of global size prod(shape) over the
 * grid for ONE view, work-items spread over OpenMP threads.  Work-items are
 * independent (one voxel each), so the thread order does not change results.
 *
 * Built twice by oracle/Makefile into oracle/_ref/:
 *   libbp_ref_off.so  -O2 -ffp-contract=off            (canonical oracle)
 *   libbp_ref_fma.so  -O2 -ffp-contract=fast -mfma     (what an FMA-contracting
 *                                                       OpenCL compiler would do)
 */
#include <cstdint>
#include "ref_shim.h"
#include REF_KERNEL_FILE

extern "C" {

/* mask: int32 image H x W, exactly what cl.image_from_array(ctx, mask_h, 1) holds for
 * dtype int32 (cl.py:215-217). labels: int32[nx*ny*nz], read-modify-write. */
void ref_carve_view(const int32_t *mask, int H, int W, int32_t *labels,
                    const float *intrinsics, const float *rot, const float *tvec,
                    const float *volinfo, const int32_t *shape) {
    shim_image im = {mask, W, H};
    const long n = (long)shape[0] * shape[1] * shape[2];
#pragma omp parallel for schedule(static)
    for (long g = 0; g < n; ++g) {
        shim_gid = (int)g;
        carve(&im, labels, (float *)intrinsics, (float *)rot, (float *)tvec,
              (float *)volinfo, (int *)shape);
    }
}

/* mask: float32 image H x W (cl.py:205-217 for dtype float32). */
void ref_average_view(const float *mask, int H, int W, float *value,
                      const float *intrinsics, const float *rot, const float *tvec,
                      const float *volinfo, const int32_t *shape) {
    shim_image im = {mask, W, H};
    const long n = (long)shape[0] * shape[1] * shape[2];
#pragma omp parallel for schedule(static)
    for (long g = 0; g < n; ++g) {
        shim_gid = (int)g;
        average(&im, value, (float *)intrinsics, (float *)rot, (float *)tvec,
                (float *)volinfo, (int *)shape);
    }
}

/* Sub-range variant used only to time a bounded sample of a large grid
 * (work-items [g0, g1) of the same global range). */
void ref_carve_view_range(const int32_t *mask, int H, int W, int32_t *labels,
                          const float *intrinsics, const float *rot, const float *tvec,
                          const float *volinfo, const int32_t *shape, long g0, long g1) {
    shim_image im = {mask, W, H};
#pragma omp parallel for schedule(static)
    for (long g = g0; g < g1; ++g) {
        shim_gid = (int)g;
        carve(&im, labels, (float *)intrinsics, (float *)rot, (float *)tvec,
              (float *)volinfo, (int *)shape);
    }
}

void ref_average_view_range(const float *mask, int H, int W, float *value,
                            const float *intrinsics, const float *rot, const float *tvec,
                            const float *volinfo, const int32_t *shape, long g0, long g1) {
    shim_image im = {mask, W, H};
#pragma omp parallel for schedule(static)
    for (long g = g0; g < g1; ++g) {
        shim_gid = (int)g;
        average(&im, value, (float *)intrinsics, (float *)rot, (float *)tvec,
                (float *)volinfo, (int *)shape);
    }
}

const char *ref_build_info(void) {
#ifdef __FP_FAST_FMAF
    return "reference kernels/backprojection.c via shim; fp-contract=fast (FMA)";
#else
    return "reference kernels/backprojection.c via shim; fp-contract=off";
#endif
}
}
