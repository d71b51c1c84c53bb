/* TEST INFRASTRUCTURE — not product code.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library.
 *
 * CPU restatement (plain C, OpenMP over grid (plane, row) pairs) of the reference's voxel
 * back-projection, one call per view exactly as Backprojection.process_view does
 * (plant3dvision/cl.py:190-227 launches one work-item per voxel per view).
 *
 * PARITY PINNING: the reference's tests hold no golden vector for this path
 * (tests/unit/test_cl.py:4-9 only constructs the object).  This restatement is
 * pinned instead against the reference's own kernel source compiled unmodified
 * (oracle/ref_driver.cpp -> oracle/_ref/libbp_ref_off.so) on the reference's
 * virtual_plant fixture: tests/test_oracle.py checks bit-equality, and the
 * outputs of that run are committed under tests/golden/ (tests/golden/make_golden.py).
 *
 * Build: -O2 -ffp-contract=off  (every product and sum rounded separately, the
 * canonical reading of the OpenCL C source).
 *
 * Arithmetic followed, with the reference line each step restates:
 *   idx -> (i,j,k), k fastest                       kernels/common.h:1-12
 *   world = origin + ijk * voxel_size               kernels/backprojection.c:46-48, 71-73
 *   cam z = r6*x + r7*y + r8*z + t2, reject z < 0   kernels/backprojection.c:11-15
 *   cam x, y likewise                               kernels/backprojection.c:17-18
 *   u = x/z*fx + cx ; v = y/z*fy + cy               kernels/backprojection.c:20-21
 *   (int) truncation, accept 0..W-1 / 0..H-1        kernels/backprojection.c:23-31
 *   carve state machine                             kernels/backprojection.c:65-83
 *   average accumulate                              kernels/backprojection.c:54
 */
#include <stdint.h>
#include <limits.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct {
    float fx, fy, cx, cy;
    float r[9];
    float t[3];
} bp_cam;

/* x86 cvttss2si semantics for (int)float, made explicit so the restatement does not
 * depend on the host ISA: NaN / out-of-range -> INT_MIN ("integer indefinite"). */
static inline int trunc_to_int(float f) {
    if (!(f > -2147483904.0f && f < 2147483648.0f)) return INT_MIN;
    return (int)f;
}

/* returns 1 and the pixel if the voxel's world point lands inside the W x H image */
static inline int project(const bp_cam *c, float x, float y, float z, int W, int H,
                          int *px, int *py) {
    float pz = c->r[6] * x + c->r[7] * y + c->r[8] * z + c->t[2];
    if (pz < 0) return 0;
    float pxf = c->r[0] * x + c->r[1] * y + c->r[2] * z + c->t[0];
    float pyf = c->r[3] * x + c->r[4] * y + c->r[5] * z + c->t[1];
    pxf = pxf / pz * c->fx + c->cx;
    pyf = pyf / pz * c->fy + c->cy;
    int ix = trunc_to_int(pxf);
    int iy = trunc_to_int(pyf);
    if (ix < 0 || ix > W - 1) return 0;
    if (iy < 0 || iy > H - 1) return 0;
    *px = ix;
    *py = iy;
    return 1;
}

static inline bp_cam load_cam(const float *intr, const float *rot, const float *tvec) {
    bp_cam c;
    c.fx = intr[0]; c.fy = intr[1]; c.cx = intr[2]; c.cy = intr[3];
    for (int a = 0; a < 9; ++a) c.r[a] = rot[a];
    for (int a = 0; a < 3; ++a) c.t[a] = tvec[a];
    return c;
}

/* Carving, one view.  mask: uint8 H x W, only (mask != 0) matters because the
 * reference widens to int32 (cl.py:215) and the kernel tests `== 0`.
 * i_offset lets a caller run a slab [i_offset, i_offset + shape[0]) of a taller grid
 * with the GLOBAL plane index in the world-point formula (multi-GPU slab parity). */
void oracle_carve_view(const uint8_t *mask, int H, int W, int32_t *state,
                       const float *intr, const float *rot, const float *tvec,
                       const float *origin, float voxel_size, const int32_t *shape,
                       int i_offset) {
    const bp_cam c = load_cam(intr, rot, tvec);
    const int nx = shape[0], ny = shape[1], nz = shape[2];
#pragma omp parallel for collapse(2) schedule(dynamic, 8)
    for (int i = 0; i < nx; ++i) {
        for (int j = 0; j < ny; ++j) {
            const float x = origin[0] + (i + i_offset) * voxel_size;
            const float y = origin[1] + j * voxel_size;
            int32_t *row = state + ((int64_t)i * ny + j) * nz;
            for (int k = 0; k < nz; ++k) {
                if (row[k] == -1) continue;
                const float z = origin[2] + k * voxel_size;
                int px, py;
                if (!project(&c, x, y, z, W, H, &px, &py)) continue;
                if (mask[(int64_t)py * W + px] == 0) row[k] = -1;
                else if (row[k] == 0) row[k] = 1;
            }
        }
    }
}

/* Averaging, one view.  mask: float32 H x W, already converted the way
 * process_view does (img_as_float32, optional log: cl.py:205-208). */
void oracle_average_view(const float *mask, int H, int W, float *value,
                         const float *intr, const float *rot, const float *tvec,
                         const float *origin, float voxel_size, const int32_t *shape,
                         int i_offset) {
    const bp_cam c = load_cam(intr, rot, tvec);
    const int nx = shape[0], ny = shape[1], nz = shape[2];
#pragma omp parallel for collapse(2) schedule(dynamic, 8)
    for (int i = 0; i < nx; ++i) {
        for (int j = 0; j < ny; ++j) {
            const float x = origin[0] + (i + i_offset) * voxel_size;
            const float y = origin[1] + j * voxel_size;
            float *row = value + ((int64_t)i * ny + j) * nz;
            for (int k = 0; k < nz; ++k) {
                const float z = origin[2] + k * voxel_size;
                int px, py;
                if (!project(&c, x, y, z, W, H, &px, &py)) continue;
                row[k] += mask[(int64_t)py * W + px];
            }
        }
    }
}

/* Same as oracle_average_view but the map stays uint8 and is looked up through a
 * 256-entry float table (the table is what img_as_float32 [+ log] gives for each
 * byte value, computed by the caller in numpy) - avoids a 4 B/pixel host copy on
 * large synthetic workloads; bit-identical by construction. */
void oracle_average_view_lut(const uint8_t *mask, const float *lut, int H, int W, float *value,
                             const float *intr, const float *rot, const float *tvec,
                             const float *origin, float voxel_size, const int32_t *shape,
                             int i_offset) {
    const bp_cam c = load_cam(intr, rot, tvec);
    const int nx = shape[0], ny = shape[1], nz = shape[2];
#pragma omp parallel for collapse(2) schedule(dynamic, 8)
    for (int i = 0; i < nx; ++i) {
        for (int j = 0; j < ny; ++j) {
            const float x = origin[0] + (i + i_offset) * voxel_size;
            const float y = origin[1] + j * voxel_size;
            float *row = value + ((int64_t)i * ny + j) * nz;
            for (int k = 0; k < nz; ++k) {
                const float z = origin[2] + k * voxel_size;
                int px, py;
                if (!project(&c, x, y, z, W, H, &px, &py)) continue;
                row[k] += lut[mask[(int64_t)py * W + px]];
            }
        }
    }
}

/* Projection alone (kernels/backprojection.c:3-34) of every voxel of the grid into one view: hit[idx] = 1 and
 * (px[idx], py[idx]) = the pixel when the voxel lands inside the W x H image, else hit[idx] = 0.  Used by the tests of
 * the conservative pixel bounds behind the view rectangles and the block cull (csrc/bp_geom.h). */
void oracle_project_view(int H, int W, int32_t *px, int32_t *py, uint8_t *hit,
                         const float *intr, const float *rot, const float *tvec,
                         const float *origin, float voxel_size, const int32_t *shape, int i_offset) {
    const bp_cam c = load_cam(intr, rot, tvec);
    const int nx = shape[0], ny = shape[1], nz = shape[2];
#pragma omp parallel for collapse(2) schedule(static)
    for (int i = 0; i < nx; ++i) {
        for (int j = 0; j < ny; ++j) {
            const float x = origin[0] + (i + i_offset) * voxel_size;
            const float y = origin[1] + j * voxel_size;
            const int64_t base = ((int64_t)i * ny + j) * nz;
            for (int k = 0; k < nz; ++k) {
                const float z = origin[2] + k * voxel_size;
                int ix = 0, iy = 0;
                const int ok = project(&c, x, y, z, W, H, &ix, &iy);
                hit[base + k] = (uint8_t)ok;
                px[base + k] = ok ? ix : -1;
                py[base + k] = ok ? iy : -1;
            }
        }
    }
}

/* Thread control for the timed CPU baseline (bench.py reports the threads that were busy and the measured
 * 1-thread -> all-thread speed-up).  The view functions parallelise over (plane, row) pairs, so a one-plane
 * slab still spreads over every thread. */
int oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void oracle_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

const char *oracle_build_info(void) {
    return "bp_oracle.c restatement; -ffp-contract=off; x86 cvttss2si truncation made explicit";
}
