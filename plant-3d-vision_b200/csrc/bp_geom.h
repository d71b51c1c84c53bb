// bp_geom.h — conservative pixel bounds of a box of voxels under the projection of kernels/backprojection.c:3-34.
// Shared by the host (view rectangles of the mask ingest, bp_view_rect / bp_box_rect) and the device (block cull of the
// carve path, bp_kernels.cu): the SAME code, so the CPU tests of its conservativeness (tests/test_view_rect.py) cover
// what the GPU evaluates.
#ifndef BP_GEOM_H
#define BP_GEOM_H
#include <cmath>

#ifdef __CUDACC__
#define BP_HD __host__ __device__
#else
#define BP_HD
#endif

struct ViewRect {
    int x0, y0, x1, y1;  // [x0, x1) x [y0, y1)
};

BP_HD inline bool bp_finite(double v) { return (v - v) == 0.0; }

// Pixel bounds of the voxels [i_lo, i_hi] x [j_lo, j_hi] x [k_lo, k_hi] (inclusive index ranges; i is the GLOBAL plane
// index) of a grid with origin (ox, oy, oz) and voxel size vs, seen by camera `cam` (fx fy cx cy | rot[9] | tvec[3]) in a
// W x H image.  On success lo[a] / hi[a] (a = 0: x, 1: y) are such that EVERY voxel of the box, as the kernels evaluate it
// in float32, has p_z > 0 and a pixel coordinate u with  lo <= floor(u) <= hi  — without clipping to the image.
// Conservative by construction: p_z is affine in the voxel index, so when it is positive at the 8 corners of the box it
// is positive inside, and the pinhole map then sends the box into the convex hull of its projected corners.  The kernel
// evaluates the same map in float32: p_z must be at least 1 % of the magnitude of its terms, and the margin added is
// twice a bound of the float32 evaluation error plus 2 pixels.  Any doubt (a corner at or behind the camera plane,
// non-finite numbers) returns false.
BP_HD inline bool bp_box_bounds(int i_lo, int i_hi, int j_lo, int j_hi, int k_lo, int k_hi, double ox, double oy, double oz,
                                double vs, const float *cam, int H, int W, double lo_out[2], double hi_out[2]) {
    const double fx = cam[0], fy = cam[1], cx = cam[2], cy = cam[3];
    const float *r = cam + 4, *t = cam + 13;
    double lo[2] = {INFINITY, INFINITY}, hi[2] = {-INFINITY, -INFINITY};
    double mag_max[3] = {0.0, 0.0, 0.0}, pz_min = INFINITY;  // magnitude of the terms of p_x, p_y, p_z
    for (int c = 0; c < 8; ++c) {
        const double o[3] = {ox, oy, oz};
        const double d[3] = {(double)((c & 4) ? i_hi : i_lo) * vs, (double)((c & 2) ? j_hi : j_lo) * vs,
                             (double)((c & 1) ? k_hi : k_lo) * vs};
        double w[3], wm[3];  // world point, and the magnitude its float32 rounding scales with
        for (int a = 0; a < 3; ++a) {
            w[a] = o[a] + d[a];
            wm[a] = fabs(o[a]) + fabs(d[a]);
        }
        double p[3], mag[3];
        for (int a = 0; a < 3; ++a) {
            p[a] = (double)r[3 * a] * w[0] + (double)r[3 * a + 1] * w[1] + (double)r[3 * a + 2] * w[2] + (double)t[a];
            mag[a] = fabs((double)r[3 * a]) * wm[0] + fabs((double)r[3 * a + 1]) * wm[1] +
                     fabs((double)r[3 * a + 2]) * wm[2] + fabs((double)t[a]);
            if (!bp_finite(mag[a])) return false;
            if (mag[a] > mag_max[a]) mag_max[a] = mag[a];
        }
        if (!(p[2] > 1e-2 * mag[2])) return false;
        if (p[2] < pz_min) pz_min = p[2];
        const double uv[2] = {p[0] / p[2] * fx + cx, p[1] / p[2] * fy + cy};
        for (int a = 0; a < 2; ++a) {
            if (!bp_finite(uv[a])) return false;
            if (uv[a] < lo[a]) lo[a] = uv[a];
            if (uv[a] > hi[a]) hi[a] = uv[a];
        }
    }
    // float32 evaluation error of u (v likewise) for a voxel that lands inside the image, |u - cx| <= W + |cx| =: U:
    // each of p_x, p_z carries <= 8 eps of the magnitude of its terms (world point, products, sums), so
    //   |du| <= 8 eps (|fx| mag_x + U mag_z) / p_z + 4 eps (U + |cx|);   twice that, plus 2 pixels, is the margin.
    const double eps = 1.1920929e-7;
    const double f[2] = {fabs(fx), fabs(fy)}, cc[2] = {fabs(cx), fabs(cy)};
    const double extent[2] = {(double)W, (double)H};
    for (int a = 0; a < 2; ++a) {
        const double U = extent[a] + cc[a];
        const double margin = 2.0 + 16.0 * eps * (f[a] * mag_max[a] + U * mag_max[2]) / pz_min + 8.0 * eps * (U + cc[a]);
        if (!bp_finite(margin)) return false;
        lo_out[a] = floor(lo[a] - margin);
        hi_out[a] = floor(hi[a] + margin);
    }
    return true;
}

// The pixels of one view that the voxels of planes [i_lo, i_hi] x [j_lo, j_hi] x [k_lo, k_hi] can project to, as a
// rectangle [x0, x1) x [y0, y1) of whole 32 x 32 tiles clipped to the image.  Any doubt returns the whole image; nothing
// to see (the box projects outside the image) returns the first tile, so that callers never meet an empty rectangle.
inline ViewRect bp_box_rect(int i_lo, int i_hi, int j_lo, int j_hi, int k_lo, int k_hi, double ox, double oy, double oz,
                            double vs, const float *cam, int H, int W) {
    const ViewRect full = {0, 0, W, H};
    if (i_hi < i_lo || j_hi < j_lo || k_hi < k_lo) return full;
    double lo[2], hi[2];
    if (!bp_box_bounds(i_lo, i_hi, j_lo, j_hi, k_lo, k_hi, ox, oy, oz, vs, cam, H, W, lo, hi)) return full;
    const int dim[2] = {W, H};
    int a0[2], a1[2];
    for (int a = 0; a < 2; ++a) {
        const double l = lo[a], u = hi[a];
        if (u < 0.0 || l > (double)(dim[a] - 1)) {  // nothing of the box is inside the image
            const ViewRect first = {0, 0, W < 32 ? W : 32, H < 32 ? H : 32};
            return first;
        }
        const int li = l < 0.0 ? 0 : (int)l;
        const int ui = u > (double)(dim[a] - 1) ? dim[a] - 1 : (int)u;
        a0[a] = (li >> 5) << 5;
        a1[a] = ((ui >> 5) + 1) << 5;
        if (a1[a] > dim[a]) a1[a] = dim[a];
    }
    const ViewRect R = {a0[0], a0[1], a1[0], a1[1]};
    return R;
}

#endif  // BP_GEOM_H
