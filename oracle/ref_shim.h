/* TEST INFRASTRUCTURE — not product code.
 *
 * Minimal OpenCL-C -> C++ shim so that the reference's own kernel source
 * (/root/reference/plant3dvision/kernels/backprojection.c + common.h) compiles
 * UNMODIFIED, from where it lies, into oracle/_ref/libbp_ref_*.so.
 * Nothing from the reference is copied: ref_driver.cpp #includes the file by
 * path (REF_KERNEL_FILE, given on the compiler command line).
 *
 * Semantics chosen for what OpenCL leaves implementation-defined:
 *   - read_imagei / read_imagef with integer coordinates: exact texel fetch
 *     mask[y][x] (the LINEAR sampler with int2 coords is spec-undefined;
 *     every implementation we know of does a texel fetch).
 *   - (int)float: whatever g++ emits on x86-64, i.e. cvttss2si
 *     (INT_MIN for NaN / out of range).
 */
#ifndef BP_REF_SHIM_H
#define BP_REF_SHIM_H

#define __global
#define __kernel static inline
#define __read_only
#define __constant const

typedef unsigned int sampler_t;
#define CLK_NORMALIZED_COORDS_FALSE 0u
#define CLK_ADDRESS_NONE 0u
#define CLK_FILTER_LINEAR 1u
#define CLK_FILTER_NEAREST 2u

struct float3 { float x, y, z; };
struct float4 { float x, y, z, w; };
struct int2 { int x, y; };
struct int3 { int x, y, z; };
struct int4 { int x, y, z, w; };

struct shim_image { const void *data; int width; int height; };
typedef const shim_image *image2d_t;

static thread_local int shim_gid = 0;
static inline int get_global_id(int) { return shim_gid; }
static inline int get_image_width(image2d_t im) { return im->width; }
static inline int get_image_height(image2d_t im) { return im->height; }

static inline int4 read_imagei(image2d_t im, sampler_t, int2 p) {
    int v = ((const int *)im->data)[(long)p.y * im->width + p.x];
    int4 r = {v, 0, 0, 1};
    return r;
}
static inline float4 read_imagef(image2d_t im, sampler_t, int2 p) {
    float v = ((const float *)im->data)[(long)p.y * im->width + p.x];
    float4 r = {v, 0.f, 0.f, 1.f};
    return r;
}
#endif
