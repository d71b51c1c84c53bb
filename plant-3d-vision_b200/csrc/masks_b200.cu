// masks_b200.cu — the mask producer on one B200, behind bp_masks_filter of include/masks_b200.h.
// Replaces per image what the reference's Masks task does on the host (plant3dvision/tasks/proc2d.py:224-248 with
// proc2d.linear / excess_green / dilation, plant3dvision/proc2d.py:69-220): rescale to [0, 1] by the image's own
// min / max, a per-pixel float64 index, a threshold, and a sequence of 3x3 binary dilations.
//
// Layout: a batch of n equal-sized uint8 images in one pool; the binary image lives as one bit per pixel
// (32 pixels of a row per word, bit i = pixel 32*w + i, so a warp's ballot IS the word), two ping-pong bit planes
// for the dilation sequence, and a final expansion to the 0 / 255 bytes the reference writes.  HBM traffic:
// C bytes read per pixel once; everything after the threshold moves 1/8 byte per pixel per step.
#include "masks_b200.h"

#include "bp_b200.h"
#include "bp_internal.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <vector>

namespace {

#define CUM(call)                                                                              \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess)                                                                 \
            return bp_fail(BP_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_),   \
                           __FILE__, __LINE__);                                                \
    } while (0)

// min / max over ALL bytes of each image (skimage.exposure.rescale_intensity, in_range='image').
__global__ void __launch_bounds__(256)
minmax_kernel(const uint8_t *__restrict__ pool, size_t img_stride, size_t img_bytes, uint32_t *__restrict__ mm) {
    const uint8_t *img = pool + (size_t)blockIdx.y * img_stride;  // img_stride is a multiple of 256
    const size_t n16 = img_bytes / 16;
    uint32_t lo = 0xffffffffu, hi = 0u;  // four packed byte lanes
    for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < n16; q += (size_t)gridDim.x * blockDim.x) {
        const uint4 v = __ldg(reinterpret_cast<const uint4 *>(img) + q);
        lo = __vminu4(lo, __vminu4(__vminu4(v.x, v.y), __vminu4(v.z, v.w)));
        hi = __vmaxu4(hi, __vmaxu4(__vmaxu4(v.x, v.y), __vmaxu4(v.z, v.w)));
    }
    uint32_t l = min(min(lo & 255u, (lo >> 8) & 255u), min((lo >> 16) & 255u, lo >> 24));
    uint32_t h = max(max(hi & 255u, (hi >> 8) & 255u), max((hi >> 16) & 255u, hi >> 24));
    if (blockIdx.x == 0 && threadIdx.x < (img_bytes & 15)) {  // tail bytes
        const uint32_t b = img[n16 * 16 + threadIdx.x];
        l = min(l, b);
        h = max(h, b);
    }
    l = __reduce_min_sync(0xffffffffu, l);
    h = __reduce_max_sync(0xffffffffu, h);
    if ((threadIdx.x & 31) == 0) {
        atomicMin(&mm[2 * blockIdx.y], l);
        atomicMax(&mm[2 * blockIdx.y + 1], h);
    }
}

// One warp = one 32-pixel word of one row at a time, WORDS_PER_WARP words in a row.  float64 throughout, every
// operation rounded separately and in the reference's order (numpy evaluates one ufunc per operator; no
// contraction).  A pixel value is a uint8, so the rescale (a float64 subtract and divide) and, for `linear`, the
// product with the channel's coefficient are tabulated once per block in shared memory: 256 entries per channel,
// the same IEEE operations on the same operands, evaluated once instead of per pixel.
constexpr int WORDS_PER_WARP = 16;

template <int TYPE>
__global__ void __launch_bounds__(256)
threshold_kernel(const uint8_t *__restrict__ pool, size_t img_stride, int H, int W, int C, int pw,
                 const uint32_t *__restrict__ mm, double c0, double c1, double c2, double threshold,
                 uint32_t *__restrict__ bits) {
    constexpr int NL = TYPE == BP_MASKS_LINEAR ? 3 : 1, LG = NL > 1 ? 1 : 0, LB = NL > 1 ? 2 : 0;  // table rows
    __shared__ double lut[NL][256];
    {
        const double imin = (double)mm[2 * blockIdx.y], imax = (double)mm[2 * blockIdx.y + 1];
        const double v = (double)threadIdx.x;
        // rescale_intensity: (image - imin) / (imax - imin) [* 1.0 + 0.0]; flat image: clip to [0, 1].  Values
        // outside [imin, imax] do not occur in the image; their entries are never read.
        const double r = (imin != imax) ? __ddiv_rn(__dsub_rn(v, imin), __dsub_rn(imax, imin)) : fmin(fmax(v, 0.0), 1.0);
        if (TYPE == BP_MASKS_LINEAR) {
            lut[0][threadIdx.x] = __dmul_rn(c0, r);
            lut[LG][threadIdx.x] = __dmul_rn(c1, r);
            lut[LB][threadIdx.x] = __dmul_rn(c2, r);
        } else {
            lut[0][threadIdx.x] = r;
        }
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const uint8_t *img = pool + (size_t)blockIdx.y * img_stride;
    const long long total = (long long)H * pw;
    long long word = ((long long)blockIdx.x * 8 + (threadIdx.x >> 5)) * WORDS_PER_WARP;  // within the image
    for (int it = 0; it < WORDS_PER_WARP && word < total; ++it, ++word) {
        const int y = (int)(word / pw), wx = (int)(word - (long long)y * pw);
        const int x = wx * 32 + lane;
        bool on = false;
        if (x < W) {
            const uint8_t *p = img + ((size_t)y * W + x) * C;
            double val;
            if (TYPE == BP_MASKS_LINEAR) {
                // coefs[0] * R + coefs[1] * G + coefs[2] * B   (proc2d.py:115)
                val = __dadd_rn(__dadd_rn(lut[0][p[0]], lut[LG][p[1]]),
                                lut[LB][p[2]]);
            } else {
                // s = img.sum(axis=2) + EPS over ALL channels, left to right; r, g, b = ./s; 2*g - r - b
                // (proc2d.py:165-169)
                const double cr = lut[0][p[0]], cg = lut[0][p[1]], cb = lut[0][p[2]];
                double s = __dadd_rn(__dadd_rn(cr, cg), cb);
                if (C == 4) s = __dadd_rn(s, lut[0][p[3]]);
                s = __dadd_rn(s, 1e-9);
                const double r = __ddiv_rn(cr, s), g = __ddiv_rn(cg, s), b = __ddiv_rn(cb, s);
                val = __dsub_rn(__dsub_rn(__dmul_rn(2.0, g), r), b);
            }
            on = val > threshold;
        }
        const uint32_t w = __ballot_sync(0xffffffffu, on);
        if (lane == 0) bits[(size_t)blockIdx.y * total + word] = w;
    }
}

// One application of scipy.ndimage.binary_dilation(., structure=3x3 footprint, border_value=0) on bit planes:
// out[y][x] = OR over set (a, b) of in[y - (a-1)][x - (b-1)].  fp: bit 3*a + b.
__global__ void __launch_bounds__(256)
dilate_kernel(const uint32_t *__restrict__ in, uint32_t *__restrict__ out, int H, int W, int pw, uint32_t fp) {
    const long long word = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (word >= (long long)H * pw) return;
    const int y = (int)(word / pw), wx = (int)(word - (long long)y * pw);
    const uint32_t *img = in + (size_t)blockIdx.y * H * pw;
    uint32_t acc = 0;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const uint32_t f = (fp >> (3 * a)) & 7u;
        const int ys = y - (a - 1);
        if (!f || ys < 0 || ys >= H) continue;
        const uint32_t *row = img + (size_t)ys * pw;
        const uint32_t cur = row[wx];
        const uint32_t left = wx > 0 ? row[wx - 1] : 0u, right = wx + 1 < pw ? row[wx + 1] : 0u;
        if (f & 1u) acc |= (cur >> 1) | (right << 31);  // b = 0: source x + 1
        if (f & 2u) acc |= cur;                         // b = 1
        if (f & 4u) acc |= (cur << 1) | (left >> 31);   // b = 2: source x - 1
    }
    if (wx == pw - 1 && (W & 31)) acc &= (1u << (W & 31)) - 1u;  // outside the picture stays 0 (border_value=0)
    out[(size_t)blockIdx.y * H * pw + word] = acc;
}

// np.array(255 * img, dtype=np.uint8)
__global__ void __launch_bounds__(256)
expand_bits_kernel(const uint32_t *__restrict__ bits, uint8_t *__restrict__ out, int H, int W, int pw) {
    const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= (long long)H * W) return;
    const int y = (int)(p / W), x = (int)(p - (long long)y * W);
    const uint32_t w = bits[((size_t)blockIdx.y * H + y) * pw + (x >> 5)];
    out[(size_t)blockIdx.y * H * W + p] = ((w >> (x & 31)) & 1u) ? 255 : 0;
}

struct Pool {
    std::vector<void *> p;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    ~Pool() {
        for (void *q : p)
            if (q) cudaFree(q);
        if (e0) cudaEventDestroy(e0);
        if (e1) cudaEventDestroy(e1);
    }
    template <typename T>
    cudaError_t alloc(T **out, size_t bytes) {
        void *q = nullptr;
        cudaError_t e = cudaMalloc(&q, bytes ? bytes : 1);
        if (e == cudaSuccess) p.push_back(q);
        *out = static_cast<T *>(q);
        return e;
    }
    void release(void *q) {
        for (void *&r : p)
            if (r == q) r = nullptr;
    }
};

}  // namespace

extern "C" int bp_masks_filter(int device, int32_t n, const uint8_t *const *images, int32_t H, int32_t W, int32_t C,
                               int type, const double *coefs, double threshold, int32_t n_footprints,
                               const uint8_t *footprints, const int32_t *repeats, uint8_t *const *masks_out,
                               void **out_dev, float *ms) {
    if (ms) *ms = 0.f;
    if (out_dev) *out_dev = nullptr;
    if (n < 0 || H < 0 || W < 0 || n_footprints < 0) return bp_fail(BP_EINVAL, "negative size");
    if (C < 3 || C > 4) return bp_fail(BP_EINVAL, "images must have 3 or 4 channels, got %d", C);
    if (type != BP_MASKS_LINEAR && type != BP_MASKS_EXCESS_GREEN)
        return bp_fail(BP_EINVAL, "unknown masking type %d", type);  // tasks/proc2d.py:221-222
    if (type == BP_MASKS_LINEAR && !coefs) return bp_fail(BP_EINVAL, "linear masks need 3 coefficients");
    if (n_footprints > 0 && (!footprints || !repeats)) return bp_fail(BP_EINVAL, "null footprints");
    if (n == 0 || H == 0 || W == 0) return BP_OK;
    if (!images || (!masks_out && !out_dev)) return bp_fail(BP_EINVAL, "null argument");
    for (int i = 0; i < n; ++i)
        if (!images[i] || (masks_out && !masks_out[i])) return bp_fail(BP_EINVAL, "image %d is null", i);
    std::vector<uint32_t> fps;
    for (int f = 0; f < n_footprints; ++f) {
        uint32_t m = 0;
        for (int e = 0; e < 9; ++e)
            if (footprints[9 * f + e]) m |= 1u << e;
        if (repeats[f] < 0) return bp_fail(BP_EINVAL, "footprint %d repeated %d times", f, repeats[f]);
        for (int r = 0; r < repeats[f]; ++r) fps.push_back(m);
    }
    int ndev = 0;
    CUM(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return bp_fail(BP_ECUDA, "CUDA device %d not available (%d visible)", device, ndev);
    CUM(cudaSetDevice(device));
    const size_t img_bytes = (size_t)H * W * C, img_stride = (img_bytes + 255) & ~(size_t)255;
    const int pw = (W + 31) / 32;
    const size_t plane_words = (size_t)n * H * pw, mask_bytes = (size_t)n * H * W;
    if ((size_t)H * pw > 0x7fffffffull * 8 || (size_t)H * W > 0x7fffffffull * 256 || n > 65535)
        return bp_fail(BP_EINVAL, "batch too large for one call");
    Pool pl;
    uint8_t *pool = nullptr, *mask = nullptr;
    uint32_t *mm = nullptr, *bits[2] = {nullptr, nullptr};
    if (pl.alloc(&pool, img_stride * n) != cudaSuccess || pl.alloc(&mm, (size_t)n * 8) != cudaSuccess ||
        pl.alloc(&bits[0], plane_words * 4) != cudaSuccess ||
        (!fps.empty() && pl.alloc(&bits[1], plane_words * 4) != cudaSuccess) ||
        pl.alloc(&mask, mask_bytes) != cudaSuccess)
        return bp_fail(BP_ENOMEM, "cudaMalloc for %d images of %d x %d x %d failed", n, H, W, C);
    for (int i = 0; i < n; ++i)
        CUM(cudaMemcpyAsync(pool + img_stride * i, images[i], img_bytes, cudaMemcpyHostToDevice, 0));
    std::vector<uint32_t> mm0((size_t)n * 2);
    for (int i = 0; i < n; ++i) {
        mm0[2 * i] = 255u;
        mm0[2 * i + 1] = 0u;
    }
    CUM(cudaMemcpyAsync(mm, mm0.data(), mm0.size() * 4, cudaMemcpyHostToDevice, 0));
    CUM(cudaEventCreate(&pl.e0));
    CUM(cudaEventCreate(&pl.e1));
    CUM(cudaEventRecord(pl.e0, 0));
    {
        const unsigned bx = (unsigned)std::min<size_t>((img_bytes / 16 + 255) / 256 + 1, 148 * 4);
        minmax_kernel<<<dim3(bx, n), 256>>>(pool, img_stride, img_bytes, mm);
        CUM(cudaGetLastError());
    }
    const double c0 = coefs ? coefs[0] : 0.0, c1 = coefs ? coefs[1] : 0.0, c2 = coefs ? coefs[2] : 0.0;
    {
        const dim3 g((unsigned)(((size_t)H * pw + 8 * WORDS_PER_WARP - 1) / (8 * WORDS_PER_WARP)), n);
        if (type == BP_MASKS_LINEAR)
            threshold_kernel<BP_MASKS_LINEAR><<<g, 256>>>(pool, img_stride, H, W, C, pw, mm, c0, c1, c2, threshold, bits[0]);
        else
            threshold_kernel<BP_MASKS_EXCESS_GREEN><<<g, 256>>>(pool, img_stride, H, W, C, pw, mm, c0, c1, c2, threshold, bits[0]);
        CUM(cudaGetLastError());
    }
    int cur = 0;
    for (uint32_t fp : fps) {
        dilate_kernel<<<dim3((unsigned)(((size_t)H * pw + 255) / 256), n), 256>>>(bits[cur], bits[cur ^ 1], H, W, pw, fp);
        CUM(cudaGetLastError());
        cur ^= 1;
    }
    expand_bits_kernel<<<dim3((unsigned)(((size_t)H * W + 255) / 256), n), 256>>>(bits[cur], mask, H, W, pw);
    CUM(cudaGetLastError());
    CUM(cudaEventRecord(pl.e1, 0));
    if (masks_out)
        for (int i = 0; i < n; ++i)
            CUM(cudaMemcpyAsync(masks_out[i], mask + (size_t)i * H * W, (size_t)H * W, cudaMemcpyDeviceToHost, 0));
    CUM(cudaStreamSynchronize(0));
    if (ms) CUM(cudaEventElapsedTime(ms, pl.e0, pl.e1));
    if (out_dev) {
        *out_dev = mask;
        pl.release(mask);
    }
    return BP_OK;
}
