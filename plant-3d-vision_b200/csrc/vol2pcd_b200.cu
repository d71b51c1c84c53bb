// vol2pcd_b200.cu — volume -> point cloud on one B200, behind include/vol2pcd_b200.h.
// Replaces proc3d.vol2pcd (reference: plant3dvision/proc3d.py:490-570) and the multiclass gates of
// the PointCloud task (plant3dvision/tasks/proc3d.py:80-115).
//
// Design (B200-first, not a translation of the scipy pipeline):
//  * The reference builds five dense float64 volumes (two EDTs, three smoothed gradients) to read them
//    at the few voxels of the level-set band.  Here the ONLY dense data is one int16 per voxel:
//    sign = class (foreground > 0), magnitude = squared distance to the nearest voxel of the OTHER
//    class, capped.  One transform serves both of the reference's EDTs (dist for foreground voxels,
//    mdist for background voxels), since each voxel only ever needs the distance to the other class.
//  * The transform is separable (k, j, i passes) and windowed: the band and every tap the smoothing
//    reads lie within `cap` voxels of the surface (proof in DESIGN.md §9), so each pass scans at most
//    +/-cap neighbours and stops as soon as d*d exceeds the best candidate.  Values <= cap^2 are the
//    exact integer squared distances scipy's EDT produces; sqrt in float64 is correctly rounded on
//    both machines, so the signed distance is bit-identical.
//  * Band voxels are found with one ballot per 32-voxel word; an exclusive scan of the word
//    popcounts gives every band voxel its output slot in C order (np.nonzero order) without sorting.
//  * Only 8x8x8 bricks that contain band voxels are processed further: the brick's 18^3 signed
//    distances are staged in shared memory, np.gradient is evaluated on the fly, and the three
//    separable Gaussian passes run in shared memory in scipy's own summation order (symmetric
//    correlate1d: centre tap first, then (a[-j] + a[+j]) * w[j] from the outermost pair inwards,
//    'reflect' boundary), each pass rounded to float64 like the intermediate arrays scipy writes.
//    All float64 arithmetic uses explicit round-to-nearest intrinsics: nothing is contracted to FMA.
//  * No CPU fallback.
#include "vol2pcd_b200.h"

#include "bp_b200.h"
#include "bp_internal.h"

#include <cuda_runtime.h>

#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <thrust/iterator/counting_iterator.h>

#include <cmath>
#include <cstdio>
#include <new>

namespace {

constexpr int kFar = 32767;   // "no voxel of the other class within the window"
constexpr int kRad = 4;       // Gaussian radius of gaussian_filter(., 1): int(4.0 * 1 + 0.5)
constexpr int kBrick = 8;     // brick edge
constexpr int kReg = kBrick + 2 * (kRad + 1);  // 18: signed-distance region edge (halo 4 + 1 for the gradient)
constexpr int kMid = kBrick + 2 * kRad;        // 16: extent still to be filtered along an axis

#define CUV(call)                                                                              \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess)                                                                 \
            return bp_fail(BP_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_),   \
                           __FILE__, __LINE__);                                                \
    } while (0)

struct P3 {
    double x, y, z;
};

// ---------------------------------------------------------------------------------------------
// binarise (proc3d.py:515): class array, +kFar foreground / -kFar background
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void binarise_kernel(const T *__restrict__ vol, int16_t *__restrict__ s, size_t n, int nz,
                                long long pitch, int op, double val) {
    const size_t o = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= n) return;
    const size_t row = o / (size_t)nz;
    const int k = (int)(o - row * (size_t)nz);
    const double v = (double)vol[row * (size_t)pitch + k];
    bool fg;
    if (op == V2P_BIN_EQ) fg = (v == val);
    else if (op == V2P_BIN_GE) fg = (v >= val);
    else fg = (v > val);
    s[o] = fg ? (int16_t)kFar : (int16_t)-kFar;
}

// ---------------------------------------------------------------------------------------------
// one separable pass of the capped, two-sided squared Euclidean distance transform.
// in/out: sign = class, |.| = squared distance to the other class over the axes processed so far.
// For the voxel at position a of its line:  out = min_d ( cost(a+d) + d*d ),  cost(q) = 0 when q is
// of the other class, |in[q]| otherwise — the lower envelope of Felzenszwalb's separable EDT,
// evaluated by direct search because the window is tiny.
// ---------------------------------------------------------------------------------------------
//
// Most voxels of a real volume are far from any class boundary and keep the value "far" through all three
// passes.  They are recognised per 8^3 brick: `near` holds, for every brick, the OR of the class bits (1 =
// foreground present, 2 = background present) over all bricks within ceil(R / 8) bricks along every axis, i.e. over
// a neighbourhood that contains every voxel within Chebyshev distance R.  A voxel whose entry is not 3 has only
// its own class within reach: every value its window could read is "far", in every pass, so its result is its
// input and the window search (2 R loads) is skipped - exactly, not approximately.
__global__ void edt_pass_kernel(const int16_t *__restrict__ in, int16_t *__restrict__ out, size_t n,
                                int axis, int nx, int ny, int nz, int R, const uint32_t *__restrict__ near,
                                int nbj, int nbk) {
    const size_t o = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= n) return;
    const size_t t = o / (size_t)nz;
    const int k = (int)(o - t * (size_t)nz);
    const int i = (int)(t / (size_t)ny), j = (int)(t - (size_t)i * ny);
    const int s = in[o];
    if (near[((size_t)(i / kBrick) * nbj + (j / kBrick)) * nbk + (k / kBrick)] != 3u) {
        out[o] = (int16_t)s;
        return;
    }
    int a, len;
    long long stride;
    if (axis == 2) {
        a = k; len = nz; stride = 1;
    } else if (axis == 1) {
        a = j; len = ny; stride = nz;
    } else {
        a = i; len = nx; stride = (long long)nz * ny;
    }
    const bool fg = s > 0;
    int best = s > 0 ? s : -s;
    for (int d = 1; d <= R; ++d) {
        const int dd = d * d;
        if (dd >= best) break;
        if (a - d >= 0) {
            const int q = in[o - (size_t)d * stride];
            const int c = ((q > 0) != fg) ? 0 : (q > 0 ? q : -q);
            best = min(best, c + dd);
        }
        if (a + d < len) {
            const int q = in[o + (size_t)d * stride];
            const int c = ((q > 0) != fg) ? 0 : (q > 0 ? q : -q);
            best = min(best, c + dd);
        }
    }
    best = min(best, kFar);
    out[o] = (int16_t)(fg ? best : -best);
}

// class bits per 8^3 brick from the binarised volume: one warp per 32-voxel word of a row (= 4 bricks along k)
__global__ void brick_summary_kernel(const int16_t *__restrict__ s, uint32_t *__restrict__ summary, size_t n_words,
                                     int ny, int nz, int nw, int nbj, int nbk) {
    const size_t w = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= n_words) return;
    const size_t row = w / (size_t)nw;
    const int kw = (int)(w - row * (size_t)nw);
    const int k = kw * 32 + lane;
    const bool valid = k < nz;
    const bool fg = valid && s[row * (size_t)nz + k] > 0;
    const uint32_t fb = __ballot_sync(0xffffffffu, fg), vb = __ballot_sync(0xffffffffu, valid);
    if (lane < 4) {
        const uint32_t f = (fb >> (8 * lane)) & 0xffu, g = ((vb & ~fb) >> (8 * lane)) & 0xffu;
        const uint32_t v = (f ? 1u : 0u) | (g ? 2u : 0u);
        if (v) {
            const int i = (int)(row / (size_t)ny), j = (int)(row - (size_t)i * ny);
            uint32_t *p = summary + ((size_t)(i / kBrick) * nbj + (j / kBrick)) * nbk + (kw * 4 + lane);
            if ((*p & v) != v) atomicOr(p, v);
        }
    }
}

// out[b] = OR of in over the bricks within r along one axis (three passes = the box neighbourhood)
__global__ void brick_or_kernel(const uint32_t *__restrict__ in, uint32_t *__restrict__ out, int nbi, int nbj, int nbk,
                                int axis, int r) {
    const size_t b = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= (size_t)nbi * nbj * nbk) return;
    const int k = (int)(b % (size_t)nbk), j = (int)((b / (size_t)nbk) % (size_t)nbj), i = (int)(b / ((size_t)nbk * nbj));
    const int a = axis == 2 ? k : axis == 1 ? j : i, len = axis == 2 ? nbk : axis == 1 ? nbj : nbi;
    const long long stride = axis == 2 ? 1 : axis == 1 ? nbk : (long long)nbk * nbj;
    uint32_t v = 0;
    for (int d = max(-r, -a); d <= min(r, len - 1 - a) && v != 3u; ++d) v |= in[(long long)b + d * stride];
    out[b] = v;
}

// signed, half-voxel-shifted distance (proc3d.py:522): dist - 0.5 inside, -mdist + 0.5 outside
__device__ __forceinline__ double signed_dist(int s) {
    if (s > 0) return __dsub_rn(sqrt((double)s), 0.5);
    return __dadd_rn(-sqrt((double)(-s)), 0.5);
}

__global__ void signed_distance_kernel(const int16_t *__restrict__ s, double *__restrict__ out, size_t n) {
    const size_t o = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (o < n) out[o] = signed_dist(s[o]);
}

// ---------------------------------------------------------------------------------------------
// band detection (proc3d.py:536): one warp per 32-voxel word of a row.  bits / counts are indexed
// [row][word]; bricks holding at least one band voxel are flagged.
// ---------------------------------------------------------------------------------------------
__global__ void band_kernel(const int16_t *__restrict__ s, uint32_t *__restrict__ bits,
                            uint32_t *__restrict__ counts, uint8_t *__restrict__ brick_flag, size_t n_words,
                            int ny, int nz, int nw, int nbj, int nbk, double lo, double hi) {
    const size_t w = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= n_words) return;
    const size_t row = w / (size_t)nw;
    const int kw = (int)(w - row * (size_t)nw);
    const int k = kw * 32 + lane;
    bool on = false;
    if (k < nz) {
        const double d = signed_dist(s[row * (size_t)nz + k]);
        on = (d > lo) && (d <= hi);
    }
    const uint32_t word = __ballot_sync(0xffffffffu, on);
    if (lane == 0) {
        bits[w] = word;
        counts[w] = __popc(word);
    }
    if (word != 0 && lane < 4 && ((word >> (8 * lane)) & 0xffu)) {
        const int i = (int)(row / (size_t)ny), j = (int)(row - (size_t)i * ny);
        brick_flag[((size_t)(i / kBrick) * nbj + (j / kBrick)) * nbk + (kw * 4 + lane)] = 1;
    }
}

// scipy 'reflect' boundary (d c b a | a b c d | d c b a), any line length
__device__ __forceinline__ int reflect(int t, int n) {
    const int p = 2 * n;
    t %= p;
    if (t < 0) t += p;
    return t >= n ? p - 1 - t : t;
}

struct PointArgs {
    const int16_t *s;
    const uint32_t *bits;
    const uint32_t *offs;
    const uint32_t *bricks;  // linear ids of active bricks
    P3 *pts, *nrm;
    uint8_t *valid;
    unsigned long long *n_invalid;
    int nx, ny, nz, nw, nbj, nbk;
    double w[2 * kRad + 1];
    double ox, oy, oz, vs, lsv, half_diag;
};

// np.gradient component `c` (edge_order 1, unit spacing) at in-volume voxel (i, j, k), from the
// staged signed distances.  sd is indexed in region coordinates (origin r0 = brick origin - 5).
template <int C>
__device__ __forceinline__ double grad_at(const double *sd, int i, int j, int k, int ri, int rj, int rk,
                                          int n_c) {
    const int a = (C == 0) ? i : (C == 1 ? j : k);
    const int step = (C == 0) ? kReg * kReg : (C == 1 ? kReg : 1);
    const double *p = sd + ((i - ri) * kReg + (j - rj)) * kReg + (k - rk);
    if (a == 0) return __dsub_rn(p[step], p[0]);
    if (a == n_c - 1) return __dsub_rn(p[0], p[-step]);
    return __dmul_rn(__dsub_rn(p[step], p[-step]), 0.5);
}

// One Gaussian-smoothed gradient component for the whole brick -> gout[512]
template <int C>
__device__ void smooth_component(const PointArgs &A, const double *sd, double *o0, double *o1, double *gout,
                                 int bi, int bj, int bk) {
    const int ri = bi - (kRad + 1), rj = bj - (kRad + 1), rk = bk - (kRad + 1);
    const int n_c = (C == 0) ? A.nx : (C == 1 ? A.ny : A.nz);
    // pass along axis 0: o0[i][jr][kr], jr / kr cover [b - 4, b + 12)
    for (int t = threadIdx.x; t < kBrick * kMid * kMid; t += blockDim.x) {
        const int kr = t % kMid, jr = (t / kMid) % kMid, il = t / (kMid * kMid);
        const int i = bi + il, j = bj - kRad + jr, k = bk - kRad + kr;
        if (i >= A.nx || j < 0 || j >= A.ny || k < 0 || k >= A.nz) continue;
        double acc = __dmul_rn(grad_at<C>(sd, i, j, k, ri, rj, rk, n_c), A.w[kRad]);
#pragma unroll
        for (int q = -kRad; q < 0; ++q) {
            const double lo = grad_at<C>(sd, reflect(i + q, A.nx), j, k, ri, rj, rk, n_c);
            const double hi = grad_at<C>(sd, reflect(i - q, A.nx), j, k, ri, rj, rk, n_c);
            acc = __dadd_rn(acc, __dmul_rn(__dadd_rn(lo, hi), A.w[q + kRad]));
        }
        o0[t] = acc;
    }
    __syncthreads();
    // pass along axis 1: o1[i][jl][kr]
    for (int t = threadIdx.x; t < kBrick * kBrick * kMid; t += blockDim.x) {
        const int kr = t % kMid, jl = (t / kMid) % kBrick, il = t / (kMid * kBrick);
        const int i = bi + il, j = bj + jl, k = bk - kRad + kr;
        if (i >= A.nx || j >= A.ny || k < 0 || k >= A.nz) continue;
        const double *col = o0 + (il * kMid) * kMid + kr;  // + jr * kMid
        const int j0 = bj - kRad;
        double acc = __dmul_rn(col[(j - j0) * kMid], A.w[kRad]);
#pragma unroll
        for (int q = -kRad; q < 0; ++q) {
            const double lo = col[(reflect(j + q, A.ny) - j0) * kMid];
            const double hi = col[(reflect(j - q, A.ny) - j0) * kMid];
            acc = __dadd_rn(acc, __dmul_rn(__dadd_rn(lo, hi), A.w[q + kRad]));
        }
        o1[t] = acc;
    }
    __syncthreads();
    // pass along axis 2: gout[i][jl][kl]
    for (int t = threadIdx.x; t < kBrick * kBrick * kBrick; t += blockDim.x) {
        const int kl = t % kBrick, jl = (t / kBrick) % kBrick, il = t / (kBrick * kBrick);
        const int i = bi + il, j = bj + jl, k = bk + kl;
        if (i >= A.nx || j >= A.ny || k >= A.nz) continue;
        const double *line = o1 + (il * kBrick + jl) * kMid;
        const int k0 = bk - kRad;
        double acc = __dmul_rn(line[k - k0], A.w[kRad]);
#pragma unroll
        for (int q = -kRad; q < 0; ++q) {
            const double lo = line[reflect(k + q, A.nz) - k0];
            const double hi = line[reflect(k - q, A.nz) - k0];
            acc = __dadd_rn(acc, __dmul_rn(__dadd_rn(lo, hi), A.w[q + kRad]));
        }
        gout[t] = acc;
    }
    __syncthreads();
}

// proc3d.py:523-560 for one active brick
__global__ void __launch_bounds__(256)
points_kernel(const __grid_constant__ PointArgs A) {
    extern __shared__ double smem[];
    double *sd = smem;                              // kReg^3
    double *o0 = sd + kReg * kReg * kReg;           // kBrick * kMid * kMid
    double *o1 = o0 + kBrick * kMid * kMid;         // kBrick * kBrick * kMid
    double *g0 = o1 + kBrick * kBrick * kMid;       // 3 x kBrick^3
    double *g1 = g0 + kBrick * kBrick * kBrick;
    double *g2 = g1 + kBrick * kBrick * kBrick;

    const uint32_t b = A.bricks[blockIdx.x];
    const int bk = (int)(b % (uint32_t)A.nbk) * kBrick;
    const int bj = (int)((b / (uint32_t)A.nbk) % (uint32_t)A.nbj) * kBrick;
    const int bi = (int)(b / ((uint32_t)A.nbk * (uint32_t)A.nbj)) * kBrick;
    const int ri = bi - (kRad + 1), rj = bj - (kRad + 1), rk = bk - (kRad + 1);
    for (int t = threadIdx.x; t < kReg * kReg * kReg; t += blockDim.x) {
        const int k = rk + t % kReg, j = rj + (t / kReg) % kReg, i = ri + t / (kReg * kReg);
        if (i < 0 || i >= A.nx || j < 0 || j >= A.ny || k < 0 || k >= A.nz) continue;
        sd[t] = signed_dist(A.s[((size_t)i * A.ny + j) * A.nz + k]);
    }
    __syncthreads();
    smooth_component<0>(A, sd, o0, o1, g0, bi, bj, bk);
    smooth_component<1>(A, sd, o0, o1, g1, bi, bj, bk);
    smooth_component<2>(A, sd, o0, o1, g2, bi, bj, bk);

    unsigned long long bad = 0;
    for (int t = threadIdx.x; t < kBrick * kBrick * kBrick; t += blockDim.x) {
        const int kl = t % kBrick, jl = (t / kBrick) % kBrick, il = t / (kBrick * kBrick);
        const int i = bi + il, j = bj + jl, k = bk + kl;
        if (i >= A.nx || j >= A.ny || k >= A.nz) continue;
        const size_t w = ((size_t)i * A.ny + j) * A.nw + (k >> 5);
        const uint32_t word = A.bits[w];
        if (!((word >> (k & 31)) & 1u)) continue;
        const uint32_t slot = A.offs[w] + __popc(word & ((1u << (k & 31)) - 1u));
        const double gx = g0[t], gy = g1[t], gz = g2[t];
        const double sq = __dadd_rn(__dadd_rn(__dmul_rn(gx, gx), __dmul_rn(gy, gy)), __dmul_rn(gz, gz));
        const double norm = sqrt(sq);
        const bool ok = norm > 0.0;
        A.valid[slot] = ok ? 1 : 0;
        if (!ok) {
            ++bad;
            continue;
        }
        const double nx_ = __ddiv_rn(gx, norm), ny_ = __ddiv_rn(gy, norm), nz_ = __ddiv_rn(gz, norm);
        const double d = sd[((i - ri) * kReg + (j - rj)) * kReg + (k - rk)];
        const double val = __dsub_rn(__dadd_rn(d, A.lsv), A.half_diag);
        const double px = __dsub_rn((double)i, __dmul_rn(nx_, val));
        const double py = __dsub_rn((double)j, __dmul_rn(ny_, val));
        const double pz = __dsub_rn((double)k, __dmul_rn(nz_, val));
        P3 p;
        p.x = __dadd_rn(__dmul_rn(A.vs, px), A.ox);
        p.y = __dadd_rn(__dmul_rn(A.vs, py), A.oy);
        p.z = __dadd_rn(__dmul_rn(A.vs, pz), A.oz);
        A.pts[slot] = p;
        // normals = -grad_normalized, then PointCloud.normalize_normals()
        const double ax = -nx_, ay = -ny_, az = -nz_;
        const double nn = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(ax, ax), __dmul_rn(ay, ay)), __dmul_rn(az, az)));
        P3 q;
        q.x = __ddiv_rn(ax, nn);
        q.y = __ddiv_rn(ay, nn);
        q.z = __ddiv_rn(az, nn);
        A.nrm[slot] = q;
    }
    if (bad) atomicAdd(A.n_invalid, bad);
}

constexpr size_t kPointsSmem =
    sizeof(double) * (kReg * kReg * kReg + kBrick * kMid * kMid + kBrick * kBrick * kMid + 3 * kBrick * kBrick * kBrick);

// ---------------------------------------------------------------------------------------------
// tasks/proc3d.py:80-115
// ---------------------------------------------------------------------------------------------
constexpr int kMaxLabels = 32;
struct LabelPtrs {
    const void *p[kMaxLabels];
};

template <typename T>
__global__ void multiclass_kernel(const __grid_constant__ LabelPtrs L, int n_labels, size_t n, int background,
                                  double prior, double min_contrast, double min_score, int label,
                                  uint8_t *__restrict__ out) {
    const size_t o = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= n) return;
    double best = 0.0, other = 0.0, mine = 0.0;
    int arg = -1;
    bool have_other = false;
    for (int l = 0; l < n_labels; ++l) {
        double v = (double)static_cast<const T *>(L.p[l])[o];
        if (l == background) v = __dmul_rn(v, prior);
        if (arg < 0 || v > best) {  // np.argmax: first occurrence of the maximum
            best = v;
            arg = l;
        }
        if (l == label) {
            mine = v;
        } else if (!have_other || v > other) {
            other = v;
            have_other = true;
        }
    }
    (void)mine;
    bool pred = (arg == label);
    // `pred_c *= (pred_c > min_contrast * pred_no_c)`: the boolean itself (1.0 / 0.0) is compared
    if (min_contrast > 1.0) pred = pred && ((pred ? 1.0 : 0.0) > __dmul_rn(min_contrast, other));
    pred = pred && ((pred ? 1.0 : 0.0) > min_score);
    out[o] = pred ? 1 : 0;
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
struct v2p_result {
    int device = 0;
    int nx = 0, ny = 0, nz = 0;
    int cap = 0;
    int64_t n_points = 0, n_band = 0, launches = 0;
    float total_ms = 0.f, edt_ms = 0.f;
    P3 *pts = nullptr, *nrm = nullptr;
    int16_t *s = nullptr;  // the signed squared-distance volume (kept for bp_vol2pcd_signed_distance)
};

namespace {

size_t vol_elem(int dtype) {
    switch (dtype) {
        case V2P_U8: case V2P_I8: return 1;
        case V2P_I32: case V2P_F32: return 4;
        case V2P_F64: return 8;
    }
    return 0;
}

struct Scratch {  // freed on every exit path
    void *p[64];
    int n = 0;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    ~Scratch() {
        for (int q = 0; q < n; ++q) cudaFree(p[q]);
        for (auto &e : ev) if (e) cudaEventDestroy(e);
    }
    template <typename T>
    cudaError_t alloc(T **out, size_t bytes) {
        void *q = nullptr;
        cudaError_t e = cudaMalloc(&q, bytes ? bytes : 1);
        if (e == cudaSuccess) p[n++] = q;
        *out = static_cast<T *>(q);
        return e;
    }
    void release(void *q) {  // ownership moves to the result
        for (int a = 0; a < n; ++a) if (p[a] == q) { p[a] = p[--n]; return; }
    }
};

int binarise(const void *vol_d, int dtype, int16_t *s, size_t n, int nz, long long pitch, int op, double val) {
    const unsigned nb = (unsigned)((n + 255) / 256);
    switch (dtype) {
        case V2P_U8: binarise_kernel<uint8_t><<<nb, 256>>>((const uint8_t *)vol_d, s, n, nz, pitch, op, val); break;
        case V2P_I8: binarise_kernel<int8_t><<<nb, 256>>>((const int8_t *)vol_d, s, n, nz, pitch, op, val); break;
        case V2P_I32: binarise_kernel<int32_t><<<nb, 256>>>((const int32_t *)vol_d, s, n, nz, pitch, op, val); break;
        case V2P_F32: binarise_kernel<float><<<nb, 256>>>((const float *)vol_d, s, n, nz, pitch, op, val); break;
        default: binarise_kernel<double><<<nb, 256>>>((const double *)vol_d, s, n, nz, pitch, op, val);
    }
    CUV(cudaGetLastError());
    return BP_OK;
}

}  // namespace

extern "C" {

int bp_vol2pcd(v2p_result **out, int device, const void *vol, int vol_dtype, uint32_t flags,
               const int32_t shape[3], int64_t row_pitch, int bin_op, double bin_value,
               const double origin[3], double voxel_size, double level_set_value, double band_hi,
               double half_diag, const double *weights, int radius) {
    if (!out || !vol || !shape || !origin || !weights) return bp_fail(BP_EINVAL, "null argument");
    *out = nullptr;
    if (vol_elem(vol_dtype) == 0) return bp_fail(BP_EINVAL, "unknown vol_dtype %d", vol_dtype);
    if (bin_op < V2P_BIN_GT || bin_op > V2P_BIN_GE) return bp_fail(BP_EINVAL, "unknown bin_op %d", bin_op);
    if (radius != kRad) return bp_fail(BP_EINVAL, "radius %d: only gaussian_filter(., sigma=1) (radius 4) is built", radius);
    const int nx = shape[0], ny = shape[1], nz = shape[2];
    // np.gradient needs at least 2 samples per axis (numpy raises ValueError below that)
    if (nx < 2 || ny < 2 || nz < 2) return bp_fail(BP_EINVAL, "shape %d x %d x %d: every axis needs >= 2 voxels (np.gradient)", nx, ny, nz);
    const long long pitch = row_pitch ? row_pitch : nz;
    if (pitch < nz) return bp_fail(BP_EINVAL, "row_pitch %lld < nz %d", pitch, nz);
    const size_t n = (size_t)nx * ny * nz;
    if (!(level_set_value == level_set_value) || !(band_hi == band_hi))
        return bp_fail(BP_EINVAL, "level_set_value is NaN");
    // cap: every voxel the extraction reads is within |offset| <= sqrt(5^2+4^2+4^2) of a band voxel, whose own
    // distance to the other class is at most max(|lo|, |hi|) + 0.5
    const double lo = -level_set_value, hi = band_hi;
    const double E = std::fmax(std::fabs(lo), std::fabs(hi));
    const double capd = std::ceil(E + 0.5 + std::sqrt(57.0) + 1.0);
    if (!(capd <= 181.0)) return bp_fail(BP_EINVAL, "level_set_value %g needs a distance cap of %g voxels (max 181)", level_set_value, capd);
    const int R = (int)capd;
    int ndev = 0;
    CUV(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return bp_fail(BP_ECUDA, "CUDA device %d not available (%d visible)", device, ndev);
    CUV(cudaSetDevice(device));

    Scratch sc;
    v2p_result *r = new (std::nothrow) v2p_result();
    if (!r) return bp_fail(BP_ENOMEM, "out of host memory");
    struct Guard { v2p_result *r; ~Guard() { if (r) bp_vol2pcd_free(r); } } guard{r};
    r->device = device; r->nx = nx; r->ny = ny; r->nz = nz; r->cap = R;

    for (auto &e : sc.ev) CUV(cudaEventCreate(&e));
    const void *vol_d = vol;
    if (!(flags & V2P_FLAG_DEVICE_VOLUME)) {
        // host volume: one H2D copy (rows may be padded)
        const size_t bytes = (size_t)nx * ny * (size_t)pitch * vol_elem(vol_dtype);
        void *tmp = nullptr;
        cudaError_t e = sc.alloc(&tmp, bytes);
        if (e != cudaSuccess) return bp_fail(BP_ENOMEM, "cudaMalloc(%zu) for the volume failed: %s", bytes, cudaGetErrorString(e));
        CUV(cudaMemcpy(tmp, vol, bytes, cudaMemcpyHostToDevice));
        vol_d = tmp;
    }
    int16_t *s0 = nullptr, *s1 = nullptr;
    if (sc.alloc(&s0, n * 2) != cudaSuccess || sc.alloc(&s1, n * 2) != cudaSuccess)
        return bp_fail(BP_ENOMEM, "cudaMalloc(2 x %zu) for the distance volumes failed", n * 2);

    CUV(cudaEventRecord(sc.ev[0]));
    int rc = binarise(vol_d, vol_dtype, s0, n, nz, pitch, bin_op, bin_value);
    if (rc) return rc;
    const unsigned nb = (unsigned)((n + 255) / 256);
    CUV(cudaEventRecord(sc.ev[1]));
    // brick class summary and its box dilation by ceil(R / 8) bricks (see edt_pass_kernel)
    const int nw = (nz + 31) / 32;
    const size_t n_words = (size_t)nx * ny * nw;
    const int nbi = (nx + kBrick - 1) / kBrick, nbj = (ny + kBrick - 1) / kBrick, nbk = (nw * 32) / kBrick;
    const size_t n_bricks = (size_t)nbi * nbj * nbk;
    if (n_words >= 0x7fffffffull || n_bricks >= 0x7fffffffull)
        return bp_fail(BP_EINVAL, "volume too large for one call (%zu words)", n_words);
    uint32_t *near0 = nullptr, *near1 = nullptr;
    if (sc.alloc(&near0, n_bricks * 4) != cudaSuccess || sc.alloc(&near1, n_bricks * 4) != cudaSuccess)
        return bp_fail(BP_ENOMEM, "cudaMalloc for the brick summary failed");
    CUV(cudaMemsetAsync(near0, 0, n_bricks * 4));
    brick_summary_kernel<<<(unsigned)((n_words * 32 + 255) / 256), 256>>>(s0, near0, n_words, ny, nz, nw, nbj, nbk);
    {
        const int rb = (R + kBrick - 1) / kBrick;
        const unsigned gb = (unsigned)((n_bricks + 255) / 256);
        brick_or_kernel<<<gb, 256>>>(near0, near1, nbi, nbj, nbk, 2, rb);
        brick_or_kernel<<<gb, 256>>>(near1, near0, nbi, nbj, nbk, 1, rb);
        brick_or_kernel<<<gb, 256>>>(near0, near1, nbi, nbj, nbk, 0, rb);
    }
    const uint32_t *near = near1;
    edt_pass_kernel<<<nb, 256>>>(s0, s1, n, 2, nx, ny, nz, R, near, nbj, nbk);
    edt_pass_kernel<<<nb, 256>>>(s1, s0, n, 1, nx, ny, nz, R, near, nbj, nbk);
    edt_pass_kernel<<<nb, 256>>>(s0, s1, n, 0, nx, ny, nz, R, near, nbj, nbk);
    CUV(cudaGetLastError());
    CUV(cudaEventRecord(sc.ev[2]));
    r->launches += 8;
    int16_t *s = s1;

    // band words, output slots, active bricks
    uint32_t *bits = nullptr, *counts = nullptr, *offs = nullptr, *bricks = nullptr;
    uint8_t *brick_flag = nullptr;
    int *n_sel = nullptr;
    if (sc.alloc(&bits, n_words * 4) != cudaSuccess || sc.alloc(&counts, (n_words + 1) * 4) != cudaSuccess ||
        sc.alloc(&offs, (n_words + 1) * 4) != cudaSuccess || sc.alloc(&brick_flag, n_bricks) != cudaSuccess ||
        sc.alloc(&bricks, n_bricks * 4) != cudaSuccess || sc.alloc(&n_sel, 16) != cudaSuccess)
        return bp_fail(BP_ENOMEM, "cudaMalloc for the band index failed");
    CUV(cudaMemsetAsync(brick_flag, 0, n_bricks));
    CUV(cudaMemsetAsync(counts + n_words, 0, 4));
    band_kernel<<<(unsigned)((n_words * 32 + 255) / 256), 256>>>(s, bits, counts, brick_flag, n_words, ny, nz, nw,
                                                                 nbj, nbk, lo, hi);
    CUV(cudaGetLastError());
    r->launches++;
    {   // exclusive scan over n_words + 1 entries: offs[n_words] = total
        size_t tmp_bytes = 0;
        CUV(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, counts, offs, (int)(n_words + 1)));
        void *tmp = nullptr;
        if (sc.alloc(&tmp, tmp_bytes) != cudaSuccess) return bp_fail(BP_ENOMEM, "cudaMalloc(scan scratch) failed");
        CUV(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, counts, offs, (int)(n_words + 1)));
        size_t sel_bytes = 0;
        thrust::counting_iterator<uint32_t> ids(0);
        CUV(cub::DeviceSelect::Flagged(nullptr, sel_bytes, ids, brick_flag, bricks, n_sel, (int)n_bricks));
        void *tmp2 = nullptr;
        if (sc.alloc(&tmp2, sel_bytes) != cudaSuccess) return bp_fail(BP_ENOMEM, "cudaMalloc(select scratch) failed");
        CUV(cub::DeviceSelect::Flagged(tmp2, sel_bytes, ids, brick_flag, bricks, n_sel, (int)n_bricks));
        r->launches += 4;
    }
    uint32_t total = 0;
    int n_active = 0;
    CUV(cudaMemcpy(&total, offs + n_words, 4, cudaMemcpyDeviceToHost));
    CUV(cudaMemcpy(&n_active, n_sel, 4, cudaMemcpyDeviceToHost));
    r->n_band = total;

    P3 *pts = nullptr, *nrm = nullptr;
    uint8_t *valid = nullptr;
    unsigned long long *n_invalid = nullptr;
    if (sc.alloc(&pts, (size_t)total * sizeof(P3)) != cudaSuccess || sc.alloc(&nrm, (size_t)total * sizeof(P3)) != cudaSuccess ||
        sc.alloc(&valid, total) != cudaSuccess || sc.alloc(&n_invalid, 8) != cudaSuccess)
        return bp_fail(BP_ENOMEM, "cudaMalloc for %u points failed", total);
    CUV(cudaMemsetAsync(n_invalid, 0, 8));
    unsigned long long bad = 0;
    if (n_active > 0) {
        PointArgs A;
        A.s = s; A.bits = bits; A.offs = offs; A.bricks = bricks; A.pts = pts; A.nrm = nrm; A.valid = valid;
        A.n_invalid = n_invalid;
        A.nx = nx; A.ny = ny; A.nz = nz; A.nw = nw; A.nbj = nbj; A.nbk = nbk;
        for (int q = 0; q < 2 * kRad + 1; ++q) A.w[q] = weights[q];
        A.ox = origin[0]; A.oy = origin[1]; A.oz = origin[2]; A.vs = voxel_size;
        A.lsv = level_set_value; A.half_diag = half_diag;
        CUV(cudaFuncSetAttribute(points_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPointsSmem));
        points_kernel<<<(unsigned)n_active, 256, kPointsSmem>>>(A);
        CUV(cudaGetLastError());
        r->launches++;
        CUV(cudaMemcpy(&bad, n_invalid, 8, cudaMemcpyDeviceToHost));
    }
    r->n_points = (int64_t)total - (int64_t)bad;
    if (bad > 0) {
        // drop the points whose smoothed gradient vanished (proc3d.py:545, :558-560), keeping the order
        P3 *pts2 = nullptr, *nrm2 = nullptr;
        if (sc.alloc(&pts2, (size_t)r->n_points * sizeof(P3)) != cudaSuccess ||
            sc.alloc(&nrm2, (size_t)r->n_points * sizeof(P3)) != cudaSuccess)
            return bp_fail(BP_ENOMEM, "cudaMalloc for %lld points failed", (long long)r->n_points);
        size_t sel_bytes = 0;
        CUV(cub::DeviceSelect::Flagged(nullptr, sel_bytes, pts, valid, pts2, n_sel, (int)total));
        void *tmp = nullptr;
        if (sc.alloc(&tmp, sel_bytes) != cudaSuccess) return bp_fail(BP_ENOMEM, "cudaMalloc(select scratch) failed");
        CUV(cub::DeviceSelect::Flagged(tmp, sel_bytes, pts, valid, pts2, n_sel, (int)total));
        CUV(cub::DeviceSelect::Flagged(tmp, sel_bytes, nrm, valid, nrm2, n_sel, (int)total));
        r->launches += 4;
        pts = pts2;
        nrm = nrm2;
    }
    CUV(cudaEventRecord(sc.ev[3]));
    CUV(cudaDeviceSynchronize());
    CUV(cudaEventElapsedTime(&r->total_ms, sc.ev[0], sc.ev[3]));
    CUV(cudaEventElapsedTime(&r->edt_ms, sc.ev[1], sc.ev[2]));
    sc.release(pts); sc.release(nrm); sc.release(s);
    r->pts = pts; r->nrm = nrm; r->s = s;
    guard.r = nullptr;
    *out = r;
    return BP_OK;
}

int bp_vol2pcd_info(const v2p_result *r, int64_t *n_points, float *total_ms, float *edt_ms,
                    int64_t *n_band_voxels, int64_t *launches) {
    if (!r) return bp_fail(BP_EINVAL, "null result");
    if (n_points) *n_points = r->n_points;
    if (total_ms) *total_ms = r->total_ms;
    if (edt_ms) *edt_ms = r->edt_ms;
    if (n_band_voxels) *n_band_voxels = r->n_band;
    if (launches) *launches = r->launches;
    return BP_OK;
}

int bp_vol2pcd_fetch(const v2p_result *r, double *points, double *normals) {
    if (!r) return bp_fail(BP_EINVAL, "null result");
    CUV(cudaSetDevice(r->device));
    const size_t bytes = (size_t)r->n_points * sizeof(P3);
    if (bytes == 0) return BP_OK;
    if (points) CUV(cudaMemcpy(points, r->pts, bytes, cudaMemcpyDeviceToHost));
    if (normals) CUV(cudaMemcpy(normals, r->nrm, bytes, cudaMemcpyDeviceToHost));
    return BP_OK;
}

int bp_vol2pcd_device(const v2p_result *r, void **points, void **normals) {
    if (!r) return bp_fail(BP_EINVAL, "null result");
    if (points) *points = r->pts;
    if (normals) *normals = r->nrm;
    return BP_OK;
}

int bp_vol2pcd_signed_distance(const v2p_result *r, double *host_dst, double *cap) {
    if (!r || !host_dst) return bp_fail(BP_EINVAL, "null argument");
    CUV(cudaSetDevice(r->device));
    const size_t n = (size_t)r->nx * r->ny * r->nz;
    double *d = nullptr;
    cudaError_t e = cudaMalloc((void **)&d, n * 8);
    if (e != cudaSuccess) return bp_fail(BP_ENOMEM, "cudaMalloc(%zu) failed: %s", n * 8, cudaGetErrorString(e));
    signed_distance_kernel<<<(unsigned)((n + 255) / 256), 256>>>(r->s, d, n);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpy(host_dst, d, n * 8, cudaMemcpyDeviceToHost);
    cudaFree(d);
    if (e != cudaSuccess) return bp_fail(BP_ECUDA, "signed distance readback failed: %s", cudaGetErrorString(e));
    if (cap) *cap = (double)r->cap;
    return BP_OK;
}

int bp_vol2pcd_free(v2p_result *r) {
    if (!r) return BP_OK;
    cudaSetDevice(r->device);
    cudaFree(r->pts);
    cudaFree(r->nrm);
    cudaFree(r->s);
    delete r;
    return BP_OK;
}

int bp_multiclass_select(int device, int n_labels, const void *const *vols, int vol_dtype, uint32_t flags,
                         const int32_t shape[3], int background, double background_prior,
                         double min_contrast, double min_score, int label, void **out_dev) {
    if (!vols || !shape || !out_dev) return bp_fail(BP_EINVAL, "null argument");
    *out_dev = nullptr;
    if (vol_dtype != V2P_F32 && vol_dtype != V2P_F64) return bp_fail(BP_EINVAL, "label volumes must be float32 or float64");
    if (n_labels < 2 || n_labels > kMaxLabels) return bp_fail(BP_EINVAL, "n_labels %d outside [2, %d]", n_labels, kMaxLabels);
    if (label < 0 || label >= n_labels || label == background) return bp_fail(BP_EINVAL, "bad label index %d", label);
    int ndev = 0;
    CUV(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return bp_fail(BP_ECUDA, "CUDA device %d not available (%d visible)", device, ndev);
    CUV(cudaSetDevice(device));
    const size_t n = (size_t)shape[0] * shape[1] * shape[2];
    const size_t es = vol_elem(vol_dtype);
    Scratch sc;
    LabelPtrs L;
    for (int l = 0; l < n_labels; ++l) {
        if (!vols[l]) return bp_fail(BP_EINVAL, "label volume %d is null", l);
        if (flags & V2P_FLAG_DEVICE_VOLUME) {
            L.p[l] = vols[l];
        } else {
            void *d = nullptr;
            if (sc.alloc(&d, n * es) != cudaSuccess) return bp_fail(BP_ENOMEM, "cudaMalloc(%zu) for label volume %d failed", n * es, l);
            CUV(cudaMemcpy(d, vols[l], n * es, cudaMemcpyHostToDevice));
            L.p[l] = d;
        }
    }
    uint8_t *out = nullptr;
    cudaError_t e = cudaMalloc((void **)&out, n ? n : 1);
    if (e != cudaSuccess) return bp_fail(BP_ENOMEM, "cudaMalloc(%zu) failed: %s", n, cudaGetErrorString(e));
    if (n) {
        const unsigned nb = (unsigned)((n + 255) / 256);
        if (vol_dtype == V2P_F32)
            multiclass_kernel<float><<<nb, 256>>>(L, n_labels, n, background, background_prior, min_contrast, min_score, label, out);
        else
            multiclass_kernel<double><<<nb, 256>>>(L, n_labels, n, background, background_prior, min_contrast, min_score, label, out);
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaDeviceSynchronize();
        if (e != cudaSuccess) {
            cudaFree(out);
            return bp_fail(BP_ECUDA, "multiclass_kernel failed: %s", cudaGetErrorString(e));
        }
    }
    *out_dev = out;
    return BP_OK;
}

int bp_device_free(int device, void *ptr) {
    if (!ptr) return BP_OK;
    CUV(cudaSetDevice(device));
    CUV(cudaFree(ptr));
    return BP_OK;
}

}  // extern "C"
