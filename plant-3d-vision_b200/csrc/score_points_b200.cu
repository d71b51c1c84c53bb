// score_points_b200.cu — point -> label-map scoring on one B200, behind bp_score_points of
// include/vol2pcd_b200.h.  Replaces the scoring loops of the reference's SegmentedPointCloud task
// (plant3dvision/tasks/proc3d.py:203-228: a Python loop over points inside a loop over files) and
// proc3d.backproject_points (plant3dvision/proc3d.py:655-659).
//
// One thread owns one point for ALL views: its world coordinates stay in registers, the views (camera + map
// pointer + label) are read from a device table, and its per-label sums live in global memory [label][point]
// (coalesced across the warp).  Maps are uint8, so the float64 sums the reference forms are exact integers:
// they are accumulated as uint32 and converted at the end.  Projection is float64 with separate multiplies
// and adds (no FMA), in the operand order of the reference's matrix products.
#include "vol2pcd_b200.h"

#include "bp_b200.h"
#include "bp_internal.h"

#include <cuda_runtime.h>

#include <cmath>
#include <vector>

namespace {

struct ScoreView {  // 160 bytes
    double r[9], t[3], fx, fy, cx, cy;
    const uint8_t *map;
    int H, W, label, pad;
};

#define CUS(call)                                                                              \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess)                                                                 \
            return bp_fail(BP_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_),   \
                           __FILE__, __LINE__);                                                \
    } while (0)

__global__ void __launch_bounds__(256)
score_points_kernel(const double *__restrict__ pts, long long n, const ScoreView *__restrict__ views, int n_views,
                    uint32_t *__restrict__ scores) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double X = pts[3 * i], Y = pts[3 * i + 1], Z = pts[3 * i + 2];
    for (int v = 0; v < n_views; ++v) {
        const ScoreView &V = views[v];
        // x = rot @ p + tvec ; x = K @ x ; x / x[2]   (proc3d.py:656-658)
        const double xc = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(V.r[0], X), __dmul_rn(V.r[1], Y)), __dmul_rn(V.r[2], Z)), V.t[0]);
        const double yc = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(V.r[3], X), __dmul_rn(V.r[4], Y)), __dmul_rn(V.r[5], Z)), V.t[1]);
        const double zc = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(V.r[6], X), __dmul_rn(V.r[7], Y)), __dmul_rn(V.r[8], Z)), V.t[2]);
        // K rows (fx 0 cx), (0 fy cy), (0 0 1): the products with the zero entries add exact zeros
        const double u = __ddiv_rn(__dadd_rn(__dmul_rn(V.fx, xc), __dmul_rn(V.cx, zc)), zc);
        const double w = __ddiv_rn(__dadd_rn(__dmul_rn(V.fy, yc), __dmul_rn(V.cy, zc)), zc);
        // np.asarray(. + 0.5, dtype=int) truncates toward zero; is_in_pict (tasks/proc3d.py:182-183).
        // Non-finite values become INT_MIN in numpy and fail the test; the comparisons below fail them too.
        const double uu = __dadd_rn(u, 0.5), ww = __dadd_rn(w, 0.5);
        if (!(uu > -1.0) || !(uu < (double)V.W) || !(ww > -1.0) || !(ww < (double)V.H)) continue;
        const int px = (int)uu, py = (int)ww;  // C casts truncate toward zero like numpy's
        scores[(size_t)V.label * (size_t)n + (size_t)i] += V.map[(size_t)py * V.W + px];
    }
}

__global__ void finish_scores_kernel(const uint32_t *__restrict__ acc, long long n, int n_labels,
                                     double *__restrict__ scores, int32_t *__restrict__ labels) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t best = 0;
    int arg = 0;
    for (int l = 0; l < n_labels; ++l) {
        const uint32_t s = acc[(size_t)l * (size_t)n + (size_t)i];
        if (scores) scores[(size_t)l * (size_t)n + (size_t)i] = (double)s;
        if (l == 0 || s > best) {  // np.argmax: first maximum
            best = s;
            arg = l;
        }
    }
    labels[i] = arg;
}

struct Frees {
    std::vector<void *> p;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    ~Frees() {
        for (void *q : p) cudaFree(q);
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
};

}  // namespace

extern "C" int bp_score_points(int device, int64_t n_points, const double *points, int32_t n_views, const double *cams,
                               const void *const *maps, const int32_t *hw, const int32_t *view_label,
                               int32_t n_labels, double *scores, int32_t *labels, float *ms) {
    if (ms) *ms = 0.f;
    if (n_points < 0 || n_views < 0 || n_labels < 1) return bp_fail(BP_EINVAL, "bad sizes");
    if (n_points == 0) return BP_OK;
    if (!points || !labels || (n_views > 0 && (!cams || !maps || !hw || !view_label)))
        return bp_fail(BP_EINVAL, "null argument");
    if ((double)n_points * n_labels >= 4.0e9) return bp_fail(BP_EINVAL, "too many points x labels for one call");
    int ndev = 0;
    CUS(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return bp_fail(BP_ECUDA, "CUDA device %d not available (%d visible)", device, ndev);
    CUS(cudaSetDevice(device));
    Frees fr;
    std::vector<ScoreView> hv((size_t)n_views);
    std::vector<size_t> off((size_t)n_views);
    size_t pool_bytes = 0;
    for (int v = 0; v < n_views; ++v) {
        const int H = hw[2 * v], W = hw[2 * v + 1];
        if (!maps[v] || H <= 0 || W <= 0) return bp_fail(BP_EINVAL, "map %d is null or empty", v);
        if (view_label[v] < 0 || view_label[v] >= n_labels) return bp_fail(BP_EINVAL, "view %d has label %d", v, view_label[v]);
        off[v] = pool_bytes;
        pool_bytes += ((size_t)H * W + 255) & ~(size_t)255;
    }
    // one allocation for all label maps (a cudaMalloc per map costs more than the scoring kernel)
    uint8_t *pool = nullptr;
    if (fr.alloc(&pool, pool_bytes) != cudaSuccess) return bp_fail(BP_ENOMEM, "cudaMalloc(%zu) for %d label maps failed", pool_bytes, n_views);
    for (int v = 0; v < n_views; ++v) {
        const int H = hw[2 * v], W = hw[2 * v + 1];
        const double *c = cams + 16 * (size_t)v;
        ScoreView &S = hv[v];
        S.fx = c[0]; S.fy = c[1]; S.cx = c[2]; S.cy = c[3];
        for (int a = 0; a < 9; ++a) S.r[a] = c[4 + a];
        for (int a = 0; a < 3; ++a) S.t[a] = c[13 + a];
        S.H = H; S.W = W; S.label = view_label[v]; S.pad = 0;
        CUS(cudaMemcpyAsync(pool + off[v], maps[v], (size_t)H * W, cudaMemcpyHostToDevice, 0));
        S.map = pool + off[v];
    }
    if (n_views > 16000000) return bp_fail(BP_EINVAL, "too many views");
    double *pts_d = nullptr, *scores_d = nullptr;
    uint32_t *acc = nullptr;
    int32_t *labels_d = nullptr;
    ScoreView *views_d = nullptr;
    const size_t cells = (size_t)n_points * (size_t)n_labels;
    if (fr.alloc(&pts_d, (size_t)n_points * 24) != cudaSuccess || fr.alloc(&acc, cells * 4) != cudaSuccess ||
        fr.alloc(&labels_d, (size_t)n_points * 4) != cudaSuccess || fr.alloc(&views_d, hv.size() * sizeof(ScoreView)) != cudaSuccess ||
        (scores && fr.alloc(&scores_d, cells * 8) != cudaSuccess))
        return bp_fail(BP_ENOMEM, "cudaMalloc for %lld points x %d labels failed", (long long)n_points, n_labels);
    CUS(cudaMemcpyAsync(pts_d, points, (size_t)n_points * 24, cudaMemcpyHostToDevice, 0));
    if (n_views) CUS(cudaMemcpyAsync(views_d, hv.data(), hv.size() * sizeof(ScoreView), cudaMemcpyHostToDevice, 0));
    CUS(cudaMemsetAsync(acc, 0, cells * 4, 0));
    CUS(cudaEventCreate(&fr.e0));
    CUS(cudaEventCreate(&fr.e1));
    const unsigned nb = (unsigned)((n_points + 255) / 256);
    CUS(cudaEventRecord(fr.e0, 0));
    if (n_views) score_points_kernel<<<nb, 256>>>(pts_d, (long long)n_points, views_d, n_views, acc);
    CUS(cudaGetLastError());
    CUS(cudaEventRecord(fr.e1, 0));
    finish_scores_kernel<<<nb, 256>>>(acc, (long long)n_points, n_labels, scores_d, labels_d);
    CUS(cudaGetLastError());
    CUS(cudaMemcpy(labels, labels_d, (size_t)n_points * 4, cudaMemcpyDeviceToHost));
    if (scores) CUS(cudaMemcpy(scores, scores_d, cells * 8, cudaMemcpyDeviceToHost));
    if (ms) CUS(cudaEventElapsedTime(ms, fr.e0, fr.e1));
    return BP_OK;
}
