// tail_static_b200.cu — bp_tail_static: the fused task tail (tail_kernel of bp_kernels.cu: raw state -> the volume the
// caller reads + min / max / NaN of the raw values) for callers that may not allocate: code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the reference.
// Nothing here calls cudaMalloc / cudaFree or any allocating routine.
//
// The per-block partials of the reduction live in ONE file-scope device array of fixed size (768 KiB) that every GPU
// that loads the library holds until exit.  Every call on a device shares it and it keeps the last call's contents:
// the kernel overwrites every partial the reduction then reads, so nothing needs clearing, but two calls that overlap
// on one GPU (different streams, no ordering) race on it.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <mutex>

#include "bp_b200.h"
#include "bp_internal.h"

namespace {

constexpr int kThreads = 256;
constexpr int kPerThread = 16;                          // values per thread
constexpr int kPerBlock = kThreads * kPerThread;        // 4096 values per block, one partial each
constexpr long long kMaxBlocks = BP_TAIL_STATIC_MAX_VALUES / kPerBlock;  // 32768
constexpr int kMaxDevices = 64;

// (min, max, nan) per block.  In HOST code this name is a host-side stand-in, not a device address: the address the
// kernels get comes from cudaGetSymbolAddress, once per device (scratch_of).
static __device__ __align__(256) double g_tail_partial[3 * kMaxBlocks];

double *g_scratch_ptr[kMaxDevices];  // per device, filled by scratch_of
std::mutex g_scratch_mutex;

// The device address of g_tail_partial on `device` (which must be current).
int scratch_of(int device, double **out) {
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (g_scratch_ptr[device] == nullptr) {
        double *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_tail_partial);
        if (e != cudaSuccess || p == nullptr)
            return bp_fail(BP_ECUDA, "bp_tail_static: cudaGetSymbolAddress on device %d failed: %s", device, cudaGetErrorString(e));
        g_scratch_ptr[device] = p;
    }
    *out = g_scratch_ptr[device];
    return BP_OK;
}

// tail_kernel of bp_kernels.cu with a fixed 4096 values per block (so the number of partials is a function of the
// size alone) - same arithmetic per value, same order-free min / max.
template <typename OutT, bool CARVE>
__global__ void __launch_bounds__(kThreads)
tail_static_kernel(const void *__restrict__ grid, OutT *__restrict__ out, long long n, int nz, int nzp, int v_carved,
                   int v_seen, int v_unseen, int exp_clip, double *__restrict__ partial) {
    double lo = INFINITY, hi = -INFINITY;
    int nan = 0;
    const long long base = (long long)blockIdx.x * kPerBlock + threadIdx.x;
#pragma unroll 4
    for (int q = 0; q < kPerThread; ++q) {
        const long long o = base + (long long)q * kThreads;
        if (o >= n) break;
        const long long row = o / nz;
        const int k = (int)(o - row * nz);
        double v;
        if (CARVE) {
            const int s = ((const int8_t *)grid)[row * nzp + k];
            v = (double)((s == 0) ? v_carved : (s == 1 ? v_seen : v_unseen));
        } else {
            v = (double)((const float *)grid)[row * nzp + k];
        }
        nan |= (v != v);
        lo = fmin(lo, v);
        hi = fmax(hi, v);
        if (exp_clip) {
            // np.exp runs in the array's dtype: round to it before the clip (tasks/cl.py:172-174)
            v = sizeof(OutT) == 8 ? exp(v) : (double)(float)exp(v);
            if (v > 1.0) v = 1.0;
        }
        out[o] = (OutT)v;
    }
    for (int d = 16; d; d >>= 1) {
        lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, d));
        hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, d));
        nan |= __shfl_xor_sync(0xffffffffu, nan, d);
    }
    __shared__ double sl[8], sh[8];
    __shared__ int sn[8];
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) {
        sl[w] = lo;
        sh[w] = hi;
        sn[w] = nan;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int a = 1; a < 8; ++a) {
            lo = fmin(lo, sl[a]);
            hi = fmax(hi, sh[a]);
            nan |= sn[a];
        }
        partial[3 * blockIdx.x] = lo;  // every partial the reduction reads is written here: no stale contents
        partial[3 * blockIdx.x + 1] = hi;
        partial[3 * blockIdx.x + 2] = (double)nan;
    }
}

// One block folds the n_partials partials into stats = {min, max} of the raw values, NaN for both if any value is NaN
// (numpy's min / max propagate NaN), as bp_get_values_tail reports them.
__global__ void __launch_bounds__(kThreads)
tail_static_reduce_kernel(const double *__restrict__ partial, int n_partials, double *__restrict__ stats) {
    double lo = INFINITY, hi = -INFINITY;
    int nan = 0;
    for (int b = threadIdx.x; b < n_partials; b += kThreads) {
        lo = fmin(lo, partial[3 * b]);
        hi = fmax(hi, partial[3 * b + 1]);
        nan |= partial[3 * b + 2] != 0.0;
    }
    for (int d = 16; d; d >>= 1) {
        lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, d));
        hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, d));
        nan |= __shfl_xor_sync(0xffffffffu, nan, d);
    }
    __shared__ double sl[8], sh[8];
    __shared__ int sn[8];
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) {
        sl[w] = lo;
        sh[w] = hi;
        sn[w] = nan;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int a = 1; a < 8; ++a) {
            lo = fmin(lo, sl[a]);
            hi = fmax(hi, sh[a]);
            nan |= sn[a];
        }
        stats[0] = nan ? NAN : lo;
        stats[1] = nan ? NAN : hi;
    }
}

}  // namespace

extern "C" int bp_tail_static(int device, int mode, const void *state_dev, int64_t n_rows, int32_t nz, int32_t nz_pitch,
                              const int32_t *codes, uint32_t flags, void *out_dev, double *stats_dev, void *stream) {
    if (!state_dev || !out_dev || !stats_dev) return bp_fail(BP_EINVAL, "bp_tail_static: null argument");
    if (mode != BP_MODE_CARVING && mode != BP_MODE_AVERAGING) return bp_fail(BP_EINVAL, "bp_tail_static: unknown mode %d", mode);
    if (flags & ~(BP_TAIL_F64 | BP_TAIL_EXP_CLIP)) return bp_fail(BP_EINVAL, "bp_tail_static: unknown tail flags 0x%x", flags);
    const bool carve = mode == BP_MODE_CARVING;
    if (carve && (flags & BP_TAIL_EXP_CLIP))
        return bp_fail(BP_EINVAL, "bp_tail_static: exp/clip applies to averaging volumes only (tasks/cl.py:172)");
    if (carve && !codes) return bp_fail(BP_EINVAL, "bp_tail_static: carving needs the three output codes");
    if (n_rows < 1 || nz < 1 || nz_pitch < nz)
        return bp_fail(BP_EINVAL, "bp_tail_static: bad shape (%lld rows of %d values, pitch %d)", (long long)n_rows, nz, nz_pitch);
    if (device < 0 || device >= kMaxDevices) return bp_fail(BP_EINVAL, "bp_tail_static: device %d out of range", device);
    if (n_rows > BP_TAIL_STATIC_MAX_VALUES / nz)
        return bp_fail(BP_EINVAL,
                       "bp_tail_static: %lld x %d values exceed the static scratch's limit of %lld values (BP_TAIL_STATIC_MAX_VALUES); "
                       "use bp_get_values_tail",
                       (long long)n_rows, nz, (long long)BP_TAIL_STATIC_MAX_VALUES);
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return bp_fail(BP_ECUDA, "bp_tail_static: cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    double *partial = nullptr;
    const int rc = scratch_of(device, &partial);
    if (rc) return rc;
    const long long n = (long long)n_rows * nz;
    const int nb = (int)((n + kPerBlock - 1) / kPerBlock);
    cudaStream_t st = (cudaStream_t)stream;
    const bool f64 = (flags & BP_TAIL_F64) != 0;
    const int ec = (flags & BP_TAIL_EXP_CLIP) ? 1 : 0;
    if (carve) {
        if (f64) tail_static_kernel<double, true><<<nb, kThreads, 0, st>>>(state_dev, (double *)out_dev, n, nz, nz_pitch, codes[0], codes[1], codes[2], 0, partial);
        else tail_static_kernel<int32_t, true><<<nb, kThreads, 0, st>>>(state_dev, (int32_t *)out_dev, n, nz, nz_pitch, codes[0], codes[1], codes[2], 0, partial);
    } else {
        if (f64) tail_static_kernel<double, false><<<nb, kThreads, 0, st>>>(state_dev, (double *)out_dev, n, nz, nz_pitch, 0, 0, 0, ec, partial);
        else tail_static_kernel<float, false><<<nb, kThreads, 0, st>>>(state_dev, (float *)out_dev, n, nz, nz_pitch, 0, 0, 0, ec, partial);
    }
    tail_static_reduce_kernel<<<1, kThreads, 0, st>>>(partial, nb, stats_dev);
    e = cudaGetLastError();
    if (e != cudaSuccess) return bp_fail(BP_ECUDA, "bp_tail_static: launch failed: %s", cudaGetErrorString(e));
    return BP_OK;
}
