/* Host memory bandwidth probe for the read-back design: how fast can the host's own threads write an
 * int32 volume - either a plain fill or the int8 -> int32 widening of a 1-byte-per-voxel state - compared with the
 * ~55 GB/s at which the GPU's DMA engine writes it over PCIe?  gcc -O3 -fopenmp tools/host_bw_probe.c -o /tmp/hbw */
#include <omp.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

int main(int argc, char **argv) {
    const size_t n = (size_t)1 << 29; /* 512 Mi voxels: 0.5 GiB in, 2 GiB out */
    int8_t *src = (int8_t *)malloc(n);
    int32_t *dst = (int32_t *)malloc(n * 4);
    if (!src || !dst) return 1;
    memset(src, 1, n);
    memset(dst, 0, n * 4); /* touch every page first */
    static const int32_t lut[4] = {-1, 1, 0, 0};
    for (int rep = 0; rep < 3; ++rep) {
        double t0 = omp_get_wtime();
#pragma omp parallel for schedule(static)
        for (size_t i = 0; i < n; ++i) dst[i] = -1;
        double t1 = omp_get_wtime();
#pragma omp parallel for schedule(static)
        for (size_t i = 0; i < n; ++i) dst[i] = lut[src[i] & 3];
        double t2 = omp_get_wtime();
        printf("threads %d: fill %.1f GB/s written; widen int8->int32 %.1f GB/s written (%.1f ms per GiB of int32)\n",
               omp_get_max_threads(), n * 4 / (t1 - t0) / 1e9, n * 4 / (t2 - t1) / 1e9, (t2 - t1) * 1e3 / (n * 4.0 / (1 << 30)));
    }
    return (int)(dst[n / 2] & 1) - 1;
}
