/* The C ABI without Python or torch: build a two-dimensional space by hand
 * (x ~ uniform(-5, 10), k ~ uniform(-3, 6, discrete)), draw n points into
 * device memory with ob2_sample, check them with ob2_contains, and print
 * them (hex floats, so a caller can compare bit for bit).
 *
 *   nvcc -O2 examples/c_abi_demo.c -Iinclude -Lorion_b200 -lob2 -o /tmp/c_abi_demo
 *   LD_LIBRARY_PATH=orion_b200 /tmp/c_abi_demo 8 20240001
 *
 * The slot table is what orion_b200/descriptor.py builds from a Space object:
 * Real.sample / _Discrete.sample / Dimension.__contains__ of the reference
 * (src/orion/algo/space.py:256-294, 150-164) then run as ob2_sample /
 * ob2_contains.  The fingerprint keys (hk) only matter for dedup; any odd
 * multipliers do for a demo.
 */
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "ob2.h"

#define CHECK_CUDA(call)                                                        \
    do {                                                                        \
        cudaError_t e_ = (call);                                                \
        if (e_ != cudaSuccess) {                                                \
            fprintf(stderr, "%s: %s\n", #call, cudaGetErrorString(e_));         \
            return 2;                                                           \
        }                                                                       \
    } while (0)

#define CHECK_OB2(call)                                                         \
    do {                                                                        \
        int rc_ = (call);                                                       \
        if (rc_ != OB2_OK) {                                                    \
            fprintf(stderr, "%s: %d %s\n", #call, rc_, ob2_last_error());       \
            return 3;                                                           \
        }                                                                       \
    } while (0)

static void fill_slot(ob2_slot_desc *d, int slot, int kind, double loc, double scale, double low, double high,
                      int col_units) {
    memset(d, 0, sizeof(*d));
    d->kind = kind;
    d->dist = OB2_DIST_UNIFORM;
    d->dim_first = slot;            /* scalar dimensions: every slot is its own dimension */
    d->flags = OB2_SLOT_DIM_LAST;
    d->col_units = col_units;
    d->t_op = OB2_OP_IDENTITY;
    d->t_units = col_units;
    d->t_width = 1;
    d->loc = loc;
    d->scale = scale;
    d->low = low;
    d->high = high;
    d->hk[0] = 0x9E3779B97F4A7C15ull * (uint64_t)(slot + 1);
    d->hk[1] = 0xD6E8FEB86659FD93ull | 0x0000000100000001ull;
    d->hk[2] = 0xC2B2AE3D27D4EB4Full | 0x0000000100000001ull;
}

int main(int argc, char **argv) {
    const long long n = argc > 1 ? atoll(argv[1]) : 8;
    const unsigned long long seed = argc > 2 ? strtoull(argv[2], NULL, 10) : 1;
    if (n < 1 || n > 1000000) {
        fprintf(stderr, "usage: %s [n in 1..1000000] [seed]\n", argv[0]);
        return 1;
    }
    const long long n_stride = (n + 31) / 32 * 32;

    ob2_slot_desc slots[2];
    fill_slot(&slots[0], 0, OB2_KIND_REAL, -5.0, 15.0, -5.0, 10.0, 0);     /* uniform(-5, 10): loc -5, scale 15 */
    fill_slot(&slots[1], 1, OB2_KIND_INTEGER, -3.0, 9.0, -3.0, 6.0, 8);    /* uniform(-3, 6), floored           */
    ob2_space *space = NULL;
    CHECK_OB2(ob2_space_create(slots, 2, NULL, 0, &space));
    const long long row_bytes = ob2_space_row_bytes(space);               /* 16 */

    void *cols = NULL;
    uint8_t *mask = NULL, *inside = NULL;
    uint32_t *n_failed = NULL;
    CHECK_CUDA(cudaMalloc(&cols, (size_t)(row_bytes * n_stride)));
    CHECK_CUDA(cudaMalloc((void **)&mask, (size_t)n_stride));
    CHECK_CUDA(cudaMalloc((void **)&inside, (size_t)n_stride));
    CHECK_CUDA(cudaMalloc((void **)&n_failed, 2 * sizeof(uint32_t)));
    CHECK_CUDA(cudaMemset(n_failed, 0, 2 * sizeof(uint32_t)));

    CHECK_OB2(ob2_sample(space, n, n_stride, seed, 0, NULL, cols, mask, NULL, n_failed, NULL));
    CHECK_OB2(ob2_contains(space, n, n_stride, cols, 0, inside, NULL));
    CHECK_CUDA(cudaDeviceSynchronize());

    double *x = (double *)malloc((size_t)n * sizeof(double));
    long long *k = (long long *)malloc((size_t)n * sizeof(long long));
    uint8_t *m = (uint8_t *)malloc((size_t)n), *in = (uint8_t *)malloc((size_t)n);
    CHECK_CUDA(cudaMemcpy(x, cols, (size_t)n * 8, cudaMemcpyDeviceToHost));                        /* column 0 */
    CHECK_CUDA(cudaMemcpy(k, (char *)cols + 8 * n_stride, (size_t)n * 8, cudaMemcpyDeviceToHost)); /* column 1 */
    CHECK_CUDA(cudaMemcpy(m, mask, (size_t)n, cudaMemcpyDeviceToHost));
    CHECK_CUDA(cudaMemcpy(in, inside, (size_t)n, cudaMemcpyDeviceToHost));
    for (long long i = 0; i < n; ++i) printf("%lld %a %lld %d %d\n", i, x[i], k[i], (int)m[i], (int)in[i]);

    /* the same draw through the single-call host entry point must give the same bytes */
    char *host_cols = (char *)malloc((size_t)(row_bytes * n_stride));
    uint32_t failed[2];
    if (n <= OB2_HOST_SAMPLE_MAX) {
        CHECK_OB2(ob2_sample_host(space, n, seed, 0, host_cols, failed, NULL));
        const int same = memcmp(host_cols, x, (size_t)n * 8) == 0 &&
                         memcmp(host_cols + 8 * n_stride, k, (size_t)n * 8) == 0 && failed[0] == 0 && failed[1] == 0;
        printf("host-entry %s\n", same ? "identical" : "DIFFERENT");
    }
    free(host_cols);
    free(x);
    free(k);
    free(m);
    free(in);
    cudaFree(cols);
    cudaFree(mask);
    cudaFree(inside);
    cudaFree(n_failed);
    ob2_space_destroy(space);
    return 0;
}
