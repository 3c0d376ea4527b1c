// Micro-benchmark: what is the HBM write ceiling for the column-streaming pattern of the
// suggest kernel (every thread stores a few bytes to each of C far-apart columns)?
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a write_ceiling.cu -o write_ceiling && ./write_ceiling
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int BYTES>  // bytes per thread per column: 16 (st.v2.f64) or 32 (st.v4.f64)
__global__ void __launch_bounds__(256) stream_cols(char *out, int cols, int64_t n_stride, int64_t n) {
    const int64_t i0 = ((int64_t)blockIdx.x * 256 + threadIdx.x) * (BYTES / 8);
    if (i0 >= n) return;
    char *p = out + i0 * 8;
    const double v = (double)i0;
    for (int c = 0; c < cols; ++c) {
        if (BYTES == 16) {
            asm volatile("st.global.cs.v2.f64 [%0], {%1, %1};" ::"l"(p), "d"(v) : "memory");
        } else {
            asm volatile("st.global.cs.v4.f64 [%0], {%1, %1, %1, %1};" ::"l"(p), "d"(v) : "memory");
        }
        p += n_stride * 8;
    }
}

// row-major comparison: each warp writes one long contiguous run
__global__ void __launch_bounds__(256) stream_flat(double *out, int64_t total) {
    const int64_t stride = (int64_t)gridDim.x * 256 * 2;
    for (int64_t i = ((int64_t)blockIdx.x * 256 + threadIdx.x) * 2; i < total; i += stride)
        asm volatile("st.global.cs.v2.f64 [%0], {%1, %1};" ::"l"(out + i), "d"(1.0) : "memory");
}

int main() {
    const int64_t n = 12500000, n_stride = (n + 31) / 32 * 32;
    const int cols = 290;  // 2320 B per candidate, like ob2_suggest on mixed64
    char *buf;
    const size_t bytes = (size_t)cols * n_stride * 8;
    cudaMalloc(&buf, bytes);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float ms;
    for (int variant = 0; variant < 3; ++variant) {
        for (int rep = 0; rep < 4; ++rep) {
            cudaEventRecord(e0);
            if (variant == 0) stream_cols<16><<<(unsigned)((n / 2 + 255) / 256), 256>>>(buf, cols, n_stride, n);
            if (variant == 1) stream_cols<32><<<(unsigned)((n / 4 + 255) / 256), 256>>>(buf, cols, n_stride, n);
            if (variant == 2) stream_flat<<<148 * 8, 256>>>((double *)buf, (int64_t)cols * n_stride);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            cudaEventElapsedTime(&ms, e0, e1);
        }
        printf("%s: %.3f ms, %.2f TB/s (%s)\n", variant == 0 ? "columns, 16 B/thread" : variant == 1 ? "columns, 32 B/thread" : "flat, 16 B/thread",
               ms, bytes / ms / 1e9, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
