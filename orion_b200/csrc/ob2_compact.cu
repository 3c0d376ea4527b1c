// Order-preserving gather of the surviving candidates into a dense column
// block: the device half of what Producer.produce does after suggest()
// (src/orion/core/worker/producer.py:47-51 turns every point into a Trial and
// registers it; only deduplicated points should travel to the host).
//
// Three launches: per-block survivor counts -> exclusive scan of the counts
// (one block) -> gather.  A block covers 512 rows (256 threads x 2), reads are
// 128-bit and coalesced, writes are contiguous because ranks are.
#include "ob2_internal.cuh"

namespace {

constexpr int kRowsPerBlock = OB2_BLOCK * OB2_CPT;

__device__ __forceinline__ unsigned block_exclusive_rank(unsigned mine, unsigned *s_warp, unsigned *total) {
    // exclusive prefix of `mine` (0..2) over the block's threads
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned up = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= (unsigned)d) incl += up;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        unsigned w = lane < (OB2_BLOCK / 32) ? s_warp[lane] : 0;
        unsigned wi = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned up = __shfl_up_sync(0xFFFFFFFFu, wi, d);
            if (lane >= (unsigned)d) wi += up;
        }
        if (lane < (OB2_BLOCK / 32)) s_warp[lane] = wi - w;  // exclusive warp offsets
        if (lane == (OB2_BLOCK / 32) - 1) *total = wi;
    }
    __syncthreads();
    return s_warp[warp] + incl - mine;
}

__global__ void __launch_bounds__(OB2_BLOCK)
compact_count_kernel(int64_t n, const uint8_t *__restrict__ keep, uint32_t *__restrict__ block_counts) {
    const int64_t i0 = ((int64_t)blockIdx.x * OB2_BLOCK + threadIdx.x) * OB2_CPT;
    unsigned mine = 0;
    if (i0 < n) mine = (keep[i0] != 0) + ((i0 + 1 < n) && keep[i0 + 1] != 0);
    const unsigned ballot_sum = __reduce_add_sync(0xFFFFFFFFu, mine);
    __shared__ unsigned s_sum;
    if (threadIdx.x == 0) s_sum = 0;
    __syncthreads();
    if ((threadIdx.x & 31) == 0 && ballot_sum) atomicAdd(&s_sum, ballot_sum);
    __syncthreads();
    if (threadIdx.x == 0) block_counts[blockIdx.x] = s_sum;
}

// one block: block_counts[b] <- exclusive prefix; *n_kept <- total
__global__ void __launch_bounds__(1024)
compact_scan_kernel(int64_t n_blocks, uint32_t *__restrict__ block_counts, unsigned long long *__restrict__ n_kept) {
    __shared__ unsigned long long s_part[1024];
    const int t = threadIdx.x;
    const int64_t per = (n_blocks + 1023) / 1024;
    const int64_t lo = t * per, hi = min(lo + per, n_blocks);
    unsigned long long sum = 0;
    for (int64_t b = lo; b < hi; ++b) sum += block_counts[b];
    s_part[t] = sum;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {  // Hillis-Steele inclusive scan of the 1024 partial sums
        const unsigned long long v = t >= d ? s_part[t - d] : 0;
        __syncthreads();
        s_part[t] += v;
        __syncthreads();
    }
    unsigned long long run = s_part[t] - sum;
    for (int64_t b = lo; b < hi; ++b) {
        const unsigned c = block_counts[b];
        block_counts[b] = (uint32_t)run;
        run += c;
    }
    if (t == 1023) *n_kept = s_part[1023];
}

// WIRE: original columns only, each written at the slot's narrowed width (descriptor `reserved`
// = (offset units << 4) | element bytes, filled in by ob2_space_create).
template <bool WIRE>
__global__ void __launch_bounds__(OB2_BLOCK)
compact_gather_kernel(const ob2_slot_desc *__restrict__ g_slots, int n_slots, int transformed, int64_t n,
                      int64_t n_stride, const char *__restrict__ cols, const uint8_t *__restrict__ keep,
                      const uint32_t *__restrict__ block_offsets, int64_t out_stride, char *__restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem[];
    ob2_slot_desc *s_slots = reinterpret_cast<ob2_slot_desc *>(smem);
    __shared__ unsigned s_warp[OB2_BLOCK / 32];
    __shared__ unsigned s_total;
    ob2_stage_tables(g_slots, n_slots, nullptr, 0, s_slots, nullptr);
    const int64_t i0 = ((int64_t)blockIdx.x * OB2_BLOCK + threadIdx.x) * OB2_CPT;
    const bool k0 = i0 < n && keep[i0] != 0;
    const bool k1 = i0 + 1 < n && keep[i0 + 1] != 0;
    const unsigned mine = (unsigned)k0 + (unsigned)k1;
    const int64_t r0 = (int64_t)block_offsets[blockIdx.x] + block_exclusive_rank(mine, s_warp, &s_total);
    const int64_t r1 = r0 + (k0 ? 1 : 0);
    if (mine == 0) return;
    for (int s = 0; s < n_slots; ++s) {
        const ob2_slot_desc &d = s_slots[s];
        if (WIRE) {
            const char *src = cols + (int64_t)d.col_units * n_stride;
            char *dst = out + (int64_t)(d.reserved >> 4) * out_stride;
            long long a, b;
            if (d.kind == OB2_KIND_CATEGORICAL) {
                int a4, b4;
                ld_i32x2(reinterpret_cast<const int *>(src) + i0, a4, b4);
                a = a4;
                b = b4;
            } else {
                ld_i64x2(reinterpret_cast<const long long *>(src) + i0, a, b);
            }
            switch (d.reserved & 15) {  // truncation keeps the value: it lies in the slot's interval
                case 1:
                    if (k0) reinterpret_cast<signed char *>(dst)[r0] = (signed char)a;
                    if (k1) reinterpret_cast<signed char *>(dst)[r1] = (signed char)b;
                    break;
                case 2:
                    if (k0) reinterpret_cast<short *>(dst)[r0] = (short)a;
                    if (k1) reinterpret_cast<short *>(dst)[r1] = (short)b;
                    break;
                case 4:
                    if (k0) reinterpret_cast<int *>(dst)[r0] = (int)a;
                    if (k1) reinterpret_cast<int *>(dst)[r1] = (int)b;
                    break;
                default:
                    if (k0) reinterpret_cast<long long *>(dst)[r0] = a;
                    if (k1) reinterpret_cast<long long *>(dst)[r1] = b;
                    break;
            }
        } else if (!transformed) {
            const char *src = cols + (int64_t)d.col_units * n_stride;
            char *dst = out + (int64_t)d.col_units * out_stride;
            if (d.kind == OB2_KIND_CATEGORICAL) {
                int a, b;
                ld_i32x2(reinterpret_cast<const int *>(src) + i0, a, b);
                if (k0) reinterpret_cast<int *>(dst)[r0] = a;
                if (k1) reinterpret_cast<int *>(dst)[r1] = b;
            } else {
                long long a, b;
                ld_i64x2(reinterpret_cast<const long long *>(src) + i0, a, b);
                if (k0) reinterpret_cast<long long *>(dst)[r0] = a;
                if (k1) reinterpret_cast<long long *>(dst)[r1] = b;
            }
        } else {
            const bool narrow = d.kind == OB2_KIND_CATEGORICAL && d.t_op != OB2_OP_ONEHOT;  // i32 column
            const char *src = cols + (int64_t)d.t_units * n_stride;
            char *dst = out + (int64_t)d.t_units * out_stride;
            for (int j = 0; j < d.t_width; ++j) {
                if (narrow) {
                    int a, b;
                    ld_i32x2(reinterpret_cast<const int *>(src) + i0, a, b);
                    if (k0) reinterpret_cast<int *>(dst)[r0] = a;
                    if (k1) reinterpret_cast<int *>(dst)[r1] = b;
                } else {
                    long long a, b;
                    ld_i64x2(reinterpret_cast<const long long *>(src) + i0, a, b);
                    if (k0) reinterpret_cast<long long *>(dst)[r0] = a;
                    if (k1) reinterpret_cast<long long *>(dst)[r1] = b;
                }
                src += 8 * n_stride;
                dst += 8 * out_stride;
            }
        }
    }
}

}  // namespace

extern "C" int64_t ob2_compact_scratch_bytes(int64_t n) {
    if (n < 0) return -1;
    const int64_t blocks = (n + kRowsPerBlock - 1) / kRowsPerBlock;
    return (blocks + 1) * (int64_t)sizeof(uint32_t);
}

namespace {
int compact_impl(const ob2_space *space, int64_t n, int64_t n_stride, const void *cols, const uint8_t *keep,
                 int32_t transformed, bool wire, int64_t out_stride, void *out_cols, void *scratch, uint64_t *n_kept,
                 void *stream) {
    int rc = ob2_check_batch(space, n, n_stride);
    if (rc != OB2_OK) return rc;
    if (cols == nullptr || keep == nullptr || out_cols == nullptr || scratch == nullptr || n_kept == nullptr ||
        out_stride < 0 || (out_stride & 31) != 0) {
        ob2_set_error("ob2_compact: bad arguments (out_stride must be a multiple of 32)");
        return OB2_EINVAL;
    }
    OB2_ALIGNED(cols, 16, "ob2_compact");
    OB2_ALIGNED(out_cols, 16, "ob2_compact");
    OB2_ALIGNED(scratch, 4, "ob2_compact");
    OB2_ALIGNED(n_kept, 8, "ob2_compact");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (n == 0) return ob2_check_cuda(cudaMemsetAsync(n_kept, 0, sizeof(uint64_t), st), "cudaMemsetAsync(n_kept)");
    const int64_t blocks = (n + kRowsPerBlock - 1) / kRowsPerBlock;
    if (blocks > 0x7FFFFFFFll) {
        ob2_set_error("batch too large for one launch");
        return OB2_ETOOBIG;
    }
    uint32_t *counts = static_cast<uint32_t *>(scratch);
    compact_count_kernel<<<(unsigned)blocks, OB2_BLOCK, 0, st>>>(n, keep, counts);
    rc = ob2_check_cuda(cudaGetLastError(), "compact_count_kernel launch");
    if (rc != OB2_OK) return rc;
    compact_scan_kernel<<<1, 1024, 0, st>>>(blocks, counts, reinterpret_cast<unsigned long long *>(n_kept));
    rc = ob2_check_cuda(cudaGetLastError(), "compact_scan_kernel launch");
    if (rc != OB2_OK) return rc;
    const size_t smem = (size_t)space->n_slots * sizeof(ob2_slot_desc);
    auto kern = wire ? compact_gather_kernel<true> : compact_gather_kernel<false>;
    if (smem > 48 * 1024) {
        rc = ob2_check_cuda(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                            "cudaFuncSetAttribute(smem)");
        if (rc != OB2_OK) return rc;
    }
    kern<<<(unsigned)blocks, OB2_BLOCK, smem, st>>>(space->d_slots, space->n_slots, transformed, n, n_stride,
                                                    (const char *)cols, keep, counts, out_stride, (char *)out_cols);
    return ob2_check_cuda(cudaGetLastError(), "compact_gather_kernel launch");
}
}  // namespace

extern "C" int ob2_compact(const ob2_space *space, int64_t n, int64_t n_stride, const void *cols,
                           const uint8_t *keep, int32_t transformed, int64_t out_stride, void *out_cols,
                           void *scratch, uint64_t *n_kept, void *stream) {
    return compact_impl(space, n, n_stride, cols, keep, transformed, false, out_stride, out_cols, scratch, n_kept, stream);
}

extern "C" int ob2_compact_wire(const ob2_space *space, int64_t n, int64_t n_stride, const void *cols,
                                const uint8_t *keep, int64_t out_stride, void *out_cols, void *scratch,
                                uint64_t *n_kept, void *stream) {
    return compact_impl(space, n, n_stride, cols, keep, 0, true, out_stride, out_cols, scratch, n_kept, stream);
}

// The host half of the hand-over: copy the first n_rows of every column of a device column block
// into a HOST block of the same layout (stride dst_stride), asynchronously on `stream`.  Adjacent
// columns of one element size form a 2-D array (pitch = element size x stride), so a whole run of
// them is ONE cudaMemcpy2DAsync: 2 copies for mixed64 instead of one per column.
extern "C" int ob2_columns_to_host(const ob2_space *space, int32_t layout, int64_t n_rows, int64_t src_stride,
                                   const void *src_cols, int64_t dst_stride, void *dst_cols, void *stream) {
    if (space == nullptr || src_cols == nullptr || dst_cols == nullptr || n_rows < 0 || n_rows > src_stride ||
        n_rows > dst_stride || layout < 0 || layout > 2) {
        ob2_set_error("ob2_columns_to_host: bad arguments (layout 0 = original, 1 = transformed, 2 = wire)");
        return OB2_EINVAL;
    }
    if (n_rows == 0) return OB2_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const ob2_slot_desc *slots = space->h_slots;
    // walk the columns of the chosen layout: (offset units, element bytes, number of columns of the slot)
    int64_t run_units = -1, run_elem = 0, run_cols = 0;
    auto flush = [&]() -> int {
        if (run_cols == 0) return OB2_OK;
        const char *src = static_cast<const char *>(src_cols) + run_units * src_stride;
        char *dst = static_cast<char *>(dst_cols) + run_units * dst_stride;
        return ob2_check_cuda(cudaMemcpy2DAsync(dst, (size_t)(run_elem * dst_stride), src, (size_t)(run_elem * src_stride),
                                                (size_t)(run_elem * n_rows), (size_t)run_cols, cudaMemcpyDeviceToHost, st),
                              "cudaMemcpy2DAsync(columns to host)");
    };
    for (int s = 0; s < space->n_slots; ++s) {
        const ob2_slot_desc &d = slots[s];
        int64_t units, elem, cols = 1;
        if (layout == 2) {
            units = space->wire_units[s];
            elem = space->wire_bytes[s];
        } else if (layout == 1) {
            units = d.t_units;
            cols = d.t_width;
            elem = (d.t_op == OB2_OP_ENUMERATE || (d.t_op == OB2_OP_IDENTITY && d.kind == OB2_KIND_CATEGORICAL)) ? 4 : 8;
        } else {
            units = d.col_units;
            elem = d.kind == OB2_KIND_CATEGORICAL ? 4 : 8;
        }
        if (run_cols > 0 && elem == run_elem && units == run_units + run_elem * run_cols) {
            run_cols += cols;
            continue;
        }
        const int rc = flush();
        if (rc != OB2_OK) return rc;
        run_units = units;
        run_elem = elem;
        run_cols = cols;
    }
    return flush();
}
