// K2: TransformedSpace.transform / .reverse as fused columnar kernels
// (Identity, Quantize, Reverse(Quantize), Enumerate, OneHotEncode in one pass
// over the batch).
//
// Same mapping as the sampler: a thread owns 2 consecutive candidates and walks
// the slots, so each warp-level access is 512 contiguous bytes of one column.
// The one-hot expansion is pure store traffic (k columns of 16 B per thread).
#include "ob2_internal.cuh"

namespace {

__device__ __forceinline__ long long quantize(double x, int t_flags) {
    // numpy.round (half to even) by default; floor when the descriptor asks
    const double r = (t_flags & OB2_TFLAG_FLOOR) ? floor(x) : rint(x);
    return __double2ll_rz(r);
}

__global__ void __launch_bounds__(OB2_BLOCK)
transform_kernel(const ob2_slot_desc *__restrict__ g_slots, int n_slots, int64_t n, int64_t n_stride,
                 const char *__restrict__ cols, char *__restrict__ tcols) {
    extern __shared__ __align__(16) unsigned char smem[];
    ob2_slot_desc *s_slots = reinterpret_cast<ob2_slot_desc *>(smem);
    ob2_stage_tables(g_slots, n_slots, nullptr, 0, s_slots, nullptr);
    const int64_t i0 = ((int64_t)blockIdx.x * OB2_BLOCK + threadIdx.x) * OB2_CPT;
    if (i0 >= n) return;
    for (int s = 0; s < n_slots; ++s) {
        const ob2_slot_desc &d = s_slots[s];
        const char *src = cols + (int64_t)d.col_units * n_stride;
        char *dst = tcols + (int64_t)d.t_units * n_stride;
        switch (d.t_op) {
            case OB2_OP_ONEHOT: {
                int c0, c1;
                ld_i32x2(reinterpret_cast<const int *>(src) + i0, c0, c1);
                if (d.t_width == 1) {  // k <= 2: a single float column
                    st_f64x2(reinterpret_cast<double *>(dst) + i0, (double)c0, (double)c1);
                } else {
                    // as in the fused sampler: only the high words of 0.0 / 1.0 are selected, and column j
                    // is addressed with one 32 x 32 + 64 multiply-add (launch_codec keeps 8 * n_stride < 2^32)
                    char *out = dst + i0 * 8;
                    const uint32_t s8 = (uint32_t)(8 * n_stride), k = (uint32_t)d.t_width;
#pragma unroll 4
                    for (uint32_t j = 0; j < k; ++j) {
                        const unsigned h0 = c0 == (int)j ? 0x3FF00000u : 0u, h1 = c1 == (int)j ? 0x3FF00000u : 0u;
                        st_onehot_hi(reinterpret_cast<double *>(out + (uint64_t)j * s8), h0, h1);
                    }
                }
                break;
            }
            case OB2_OP_QUANTIZE: {
                double x0, x1;
                ld_f64x2(reinterpret_cast<const double *>(src) + i0, x0, x1);
                st_i64x2(reinterpret_cast<long long *>(dst) + i0, quantize(x0, d.t_flags), quantize(x1, d.t_flags));
                break;
            }
            case OB2_OP_TO_REAL: {
                long long q0, q1;
                ld_i64x2(reinterpret_cast<const long long *>(src) + i0, q0, q1);
                st_f64x2(reinterpret_cast<double *>(dst) + i0, __ll2double_rn(q0), __ll2double_rn(q1));
                break;
            }
            default: {  // IDENTITY / ENUMERATE: copy with the slot's own width
                if (d.kind == OB2_KIND_CATEGORICAL) {
                    int c0, c1;
                    ld_i32x2(reinterpret_cast<const int *>(src) + i0, c0, c1);
                    st_i32x2(reinterpret_cast<int *>(dst) + i0, c0, c1);
                } else {
                    long long q0, q1;
                    ld_i64x2(reinterpret_cast<const long long *>(src) + i0, q0, q1);
                    st_i64x2(reinterpret_cast<long long *>(dst) + i0, q0, q1);
                }
                break;
            }
        }
    }
}

// WITH_FP: also fingerprint the rows it writes.  reverse is bandwidth bound with ~3/4 of its
// issue slots idle, so the fingerprint arithmetic (which costs the issue-bound sampler 17 %)
// rides along for free here.
template <bool WITH_FP>
__global__ void __launch_bounds__(OB2_BLOCK)
reverse_kernel(const ob2_slot_desc *__restrict__ g_slots, int n_slots, int64_t n, int64_t n_stride,
               const char *__restrict__ tcols, char *__restrict__ cols, uint64_t *__restrict__ fp) {
    extern __shared__ __align__(16) unsigned char smem[];
    ob2_slot_desc *s_slots = reinterpret_cast<ob2_slot_desc *>(smem);
    ob2_stage_tables(g_slots, n_slots, nullptr, 0, s_slots, nullptr);
    const int64_t i0 = ((int64_t)blockIdx.x * OB2_BLOCK + threadIdx.x) * OB2_CPT;
    if (i0 >= n) return;
    FpAcc acc0{0, 0}, acc1{0, 0};
    for (int s = 0; s < n_slots; ++s) {
        const ob2_slot_desc &d = s_slots[s];
        const char *src = tcols + (int64_t)d.t_units * n_stride;
        char *dst = cols + (int64_t)d.col_units * n_stride;
        uint64_t w0, w1;  // the values written, as canonical fingerprint words
        switch (d.t_op) {
            case OB2_OP_ONEHOT: {
                int c0 = 0, c1 = 0;
                const double *in = reinterpret_cast<const double *>(src) + i0;
                if (d.t_width == 1) {
                    if (d.n_cats == 2) {  // (v > 0.5).astype(int)
                        double v0, v1;
                        ld_f64x2(in, v0, v1);
                        c0 = v0 > 0.5 ? 1 : 0;
                        c1 = v1 > 0.5 ? 1 : 0;
                    }                     // k == 1: always category 0
                } else {                  // argmax over the trailing axis, first max wins
                    double b0, b1;
                    ld_f64x2(in, b0, b1);
                    // 4 loads in flight per step to cover HBM latency on wide encodings
                    int j = 1;
                    for (; j + 3 < d.t_width; j += 4) {
                        double v[8];
#pragma unroll
                        for (int q = 0; q < 4; ++q) ld_f64x2(in + (int64_t)(j + q) * n_stride, v[2 * q], v[2 * q + 1]);
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            if (v[2 * q] > b0) {
                                b0 = v[2 * q];
                                c0 = j + q;
                            }
                            if (v[2 * q + 1] > b1) {
                                b1 = v[2 * q + 1];
                                c1 = j + q;
                            }
                        }
                    }
                    for (; j < d.t_width; ++j) {
                        double v0, v1;
                        ld_f64x2(in + (int64_t)j * n_stride, v0, v1);
                        if (v0 > b0) {
                            b0 = v0;
                            c0 = j;
                        }
                        if (v1 > b1) {
                            b1 = v1;
                            c1 = j;
                        }
                    }
                }
                st_i32x2(reinterpret_cast<int *>(dst) + i0, c0, c1);
                w0 = (uint64_t)(int64_t)c0;
                w1 = (uint64_t)(int64_t)c1;
                break;
            }
            case OB2_OP_QUANTIZE: {  // Quantize.reverse: cast back to float
                long long q0, q1;
                ld_i64x2(reinterpret_cast<const long long *>(src) + i0, q0, q1);
                const double x0 = __ll2double_rn(q0), x1 = __ll2double_rn(q1);
                st_f64x2(reinterpret_cast<double *>(dst) + i0, x0, x1);
                w0 = canon_f64(x0);
                w1 = canon_f64(x1);
                break;
            }
            case OB2_OP_TO_REAL: {  // Reverse(Quantize).reverse = Quantize.transform
                double x0, x1;
                ld_f64x2(reinterpret_cast<const double *>(src) + i0, x0, x1);
                const long long q0 = quantize(x0, d.t_flags), q1 = quantize(x1, d.t_flags);
                st_i64x2(reinterpret_cast<long long *>(dst) + i0, q0, q1);
                w0 = (uint64_t)q0;
                w1 = (uint64_t)q1;
                break;
            }
            default: {
                if (d.kind == OB2_KIND_CATEGORICAL) {
                    int c0, c1;
                    ld_i32x2(reinterpret_cast<const int *>(src) + i0, c0, c1);
                    st_i32x2(reinterpret_cast<int *>(dst) + i0, c0, c1);
                    w0 = (uint64_t)(int64_t)c0;
                    w1 = (uint64_t)(int64_t)c1;
                } else {
                    long long q0, q1;
                    ld_i64x2(reinterpret_cast<const long long *>(src) + i0, q0, q1);
                    st_i64x2(reinterpret_cast<long long *>(dst) + i0, q0, q1);
                    if (d.kind == OB2_KIND_REAL) {
                        w0 = canon_f64(__longlong_as_double(q0));
                        w1 = canon_f64(__longlong_as_double(q1));
                    } else {
                        w0 = (uint64_t)q0;
                        w1 = (uint64_t)q1;
                    }
                }
                break;
            }
        }
        if (WITH_FP) {
            fp_add(acc0, d, w0);
            fp_add(acc1, d, w1);
        }
    }
    if (WITH_FP) {
        uint64_t lo0, hi0, lo1, hi1;
        fp_finish(acc0, n_slots, lo0, hi0);
        fp_finish(acc1, n_slots, lo1, hi1);
        if (i0 + 1 >= n) lo1 = hi1 = 0;  // padding entry
        st_u64x4(fp + 2 * i0, lo0, hi0, lo1, hi1);
    }
}

template <typename K, typename... Extra>
int launch_codec(K kernel, const char *what, size_t smem_floor, const ob2_space *sp, int64_t n, int64_t n_stride,
                 const void *in, void *out, void *stream, Extra... extra) {
    int rc = ob2_check_batch(sp, n, n_stride);
    if (rc != OB2_OK) return rc;
    if (in == nullptr || out == nullptr) {
        ob2_set_error("%s: null buffer", what);
        return OB2_EINVAL;
    }
    OB2_ALIGNED(in, 16, what);
    OB2_ALIGNED(out, 16, what);
    if (n == 0) return OB2_OK;
    if (n_stride >= (1ll << 29)) {  // the one-hot loop steps through columns with 32-bit byte strides
        ob2_set_error("%s: n_stride %lld too large for one launch (limit 2^29 - 1); split the batch", what,
                      (long long)n_stride);
        return OB2_ETOOBIG;
    }
    const int64_t threads = (n + OB2_CPT - 1) / OB2_CPT;
    const int64_t blocks = (threads + OB2_BLOCK - 1) / OB2_BLOCK;
    size_t smem = (size_t)sp->n_slots * sizeof(ob2_slot_desc);
    if (smem < smem_floor) smem = smem_floor;
    if (smem > 48 * 1024) {
        rc = ob2_check_cuda(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                            "cudaFuncSetAttribute(smem)");
        if (rc != OB2_OK) return rc;
    }
    kernel<<<(unsigned)blocks, OB2_BLOCK, smem, static_cast<cudaStream_t>(stream)>>>(
        sp->d_slots, sp->n_slots, n, n_stride, (const char *)in, (char *)out, extra...);
    return ob2_check_cuda(cudaGetLastError(), what);
}

}  // namespace

extern "C" int ob2_transform(const ob2_space *space, int64_t n, int64_t n_stride, const void *cols, void *tcols,
                             void *stream) {
    // The one-hot expansion is a pure burst of stores to t_width far-apart columns per slot; with the
    // 4 resident CTAs per SM that a 56 KB shared-memory request leaves (instead of 8) fewer column
    // streams are open at once and cat32x100 runs 5 % faster (4.26 -> 4.04 ms per 1M-point chunk).
    const size_t floor = space != nullptr && space->has_onehot ? 56 * 1024 : 0;
    return launch_codec(transform_kernel, "transform_kernel", floor, space, n, n_stride, cols, tcols, stream);
}

extern "C" int ob2_reverse(const ob2_space *space, int64_t n, int64_t n_stride, const void *tcols, void *cols,
                           void *stream) {
    return launch_codec(reverse_kernel<false>, "reverse_kernel", 0, space, n, n_stride, tcols, cols, stream,
                        static_cast<uint64_t *>(nullptr));
}

extern "C" int ob2_reverse_fingerprint(const ob2_space *space, int64_t n, int64_t n_stride, const void *tcols,
                                       void *cols, uint64_t *fp, void *stream) {
    if (fp == nullptr) {
        ob2_set_error("ob2_reverse_fingerprint: null fingerprint buffer");
        return OB2_EINVAL;
    }
    OB2_ALIGNED(fp, 16, "ob2_reverse_fingerprint");
    return launch_codec(reverse_kernel<true>, "reverse_kernel", 0, space, n, n_stride, tcols, cols, stream, fp);
}
