// K1: candidate sampling (Philox4x32-10 or injected uniforms -> per-slot
// inverse CDF), fused with the bounds check, the <=4-attempt rejection, the row
// fingerprint and -- optionally -- the transform chain and its reverse (the
// one-kernel ob2_suggest); the standalone Space.__contains__ kernel; and the
// *_host entry points for the small calls the reference's callers make.
//
// Mapping: one thread owns OB2_CPT = 2 consecutive candidates and walks the
// slots in space order.  At any instant a warp therefore works on ONE slot
// (no divergence on the prior / kind switch) and stores 32 x 16 B = 512
// contiguous bytes of that slot's column.  Output is columnar [S][n_stride].
//
// On its own the sampler is instruction-issue bound (Philox + fp64 inverse CDFs
// + fingerprint multiplies, ~8-10k instructions per candidate of mixed64) while
// transform and reverse are pure memory traffic; fusing them in (WITH_T, WITH_B)
// lets the stores ride under the arithmetic and removes every re-read: the fused
// kernel ends up at the HBM write ceiling (profiles/r1g_full_suggest.md).
#include <cstring>

#include "ob2_internal.cuh"

#ifndef OB2_PARAM_SLOTS
#define OB2_PARAM_SLOTS 192  // 24 KB of the 32 KB a launch may carry in its parameters
#endif
#ifndef OB2_SUGGEST_CTAS
#define OB2_SUGGEST_CTAS 4  // resident CTAs per SM the fused suggest kernel is compiled for
#endif

namespace {

__device__ __forceinline__ long long quantize_ll(double x, int t_flags) {
    return __double2ll_rz((t_flags & OB2_TFLAG_FLOOR) ? floor(x) : rint(x));
}

// Categorical draw = #{j : cdf[j] <= u}.  With equal probabilities floor(u * k)
// is within one of the answer, so start there and walk to the exact index
// (exact for ANY non-decreasing cdf; the guess only decides how far the walk is).
// Otherwise the branch-free descent.
__device__ __forceinline__ int draw_cat_fast(const double *cdf, int k, double u, bool uniform) {
    if (!uniform) return draw_cat(cdf, k, u);
    int idx = __double2int_rz(u * (double)k);
    idx = min(max(idx, 0), k);
    while (idx > 0 && cdf[idx - 1] > u) --idx;
    while (idx < k && cdf[idx] <= u) ++idx;
    return idx;
}

// Real.sample (space.py:263-270): a draw of dimension [first, s_last] landed
// outside [low, high): redraw the whole (possibly shaped) value for candidate i,
// attempts 1..3, overwriting its columns with scalar stores.  Rare path, inlined
// in a cold branch of the walk: the hot loop carries no redraw state at all --
// the fingerprint lanes are wrapping sums, so the contribution of the rejected
// values (read back from the columns this thread just wrote) is subtracted and
// the redrawn one added.
// WITH_B: the fingerprint is over the handed-out row (`back`, at cols + bdelta), which is also rewritten.
template <bool WITH_FP, bool WITH_T, bool WITH_B, bool EXT>
__device__ __forceinline__ bool redraw_dim(const ob2_slot_desc *s_slots, int s_last, uint64_t index, int64_t i,
                                           int64_t n_stride, char *cols, char *tcols, int64_t bdelta, uint32_t k0,
                                           uint32_t k1, uint32_t pm0, uint32_t pm1, FpAcc &acc) {
    const int first = s_slots[s_last].dim_first;
    if (WITH_FP) {
        FpAcc old{0, 0};
        for (int e = first; e <= s_last; ++e) {
            const ob2_slot_desc &d = s_slots[e];
            const char *col = cols + (int64_t)d.col_units * n_stride + (WITH_B ? bdelta : 0);
            const uint64_t w = d.kind == OB2_KIND_INTEGER
                                   ? (uint64_t) reinterpret_cast<const volatile long long *>(col)[i]
                                   : canon_f64(reinterpret_cast<const volatile double *>(col)[i]);
            fp_add(old, d, w);
        }
        acc.a -= old.a;
        acc.b -= old.b;
    }
    bool ok = false;
    FpAcc t{0, 0};
    for (uint32_t a = 1; a < 4 && !ok; ++a) {
        ok = true;
        t = FpAcc{0, 0};
        for (int e = first; e <= s_last; ++e) {
            const ob2_slot_desc &d = s_slots[e];
            const Philox4 r = philox4x32_10((uint32_t)index, (uint32_t)(index >> 32), (uint32_t)(e >> 1), a, k0, k1, pm0, pm1);
            const double u = (e & 1) ? u53(r.z, r.w) : u53(r.x, r.y);
            double x = draw_real<EXT>(d, u);
            char *col = cols + (int64_t)d.col_units * n_stride;
            char *dst = tcols + (int64_t)d.t_units * n_stride;
            uint64_t w;
            if (d.kind == OB2_KIND_INTEGER) {
                x = floor(x);
                const long long q = __double2ll_rz(x);
                reinterpret_cast<long long *>(col)[i] = q;
                long long bq = q;
                if (WITH_T) {
                    if (d.t_op == OB2_OP_TO_REAL) {
                        reinterpret_cast<double *>(dst)[i] = x;
                        if (WITH_B) bq = quantize_ll(x, d.t_flags);
                    } else {
                        reinterpret_cast<long long *>(dst)[i] = q;
                    }
                }
                if (WITH_B && bdelta != 0) reinterpret_cast<long long *>(col + bdelta)[i] = bq;
                w = (uint64_t)bq;
            } else {
                reinterpret_cast<double *>(col)[i] = x;
                double bx = x;
                if (WITH_T) {
                    if (d.t_op == OB2_OP_QUANTIZE) {
                        const long long t = quantize_ll(x, d.t_flags);
                        reinterpret_cast<long long *>(dst)[i] = t;
                        bx = __ll2double_rn(t);
                    } else {
                        reinterpret_cast<double *>(dst)[i] = x;
                    }
                }
                if (WITH_B && (bdelta != 0 || d.t_op == OB2_OP_QUANTIZE)) reinterpret_cast<double *>(col + bdelta)[i] = bx;
                w = canon_f64(WITH_B ? bx : x);
            }
            ok &= (x >= d.low) && (x < d.high);
            if (WITH_FP) fp_add(t, d, w);
        }
    }
    if (WITH_FP) {
        acc.a += t.a;
        acc.b += t.b;
    }
    return ok;
}

// pm0 / pm1 are the Philox multipliers, passed as kernel parameters so that the
// multiplies read them straight from the constant bank.
//
// Hot-loop economy (the kernel is issue bound): column pointers are running
// pointers advanced by 4*n_stride / 8*n_stride instead of being recomputed from
// the descriptor; the fingerprint accumulates straight into its lanes; the
// second candidate of a trailing odd pair replays the first (no `has1` tests
// until the final stores).
//
// WITH_B (Random.suggest behind PrimaryAlgo, in one pass): also evaluate reverse() on the
// transformed values while they are still in registers and store the handed-out row to
// `back` = cols + bdelta (same column layout); the fingerprint is then that of the `back`
// row, exactly what reverse_kernel<true> computes from the materialised tcols.  Per op:
// Quantize of the stored double, the cast back of the stored integer and (v > 0.5) for k = 2
// are evaluated on the transformed value; Identity / Enumerate pass the value through; the
// argmax over a k > 2 one-hot block is the position of its single 1.0, the drawn index.
//
// sample_body is the whole per-thread job.  `s_slots` is either the table staged in shared
// memory (sample_kernel) or the table passed BY VALUE as a kernel parameter (sample_kernel_p):
// the parameter lives in the constant bank, so every descriptor field is read with a uniform
// load into a uniform register and the kind / prior / op branches are known to be warp-uniform
// (no convergence barriers around them, no per-thread LDS of the descriptors).
template <bool PHILOX, bool WITH_FP, bool WITH_T, bool WITH_B, bool EXT>
__device__ __forceinline__ void sample_body(const ob2_slot_desc *s_slots, const double *s_cdf, int n_slots, int64_t n,
                                            int64_t n_stride, int64_t stride4, int64_t stride8, uint32_t k0,
                                            uint32_t k1, uint32_t pm0, uint32_t pm1, uint64_t base_index,
                                            const uint64_t *__restrict__ index, const double *__restrict__ U,
                                            char *__restrict__ cols, char *__restrict__ tcols, int64_t bdelta,
                                            uint8_t *__restrict__ mask, uint64_t *__restrict__ fp,
                                            uint32_t *__restrict__ n_failed) {
    const int64_t i0 = ((int64_t)blockIdx.x * OB2_BLOCK + threadIdx.x) * OB2_CPT;
    if (i0 >= n) return;
    const bool has1 = i0 + 1 < n;
    uint64_t idx0 = base_index + (uint64_t)i0, idx1 = idx0 + 1;
    if (PHILOX && index != nullptr) {  // explicit counters (replay / planted duplicates)
        idx0 = __ldg(index + i0);
        idx1 = has1 ? __ldg(index + i0 + 1) : idx0;
    }
    if (!has1) idx1 = idx0;  // the padding candidate replays its neighbour: always well defined

    // back == cols (bdelta == 0): the caller keeps only the handed-out rows; the second store of every
    // value is skipped (reverse(transform(x)) == x bit for bit except through a lossy Quantize)
    const bool split = WITH_B && bdelta != 0;
    FpAcc acc0{0, 0}, acc1{0, 0};
    bool ok0 = true, ok1 = true;    // current dimension
    bool all0 = true, all1 = true;  // whole candidate
    Philox4 r0{0, 0, 0, 0}, r1{0, 0, 0, 0};
    // stride4 / stride8 = 4 * n_stride / 8 * n_stride arrive as kernel parameters: the pointer bumps then
    // take them straight from the constant bank instead of re-deriving them every slot
    char *col = cols + i0 * 8;                          // running pointers: element i0 of the current column
    char *tcol = WITH_T ? tcols + i0 * 8 : nullptr;     // (a 4-byte column is addressed at half the offset)
    const double *ucol = PHILOX ? nullptr : U + i0;

    for (int s = 0; s < n_slots; ++s) {
        const ob2_slot_desc &d = s_slots[s];
        double u0, u1;
        if (PHILOX) {
            if ((s & 1) == 0) {
                r0 = philox4x32_10((uint32_t)idx0, (uint32_t)(idx0 >> 32), (uint32_t)(s >> 1), 0u, k0, k1, pm0, pm1);
                r1 = philox4x32_10((uint32_t)idx1, (uint32_t)(idx1 >> 32), (uint32_t)(s >> 1), 0u, k0, k1, pm0, pm1);
                u0 = u53(r0.x, r0.y);
                u1 = u53(r1.x, r1.y);
            } else {
                u0 = u53(r0.z, r0.w);
                u1 = u53(r1.z, r1.w);
            }
        } else {
            ld_f64x2(ucol, u0, u1);
            ucol += n_stride;
        }
        uint64_t w0, w1;
        if (d.kind == OB2_KIND_CATEGORICAL) {
            const double *cdf = s_cdf + d.cdf_off;
            const bool uni = (d.flags & OB2_SLOT_CAT_UNIFORM) != 0;
            const int c0 = draw_cat_fast(cdf, d.n_cats, u0, uni);
            const int c1 = draw_cat_fast(cdf, d.n_cats, u1, uni);
            st_i32x2(reinterpret_cast<int *>(col - i0 * 4), c0, c1);
            int *bcol = reinterpret_cast<int *>(col - i0 * 4 + bdelta);
            col += stride4;
            int b0 = c0, b1 = c1;  // WITH_B: reverse() of the transformed value, evaluated on what is stored
            if (WITH_T) {
                if (d.t_op != OB2_OP_ONEHOT) {  // Enumerate / Identity: the index both ways
                    st_i32x2(reinterpret_cast<int *>(tcol - i0 * 4), c0, c1);
                    tcol += stride4;
                } else if (d.t_width == 1) {
                    const double t0 = (double)c0, t1 = (double)c1;
                    st_f64x2(reinterpret_cast<double *>(tcol), t0, t1);
                    tcol += stride8;
                    if (WITH_B) {  // k == 2: (v > 0.5); k == 1: always category 0
                        b0 = (d.n_cats == 2 && t0 > 0.5) ? 1 : 0;
                        b1 = (d.n_cats == 2 && t1 > 0.5) ? 1 : 0;
                    }
                } else {
                    // column j is at tcol + j * stride8; the launcher guarantees stride8 < 2^32,
                    // so each address is ONE 32x32+64 multiply-add (IMAD.WIDE.U32)
                    const uint32_t s8 = (uint32_t)stride8, k = (uint32_t)d.t_width;
                    // WITH_B: reverse_kernel's argmax (first maximum) over these k values is the position
                    // of the single 1.0 stored below, i.e. the index itself: the Philox draw is always in
                    // [0, k).  Tracking the maximum while storing (tried: 6 more instructions per column
                    // pair) costs the kernel 4.5 % for the same bytes, so the identity is used; the staged
                    // kernels and tests/test_gpu_suggest_fused.py hold it to account.
#pragma unroll 4
                    for (uint32_t j = 0; j < k; ++j) {
                        const unsigned h0 = c0 == (int)j ? 0x3FF00000u : 0u, h1 = c1 == (int)j ? 0x3FF00000u : 0u;
                        st_onehot_hi(reinterpret_cast<double *>(tcol + (uint64_t)j * s8), h0, h1);
                    }
                    tcol += (uint64_t)k * s8;
                }
            }
            if (WITH_B && split) st_i32x2(bcol, b0, b1);
            w0 = (uint64_t)(int64_t)b0;
            w1 = (uint64_t)(int64_t)b1;
            if (!PHILOX) {  // an injected u >= 1 can run off the table; a Philox u cannot
                ok0 &= c0 < d.n_cats;
                ok1 &= c1 < d.n_cats;
            }
        } else {
            double x0 = draw_real<EXT>(d, u0);
            double x1 = draw_real<EXT>(d, u1);
            if (d.kind == OB2_KIND_INTEGER) {
                x0 = floor(x0);  // _Discrete.sample, space.py:294
                x1 = floor(x1);
                const long long q0 = __double2ll_rz(x0), q1 = __double2ll_rz(x1);
                st_i64x2(reinterpret_cast<long long *>(col), q0, q1);
                long long bq0 = q0, bq1 = q1;
                if (WITH_T) {
                    if (d.t_op == OB2_OP_TO_REAL) {  // the floored double is exactly (double)q
                        st_f64x2(reinterpret_cast<double *>(tcol), x0, x1);
                        if (WITH_B) {  // Reverse(Quantize).reverse = Quantize.transform of the stored double
                            bq0 = quantize_ll(x0, d.t_flags);
                            bq1 = quantize_ll(x1, d.t_flags);
                        }
                    } else {
                        st_i64x2(reinterpret_cast<long long *>(tcol), q0, q1);
                    }
                    tcol += stride8;
                }
                if (WITH_B && split) st_i64x2(reinterpret_cast<long long *>(col + bdelta), bq0, bq1);
                w0 = (uint64_t)bq0;
                w1 = (uint64_t)bq1;
            } else {
                st_f64x2(reinterpret_cast<double *>(col), x0, x1);
                double bx0 = x0, bx1 = x1;
                if (WITH_T) {
                    if (d.t_op == OB2_OP_QUANTIZE) {
                        const long long t0 = quantize_ll(x0, d.t_flags), t1 = quantize_ll(x1, d.t_flags);
                        st_i64x2(reinterpret_cast<long long *>(tcol), t0, t1);
                        if (WITH_B) {
                            bx0 = __ll2double_rn(t0);
                            bx1 = __ll2double_rn(t1);
                        }
                    } else {
                        st_f64x2(reinterpret_cast<double *>(tcol), x0, x1);
                    }
                    tcol += stride8;
                }
                // (aliased: only a lossy Quantize round trip differs from what was just stored)
                if (WITH_B && (split || d.t_op == OB2_OP_QUANTIZE))
                    st_f64x2(reinterpret_cast<double *>(col + bdelta), bx0, bx1);
                w0 = canon_f64(bx0);
                w1 = canon_f64(bx1);
            }
            col += stride8;
            ok0 &= (x0 >= d.low) & (x0 < d.high);  // half-open, space.py:164
            ok1 &= (x1 >= d.low) & (x1 < d.high);
        }
        if (WITH_FP) {
            fp_add(acc0, d, w0);
            fp_add(acc1, d, w1);
        }
        if (d.flags & OB2_SLOT_DIM_LAST) {
            if (!(ok0 & ok1)) {  // cold: only Real/Integer slots can clear ok
                if (PHILOX) {
                    if (!ok0)
                        ok0 = redraw_dim<WITH_FP, WITH_T, WITH_B, EXT>(s_slots, s, idx0, i0, n_stride, cols, tcols, bdelta, k0, k1, pm0, pm1, acc0);
                    if (!ok1)
                        ok1 = redraw_dim<WITH_FP, WITH_T, WITH_B, EXT>(s_slots, s, idx1, i0 + 1, n_stride, cols, tcols, bdelta, k0, k1, pm0, pm1, acc1);
                    if (n_failed != nullptr) {  // 4 misses: the reference raises "Improbable bounds"
                        const unsigned bad = (ok0 ? 0u : 1u) + ((ok1 || !has1) ? 0u : 1u);
                        if (bad) atomicAdd(n_failed + d.dim_first, bad);
                    }
                }
                all0 &= ok0;
                all1 &= ok1;
                ok0 = ok1 = true;
            }
        }
    }
    if (mask != nullptr) {
        *reinterpret_cast<uchar2 *>(mask + i0) = make_uchar2(all0 ? 1 : 0, (all1 && has1) ? 1 : 0);
    }
    if (WITH_FP) {
        uint64_t lo0, hi0, lo1, hi1;
        fp_finish(acc0, n_slots, lo0, hi0);
        fp_finish(acc1, n_slots, lo1, hi1);
        if (!has1) lo1 = hi1 = 0;  // padding entry
        st_u64x4(fp + 2 * i0, lo0, hi0, lo1, hi1);
    }
}

template <bool PHILOX, bool WITH_FP, bool WITH_T, bool WITH_B, bool EXT>
__global__ void __launch_bounds__(OB2_BLOCK, WITH_B ? OB2_SUGGEST_CTAS : 4)
sample_kernel(const ob2_slot_desc *__restrict__ g_slots, int n_slots, const double *__restrict__ g_cdf,
              int cdf_len, int64_t n, int64_t n_stride, int64_t stride4, int64_t stride8, uint32_t k0, uint32_t k1,
              uint32_t pm0, uint32_t pm1,
              uint64_t base_index, const uint64_t *__restrict__ index, const double *__restrict__ U,
              char *__restrict__ cols, char *__restrict__ tcols, int64_t bdelta, uint8_t *__restrict__ mask,
              uint64_t *__restrict__ fp, uint32_t *__restrict__ n_failed) {
    extern __shared__ __align__(16) unsigned char smem[];
    ob2_slot_desc *s_slots = reinterpret_cast<ob2_slot_desc *>(smem);
    double *s_cdf = reinterpret_cast<double *>(smem + (size_t)n_slots * sizeof(ob2_slot_desc));
    ob2_stage_tables(g_slots, n_slots, g_cdf, cdf_len, s_slots, s_cdf);
    sample_body<PHILOX, WITH_FP, WITH_T, WITH_B, EXT>(s_slots, s_cdf, n_slots, n, n_stride, stride4, stride8, k0, k1, pm0,
                                                      pm1, base_index, index, U, cols, tcols, bdelta, mask, fp, n_failed);
}

// The same kernel for spaces of up to OB2_PARAM_SLOTS slots: the descriptor table travels in the
// launch parameters (constant bank); only the categorical cdf tables are staged in shared memory.
struct ob2_slot_table {
    ob2_slot_desc s[OB2_PARAM_SLOTS];
};

template <bool PHILOX, bool WITH_FP, bool WITH_T, bool WITH_B, bool EXT>
__global__ void __launch_bounds__(OB2_BLOCK, WITH_B ? OB2_SUGGEST_CTAS : 4)
sample_kernel_p(const __grid_constant__ ob2_slot_table tbl, int n_slots, const double *__restrict__ g_cdf, int cdf_len,
                int64_t n, int64_t n_stride, int64_t stride4, int64_t stride8, uint32_t k0, uint32_t k1, uint32_t pm0,
                uint32_t pm1, uint64_t base_index, const uint64_t *__restrict__ index,
                const double *__restrict__ U, char *__restrict__ cols, char *__restrict__ tcols, int64_t bdelta,
                uint8_t *__restrict__ mask, uint64_t *__restrict__ fp, uint32_t *__restrict__ n_failed) {
    extern __shared__ __align__(16) unsigned char smem[];
    double *s_cdf = reinterpret_cast<double *>(smem);
    for (int i = threadIdx.x; i < cdf_len; i += blockDim.x) s_cdf[i] = g_cdf[i];
    __syncthreads();
    sample_body<PHILOX, WITH_FP, WITH_T, WITH_B, EXT>(tbl.s, s_cdf, n_slots, n, n_stride, stride4, stride8, k0, k1, pm0,
                                                      pm1, base_index, index, U, cols, tcols, bdelta, mask, fp, n_failed);
}

// Space.__contains__ over columns.
__global__ void __launch_bounds__(OB2_BLOCK)
contains_kernel(const ob2_slot_desc *__restrict__ g_slots, int n_slots, int64_t n, int64_t n_stride,
                const char *__restrict__ cols, int int_as_f64, uint8_t *__restrict__ mask) {
    extern __shared__ __align__(16) unsigned char smem[];
    ob2_slot_desc *s_slots = reinterpret_cast<ob2_slot_desc *>(smem);
    ob2_stage_tables(g_slots, n_slots, nullptr, 0, s_slots, nullptr);
    const int64_t i0 = ((int64_t)blockIdx.x * OB2_BLOCK + threadIdx.x) * OB2_CPT;
    if (i0 >= n) return;
    const bool has1 = i0 + 1 < n;
    bool ok0 = true, ok1 = true;
    for (int s = 0; s < n_slots; ++s) {
        const ob2_slot_desc &d = s_slots[s];
        // with int_as_f64 the Integer columns are 8-byte doubles: same width, same offsets
        const char *col = cols + (int64_t)d.col_units * n_stride;
        if (d.kind == OB2_KIND_CATEGORICAL) {
            int c0, c1;
            ld_i32x2(reinterpret_cast<const int *>(col) + i0, c0, c1);
            ok0 &= (c0 >= 0) && (c0 < d.n_cats);
            ok1 &= (c1 >= 0) && (c1 < d.n_cats);
        } else if (d.kind == OB2_KIND_INTEGER && !int_as_f64) {
            long long q0, q1;
            ld_i64x2(reinterpret_cast<const long long *>(col) + i0, q0, q1);
            const double x0 = __ll2double_rn(q0), x1 = __ll2double_rn(q1);
            ok0 &= (x0 >= d.low) && (x0 < d.high);
            ok1 &= (x1 >= d.low) && (x1 < d.high);
        } else {
            double x0, x1;
            ld_f64x2(reinterpret_cast<const double *>(col) + i0, x0, x1);
            if (d.kind == OB2_KIND_INTEGER) {  // numpy.mod(x, 1) == 0, space.py:344
                ok0 &= (x0 == floor(x0)) && !isinf(x0);
                ok1 &= (x1 == floor(x1)) && !isinf(x1);
            }
            ok0 &= (x0 >= d.low) && (x0 < d.high);
            ok1 &= (x1 >= d.low) && (x1 < d.high);
        }
    }
    *reinterpret_cast<uchar2 *>(mask + i0) = make_uchar2(ok0 ? 1 : 0, (ok1 && has1) ? 1 : 0);
}

template <typename K>
int prepare_smem(K kernel, size_t smem) {
    if (smem > 48 * 1024) {
        return ob2_check_cuda(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                              "cudaFuncSetAttribute(smem)");
    }
    return OB2_OK;
}

struct SampleArgs {
    const ob2_space *sp;
    int64_t n, n_stride;
    uint64_t seed, base_index;
    const uint64_t *index;
    const double *U;
    void *cols, *tcols, *back;
    uint8_t *mask;
    uint64_t *fp;
    uint32_t *n_failed;
    void *stream;
};

template <bool PHILOX, bool WITH_FP, bool WITH_T, bool EXT, bool WITH_B = false>
int launch_one(const SampleArgs &a, unsigned blocks) {
    auto kern = sample_kernel<PHILOX, WITH_FP, WITH_T, WITH_B, EXT>;
    const int64_t bdelta = WITH_B ? (const char *)a.back - (const char *)a.cols : 0;
    const ob2_space *sp = a.sp;
    // Descriptors in the launch parameters: measured -1.5 % on the fused kernels (suggest 4.39 -> 4.33 ms,
    // sample+transform 3.46 -> 3.41 ms) but +5 % on the issue-bound plain sampler (real16: 4.99 -> 5.23 ms
    // per 1e8), so only the store-heavy variants use it.
    if ((WITH_B || WITH_T) && sp->n_slots <= OB2_PARAM_SLOTS && sp->h_slots != nullptr) {
        auto kern_p = sample_kernel_p<PHILOX, WITH_FP, WITH_T, WITH_B, EXT>;
        static thread_local ob2_slot_table tbl;  // (24 KB: not on the stack)
        memcpy(tbl.s, sp->h_slots, (size_t)sp->n_slots * sizeof(ob2_slot_desc));
        const size_t smem_p = (size_t)(sp->cdf_len > 0 ? sp->cdf_len : 1) * sizeof(double);
        int rc_p = prepare_smem(kern_p, smem_p);
        if (rc_p != OB2_OK) return rc_p;
        kern_p<<<blocks, OB2_BLOCK, smem_p, static_cast<cudaStream_t>(a.stream)>>>(
            tbl, sp->n_slots, sp->d_cdf, sp->cdf_len, a.n, a.n_stride, 4 * a.n_stride, 8 * a.n_stride, (uint32_t)a.seed,
            (uint32_t)(a.seed >> 32), 0xD2511F53u, 0xCD9E8D57u, a.base_index, a.index, a.U, (char *)a.cols,
            (char *)a.tcols, bdelta, a.mask, a.fp, a.n_failed);
        return ob2_check_cuda(cudaGetLastError(), "sample_kernel_p launch");
    }
    int rc = prepare_smem(kern, sp->smem_bytes);
    if (rc != OB2_OK) return rc;
    kern<<<blocks, OB2_BLOCK, sp->smem_bytes, static_cast<cudaStream_t>(a.stream)>>>(
        sp->d_slots, sp->n_slots, sp->d_cdf, sp->cdf_len, a.n, a.n_stride, 4 * a.n_stride, 8 * a.n_stride, (uint32_t)a.seed,
        (uint32_t)(a.seed >> 32), 0xD2511F53u, 0xCD9E8D57u, a.base_index, a.index, a.U, (char *)a.cols, (char *)a.tcols, bdelta, a.mask, a.fp,
        a.n_failed);
    return ob2_check_cuda(cudaGetLastError(), "sample_kernel launch");
}

template <bool PHILOX>
int launch_sample(const SampleArgs &a) {
    int rc = ob2_check_batch(a.sp, a.n, a.n_stride);
    if (rc != OB2_OK) return rc;
    if (a.sp->unsamplable_slot >= 0) {
        ob2_set_error("sample: slot %d has a prior without a device inverse CDF (no CPU fallback)", a.sp->unsamplable_slot);
        return OB2_EINVAL;
    }
    if (a.cols == nullptr || (!PHILOX && a.U == nullptr)) {
        ob2_set_error("sample: null buffer");
        return OB2_EINVAL;
    }
    OB2_ALIGNED_AS(a.cols, "cols", 16, "sample");
    OB2_ALIGNED_AS(a.tcols, "tcols", 16, "sample");
    OB2_ALIGNED_AS(a.back, "back", 16, "sample");
    OB2_ALIGNED_AS(a.fp, "fp", 16, "sample");
    OB2_ALIGNED_AS(a.U, "U", 16, "sample");
    OB2_ALIGNED_AS(a.mask, "mask", 2, "sample");
    OB2_ALIGNED_AS(a.index, "index", 8, "sample");
    OB2_ALIGNED_AS(a.n_failed, "n_failed", 4, "sample");
    if (a.n == 0) return OB2_OK;
    if (a.n_stride >= (1ll << 29)) {  // the kernel steps through columns with 32-bit byte strides
        ob2_set_error("sample: n_stride %lld too large for one launch (limit 2^29 - 1); split the batch",
                      (long long)a.n_stride);
        return OB2_ETOOBIG;
    }
    const int64_t threads = (a.n + OB2_CPT - 1) / OB2_CPT;
    const int64_t blocks = (threads + OB2_BLOCK - 1) / OB2_BLOCK;
    if (blocks > 0x7FFFFFFFll) {
        ob2_set_error("batch too large for one launch");
        return OB2_ETOOBIG;
    }
    const unsigned b = (unsigned)blocks;
    const bool with_fp = a.fp != nullptr, with_t = a.tcols != nullptr;
    if (a.back != nullptr) {  // ob2_suggest: Philox, transform, reverse and fingerprint, all in the one pass
        if (a.sp->has_extended_prior) return launch_one<PHILOX, true, true, true, PHILOX>(a, b);
        return launch_one<PHILOX, true, true, false, PHILOX>(a, b);
    }
    // EXT = the space uses one of the extra closed-form priors: those inverse CDFs live in a
    // separate instantiation so that the common kernel keeps its registers and code size
    if (a.sp->has_extended_prior) {
        if (with_fp && with_t) return launch_one<PHILOX, true, true, true>(a, b);
        if (with_fp) return launch_one<PHILOX, true, false, true>(a, b);
        if (with_t) return launch_one<PHILOX, false, true, true>(a, b);
        return launch_one<PHILOX, false, false, true>(a, b);
    }
    if (with_fp && with_t) return launch_one<PHILOX, true, true, false>(a, b);
    if (with_fp) return launch_one<PHILOX, true, false, false>(a, b);
    if (with_t) return launch_one<PHILOX, false, true, false>(a, b);
    return launch_one<PHILOX, false, false, false>(a, b);
}

}  // namespace

extern "C" int ob2_sample(const ob2_space *space, int64_t n, int64_t n_stride, uint64_t seed, uint64_t base_index,
                          const uint64_t *index, void *cols, uint8_t *mask, uint64_t *fp, uint32_t *n_failed,
                          void *stream) {
    const SampleArgs a{space, n, n_stride, seed, base_index, index, nullptr, cols, nullptr, nullptr, mask, fp, n_failed, stream};
    return launch_sample<true>(a);
}

extern "C" int ob2_sample_transform(const ob2_space *space, int64_t n, int64_t n_stride, uint64_t seed,
                                    uint64_t base_index, const uint64_t *index, void *cols, void *tcols,
                                    uint8_t *mask, uint64_t *fp, uint32_t *n_failed, void *stream) {
    if (tcols == nullptr) {
        ob2_set_error("ob2_sample_transform: null transformed buffer");
        return OB2_EINVAL;
    }
    const SampleArgs a{space, n, n_stride, seed, base_index, index, nullptr, cols, tcols, nullptr, mask, fp, n_failed, stream};
    return launch_sample<true>(a);
}

extern "C" int ob2_suggest(const ob2_space *space, int64_t n, int64_t n_stride, uint64_t seed, uint64_t base_index,
                           const uint64_t *index, void *cols, void *tcols, void *back, uint8_t *mask, uint64_t *fp,
                           uint32_t *n_failed, void *stream) {
    if (tcols == nullptr || back == nullptr || fp == nullptr) {
        ob2_set_error("ob2_suggest: null transformed / handed-out / fingerprint buffer");
        return OB2_EINVAL;
    }
    const SampleArgs a{space, n, n_stride, seed, base_index, index, nullptr, cols, tcols, back, mask, fp, n_failed, stream};
    return launch_sample<true>(a);
}

namespace {
size_t stage_counter_bytes(const ob2_space *space) {
    return ((size_t)space->n_slots * sizeof(uint32_t) + 255) / 256 * 256;
}

// staging block of the *_host entry points: [columns of up to OB2_HOST_SAMPLE_MAX rows][counters or mask]
int ensure_stage(ob2_space *space) {
    if (space->d_stage != nullptr) return OB2_OK;
    // behind the columns: rejection counters, then per row a verdict byte, a fingerprint and a scratch word
    const size_t bytes = (size_t)space->row_bytes * OB2_HOST_SAMPLE_MAX + stage_counter_bytes(space) +
                         (size_t)OB2_HOST_SAMPLE_MAX * (1 + 16 + 4);
    OB2_CUDA(cudaMalloc(&space->d_stage, bytes));
    const int rc = ob2_check_cuda(cudaMallocHost(&space->h_stage, bytes), "cudaMallocHost(stage)");
    if (rc != OB2_OK) {
        cudaFree(space->d_stage);
        space->d_stage = nullptr;
    }
    return rc;
}
}  // namespace

// Small draws straight into host memory: Space.sample(n) as the reference's callers use it
// (Producer.produce asks for a handful of points at a time, producer.py:45).  One kernel, one
// device->host copy, one stream synchronisation -- instead of the allocations, the counter
// read-back and the pageable copy the general tensor path pays (~340 us -> tens of us per call).
extern "C" int ob2_sample_host(ob2_space *space, int64_t n, uint64_t seed, uint64_t base_index, void *host_cols,
                               uint32_t *host_n_failed, void *stream) {
    if (space == nullptr || host_cols == nullptr || host_n_failed == nullptr) {
        ob2_set_error("ob2_sample_host: null argument");
        return OB2_EINVAL;
    }
    if (n < 0 || n > OB2_HOST_SAMPLE_MAX) {
        ob2_set_error("ob2_sample_host: n = %lld outside [0, %d]; use ob2_sample for bulk draws", (long long)n,
                      (int)OB2_HOST_SAMPLE_MAX);
        return OB2_ETOOBIG;
    }
    const int64_t n_stride = (n + 31) / 32 * 32;
    int rc = ob2_check_batch(space, n, n_stride);
    if (rc != OB2_OK) return rc;
    const size_t fail_bytes = (size_t)space->n_slots * sizeof(uint32_t);
    std::lock_guard<std::mutex> lock(*space->stage_mutex);
    rc = ensure_stage(space);
    if (rc != OB2_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char *d = static_cast<char *>(space->d_stage), *h = static_cast<char *>(space->h_stage);
    const size_t used = (size_t)space->row_bytes * n_stride;
    // the counters sit right behind the columns in use, so ONE copy brings both back
    uint32_t *d_failed = reinterpret_cast<uint32_t *>(d + used);
    OB2_CUDA(cudaMemsetAsync(d_failed, 0, fail_bytes, st));
    if (n > 0) {
        const SampleArgs a{space, n, n_stride, seed, base_index, nullptr, nullptr, d, nullptr, nullptr, nullptr, nullptr, d_failed, st};
        rc = launch_sample<true>(a);
        if (rc != OB2_OK) return rc;
    }
    OB2_CUDA(cudaMemcpyAsync(h, d, used + fail_bytes, cudaMemcpyDeviceToHost, st));
    OB2_CUDA(cudaStreamSynchronize(st));
    memcpy(host_cols, h, used);
    memcpy(host_n_failed, h + used, fail_bytes);
    return OB2_OK;
}

// Random.suggest with the producer's dedup for a handful of points: draw, fingerprint, dedup
// against the history set, and bring columns + verdicts + counters back with ONE copy.
extern "C" int ob2_sample_dedup_host(ob2_space *space, ob2_set *set, int64_t n, uint64_t seed, uint64_t base_index,
                                     void *host_cols, uint8_t *host_keep, uint32_t *host_n_failed, void *stream) {
    if (space == nullptr || set == nullptr || host_cols == nullptr || host_keep == nullptr || host_n_failed == nullptr) {
        ob2_set_error("ob2_sample_dedup_host: null argument");
        return OB2_EINVAL;
    }
    if (n < 0 || n > OB2_HOST_SAMPLE_MAX) {
        ob2_set_error("ob2_sample_dedup_host: n = %lld outside [0, %d]; use ob2_sample + ob2_set_dedup for bulk",
                      (long long)n, (int)OB2_HOST_SAMPLE_MAX);
        return OB2_ETOOBIG;
    }
    const size_t fail_bytes = (size_t)space->n_slots * sizeof(uint32_t);
    if (n == 0) {
        memset(host_n_failed, 0, fail_bytes);
        return OB2_OK;
    }
    const int64_t n_stride = (n + 31) / 32 * 32;
    int rc = ob2_check_batch(space, n, n_stride);
    if (rc != OB2_OK) return rc;
    std::lock_guard<std::mutex> lock(*space->stage_mutex);
    rc = ensure_stage(space);
    if (rc != OB2_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char *d = static_cast<char *>(space->d_stage), *h = static_cast<char *>(space->h_stage);
    const size_t used = (size_t)space->row_bytes * n_stride, counters = stage_counter_bytes(space);
    uint32_t *d_failed = reinterpret_cast<uint32_t *>(d + used);
    uint8_t *d_keep = reinterpret_cast<uint8_t *>(d + used + counters);
    uint64_t *d_fp = reinterpret_cast<uint64_t *>(d + used + counters + n_stride);
    uint32_t *d_scratch = reinterpret_cast<uint32_t *>(d + used + counters + n_stride + 16 * n_stride);
    OB2_CUDA(cudaMemsetAsync(d_failed, 0, fail_bytes, st));
    const SampleArgs a{space, n, n_stride, seed, base_index, nullptr, nullptr, d, nullptr, nullptr, nullptr, d_fp, d_failed, st};
    rc = launch_sample<true>(a);
    if (rc != OB2_OK) return rc;
    rc = ob2_set_dedup(set, n, d_fp, nullptr, d_scratch, d_keep, st);
    if (rc != OB2_OK) return rc;
    OB2_CUDA(cudaMemcpyAsync(h, d, used + counters + (size_t)n_stride, cudaMemcpyDeviceToHost, st));
    OB2_CUDA(cudaStreamSynchronize(st));
    memcpy(host_cols, h, used);
    memcpy(host_n_failed, h + used, fail_bytes);
    memcpy(host_keep, h + used + counters, (size_t)n);
    return OB2_OK;
}

// `point in space` for a handful of host points (PrimaryAlgo.suggest / observe / score / judge,
// primary_algo.py:53-54,64-66,78-84): copy in, check, copy the verdicts out, one synchronisation.
extern "C" int ob2_contains_host(ob2_space *space, int64_t n, const void *host_cols, int32_t int_as_f64,
                                 uint8_t *host_mask, void *stream) {
    if (space == nullptr || host_cols == nullptr || host_mask == nullptr) {
        ob2_set_error("ob2_contains_host: null argument");
        return OB2_EINVAL;
    }
    if (n < 0 || n > OB2_HOST_SAMPLE_MAX) {
        ob2_set_error("ob2_contains_host: n = %lld outside [0, %d]; use ob2_contains for bulk checks", (long long)n,
                      (int)OB2_HOST_SAMPLE_MAX);
        return OB2_ETOOBIG;
    }
    if (n == 0) return OB2_OK;
    const int64_t n_stride = (n + 31) / 32 * 32;
    int rc = ob2_check_batch(space, n, n_stride);
    if (rc != OB2_OK) return rc;
    std::lock_guard<std::mutex> lock(*space->stage_mutex);
    rc = ensure_stage(space);
    if (rc != OB2_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char *d = static_cast<char *>(space->d_stage), *h = static_cast<char *>(space->h_stage);
    const size_t used = (size_t)space->row_bytes * n_stride;
    memcpy(h, host_cols, used);
    OB2_CUDA(cudaMemcpyAsync(d, h, used, cudaMemcpyHostToDevice, st));
    rc = ob2_contains(space, n, n_stride, d, int_as_f64, reinterpret_cast<uint8_t *>(d + used), st);
    if (rc != OB2_OK) return rc;
    OB2_CUDA(cudaMemcpyAsync(h + used, d + used, (size_t)n_stride, cudaMemcpyDeviceToHost, st));
    OB2_CUDA(cudaStreamSynchronize(st));
    memcpy(host_mask, h + used, (size_t)n);
    return OB2_OK;
}

extern "C" int ob2_sample_from_uniform(const ob2_space *space, int64_t n, int64_t n_stride, const double *U,
                                       void *cols, uint8_t *ok, uint64_t *fp, void *stream) {
    const SampleArgs a{space, n, n_stride, 0, 0, nullptr, U, cols, nullptr, nullptr, ok, fp, nullptr, stream};
    return launch_sample<false>(a);
}

extern "C" int ob2_contains(const ob2_space *space, int64_t n, int64_t n_stride, const void *cols,
                            int32_t int_as_f64, uint8_t *mask, void *stream) {
    int rc = ob2_check_batch(space, n, n_stride);
    if (rc != OB2_OK) return rc;
    if (cols == nullptr || mask == nullptr) {
        ob2_set_error("contains: null buffer");
        return OB2_EINVAL;
    }
    OB2_ALIGNED(cols, 16, "contains");
    OB2_ALIGNED(mask, 2, "contains");
    if (n == 0) return OB2_OK;
    const int64_t threads = (n + OB2_CPT - 1) / OB2_CPT;
    const int64_t blocks = (threads + OB2_BLOCK - 1) / OB2_BLOCK;
    const size_t smem = (size_t)space->n_slots * sizeof(ob2_slot_desc);
    rc = prepare_smem(contains_kernel, smem);
    if (rc != OB2_OK) return rc;
    contains_kernel<<<(unsigned)blocks, OB2_BLOCK, smem, static_cast<cudaStream_t>(stream)>>>(
        space->d_slots, space->n_slots, n, n_stride, (const char *)cols, int_as_f64, mask);
    return ob2_check_cuda(cudaGetLastError(), "contains_kernel launch");
}
