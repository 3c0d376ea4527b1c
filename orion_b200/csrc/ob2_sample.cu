// K1: candidate sampling (Philox4x32 or injected uniforms -> per-slot
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
// + fingerprint multiplies: ~3.1k warp instructions per candidate of mixed64,
// 75 % issue utilisation, profiles/r2f_sampler.md) while transform and reverse are
// pure memory traffic; fusing them in (WITH_T, WITH_B) lets the stores ride under
// the arithmetic and removes every re-read: the fused kernel runs at the HBM
// write ceiling (profiles/r2k_full_step.md).
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
// Otherwise the branch-free descent.  (53-bit / injected uniforms.)
__device__ __forceinline__ int draw_cat_fast(const double *cdf, int k, double u, bool uniform) {
    if (!uniform) return draw_cat(cdf, k, u);
    int idx = __double2int_rz(u * (double)k);
    idx = min(max(idx, 0), k);
    while (idx > 0 && cdf[idx - 1] > u) --idx;
    while (idx < k && cdf[idx] <= u) ++idx;
    return idx;
}

// The same count on a 32-bit word w (u = w / 2^32) against the integer thresholds
// thr[j] = ceil(cdf[j] * 2^32), j < k - 1:  cdf[j] <= u  <=>  thr[j] <= w.
__device__ __forceinline__ int draw_cat_u32(const uint32_t *thr, int k, uint32_t w, bool uniform) {
    const int m = k - 1;  // thresholds that can be met (cdf[k - 1] = 1 > u)
    if (uniform) {
        int idx = min((int)__umulhi(w, (uint32_t)k), m);
        while (idx > 0 && thr[idx - 1] > w) --idx;
        while (idx < m && thr[idx] <= w) ++idx;
        return idx;
    }
    int lo = 0;
    if (m <= 0) return 0;
    for (int step = 1 << (31 - __clz(m)); step > 0; step >>= 1) {
        const int probe = lo + step;
        const int at = probe <= m ? probe : m;
        const bool take = (probe <= m) && (thr[at - 1] <= w);
        lo = take ? probe : lo;
    }
    return lo;
}

__device__ __forceinline__ uint32_t pick_word(const Philox4 &r, uint32_t word) {
    return word == 0 ? r.x : word == 1 ? r.y : word == 2 ? r.z : r.w;
}

// Real.sample (space.py:263-270): a draw that lands outside [low, high) is redrawn -- the whole
// (possibly shaped) value of that dimension, up to 4 attempts.  The walk itself only ANDs the
// bounds bits; a candidate that missed somewhere (rare) comes here afterwards, ONCE: the first
// attempt of every Real / Integer dimension is recomputed (same arithmetic, nothing stored; not
// read back, because with `back` aliased to `cols` a lossy Quantize has overwritten the draw),
// and those that missed are redrawn with attempts 1..3 (counter word 3), overwriting their
// columns with scalar stores.  The fingerprint lanes are wrapping sums, so the contribution of
// the rejected values is subtracted and the redrawn one added.  Out of line and everything by
// value: this is the only call in the kernel, made when the walk's state is dead, so that the
// hot loops keep their registers.  Only slots drawn from a 53-bit uniform can miss (the 32-bit
// classes are in bounds by construction), so every slot here owns words (0, 1) or (2, 3) of its block.
// WITH_B: the fingerprint is over the handed-out row (`back`, at cols + bdelta), which is also rewritten.
struct Fixed {
    FpAcc acc;
    bool ok;
};

// Attempt `a` of dimension [first, s_last] for one candidate: its values, optionally stored,
// their fingerprint contribution in `t`; returns whether every element is in bounds.
template <bool STORE, bool WITH_FP, bool WITH_T, bool WITH_B, bool EXT, int ROUNDS>
__device__ __forceinline__ bool draw_dimension(const ob2_slot_desc *s_slots, const ob2_slot_plan *s_plan, int first, int s_last,
                                               uint32_t a, uint64_t index, int64_t i, int64_t n_stride, char *cols, char *tcols,
                                               int64_t bdelta, uint32_t k0, uint32_t k1, uint32_t pm0, uint32_t pm1, FpAcc &t) {
    bool ok = true;
    t = FpAcc{0, 0};
    for (int e = first; e <= s_last; ++e) {
        const ob2_slot_desc &d = s_slots[e];
        const Philox4 r = philox4x32<ROUNDS>((uint32_t)index, (uint32_t)(index >> 32), (uint32_t)s_plan[e].block, a, k0, k1,
                                             pm0, pm1);
        const double u = s_plan[e].word ? u53(r.z, r.w) : u53(r.x, r.y);
        double x = draw_real<EXT>(d, u);
        char *col = cols + (int64_t)d.col_units * n_stride;
        char *dst = tcols + (int64_t)d.t_units * n_stride;
        uint64_t w;
        if (d.kind == OB2_KIND_INTEGER) {
            x = floor(x);
            const long long q = __double2ll_rz(x);
            if (STORE) reinterpret_cast<long long *>(col)[i] = q;
            long long bq = q;
            if (WITH_T) {
                if (d.t_op == OB2_OP_TO_REAL) {
                    if (STORE) reinterpret_cast<double *>(dst)[i] = x;
                    if (WITH_B) bq = quantize_ll(x, d.t_flags);
                } else if (STORE) {
                    reinterpret_cast<long long *>(dst)[i] = q;
                }
            }
            if (STORE && WITH_B && bdelta != 0) reinterpret_cast<long long *>(col + bdelta)[i] = bq;
            w = (uint64_t)bq;
        } else {
            if (STORE) reinterpret_cast<double *>(col)[i] = x;
            double bx = x;
            if (WITH_T) {
                if (d.t_op == OB2_OP_QUANTIZE) {
                    const long long tq = quantize_ll(x, d.t_flags);
                    if (STORE) reinterpret_cast<long long *>(dst)[i] = tq;
                    bx = __ll2double_rn(tq);
                } else if (STORE) {
                    reinterpret_cast<double *>(dst)[i] = x;
                }
            }
            if (STORE && WITH_B && (bdelta != 0 || d.t_op == OB2_OP_QUANTIZE)) reinterpret_cast<double *>(col + bdelta)[i] = bx;
            w = canon_f64(WITH_B ? bx : x);
        }
        ok &= (x >= d.low) && (x < d.high);
        if (WITH_FP) fp_add(t, d, w);
    }
    return ok;
}

template <bool WITH_FP, bool WITH_T, bool WITH_B, bool EXT, int ROUNDS>
static __device__ __noinline__ Fixed fixup_candidate(const ob2_slot_desc *s_slots, const ob2_slot_plan *s_plan, int n_slots,
                                                     uint64_t index, int64_t i, int64_t n_stride, char *cols, char *tcols,
                                                     int64_t bdelta, uint32_t k0, uint32_t k1, uint32_t pm0, uint32_t pm1,
                                                     FpAcc acc, uint32_t *n_failed) {
    bool all = true;
    for (int first = 0; first < n_slots;) {
        int s_last = first;
        while (!(s_slots[s_last].flags & OB2_SLOT_DIM_LAST)) ++s_last;
        const uint32_t cls = s_plan[first].cls;
        if (cls <= OB2_CLS_NORM || (cls == OB2_CLS_GENERIC && s_slots[first].kind != OB2_KIND_CATEGORICAL)) {
            FpAcc t;
            bool ok = draw_dimension<false, WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>(s_slots, s_plan, first, s_last, 0u, index, i,
                                                                                 n_stride, cols, tcols, bdelta, k0, k1, pm0, pm1, t);
            if (!ok) {
                if (WITH_FP) {  // the rejected first attempt leaves the fingerprint
                    acc.a -= t.a;
                    acc.b -= t.b;
                }
                for (uint32_t a = 1; a < 4 && !ok; ++a)
                    ok = draw_dimension<true, WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>(s_slots, s_plan, first, s_last, a, index, i,
                                                                                   n_stride, cols, tcols, bdelta, k0, k1, pm0, pm1, t);
                if (WITH_FP) {  // values of a 4-times-failed dimension stay as the last attempt
                    acc.a += t.a;
                    acc.b += t.b;
                }
                if (!ok && n_failed != nullptr) atomicAdd(n_failed + first, 1u);  // the reference raises "Improbable bounds"
                all &= ok;
            }
        }
        first = s_last + 1;
    }
    return Fixed{acc, all};
}

// Per-thread state of the slot walk (two candidates).  Column pointers are running pointers
// advanced by 4 * n_stride / 8 * n_stride; the fingerprint accumulates straight into its lanes.
struct Walk {
    FpAcc acc0, acc1;
    bool all0, all1;  // every bounds check so far passed (first attempt)
    char *col;        // element i0 of the current column (a 4-byte column is addressed at half the offset)
    char *tcol;
    Philox4 r0, r1;   // the current Philox block of the two candidates
};

struct WalkConst {  // loop invariants
    int64_t i0, n_stride, stride4, stride8, bdelta;
    bool split, has1;
    uint64_t idx0, idx1;
    uint32_t k0, k1, pm0, pm1;
    bool base_normal;  // injected mode only: norm slots receive standard normal variates, not uniforms
};

// ---- what a drawn pair of values costs downstream: stores (sampled / transformed / handed-out), the
// fingerprint and the bounds bit.  WITH_B: reverse() of the transformed value, evaluated on what is stored.
template <bool WITH_FP, bool WITH_T, bool WITH_B>
__device__ __forceinline__ void emit_real(const ob2_slot_desc &d, double x0, double x1, Walk &w, const WalkConst &c) {
    st_f64x2(reinterpret_cast<double *>(w.col), x0, x1);
    double bx0 = x0, bx1 = x1;
    if (WITH_T) {
        if (d.t_op == OB2_OP_QUANTIZE) {
            const long long t0 = quantize_ll(x0, d.t_flags), t1 = quantize_ll(x1, d.t_flags);
            st_i64x2(reinterpret_cast<long long *>(w.tcol), t0, t1);
            if (WITH_B) {
                bx0 = __ll2double_rn(t0);
                bx1 = __ll2double_rn(t1);
            }
        } else {
            st_f64x2(reinterpret_cast<double *>(w.tcol), x0, x1);
        }
        w.tcol += c.stride8;
    }
    // (aliased: only a lossy Quantize round trip differs from what was just stored)
    if (WITH_B && (c.split || d.t_op == OB2_OP_QUANTIZE)) st_f64x2(reinterpret_cast<double *>(w.col + c.bdelta), bx0, bx1);
    if (WITH_FP) {
        fp_add(w.acc0, d, canon_f64(bx0));
        fp_add(w.acc1, d, canon_f64(bx1));
    }
    w.col += c.stride8;
    w.all0 &= (x0 >= d.low) & (x0 < d.high);  // half-open, space.py:164
    w.all1 &= (x1 >= d.low) & (x1 < d.high);
}

// Integer slot from the floored double (general priors)
template <bool WITH_FP, bool WITH_T, bool WITH_B>
__device__ __forceinline__ void emit_int_f(const ob2_slot_desc &d, double x0, double x1, Walk &w, const WalkConst &c) {
    const long long q0 = __double2ll_rz(x0), q1 = __double2ll_rz(x1);
    st_i64x2(reinterpret_cast<long long *>(w.col), q0, q1);
    long long bq0 = q0, bq1 = q1;
    if (WITH_T) {
        if (d.t_op == OB2_OP_TO_REAL) {  // the floored double is exactly (double)q
            st_f64x2(reinterpret_cast<double *>(w.tcol), x0, x1);
            if (WITH_B) {  // Reverse(Quantize).reverse = Quantize.transform of the stored double
                bq0 = quantize_ll(x0, d.t_flags);
                bq1 = quantize_ll(x1, d.t_flags);
            }
        } else {
            st_i64x2(reinterpret_cast<long long *>(w.tcol), q0, q1);
        }
        w.tcol += c.stride8;
    }
    if (WITH_B && c.split) st_i64x2(reinterpret_cast<long long *>(w.col + c.bdelta), bq0, bq1);
    if (WITH_FP) {
        fp_add(w.acc0, d, (uint64_t)bq0);
        fp_add(w.acc1, d, (uint64_t)bq1);
    }
    w.col += c.stride8;
    w.all0 &= (x0 >= d.low) & (x0 < d.high);
    w.all1 &= (x1 >= d.low) & (x1 < d.high);
}

// Integer slot from the exact integer (OB2_CLS_INT_MULHI: |q| < 2^21, in bounds by construction;
// the cast to double and the Quantize of it back are exact, so the handed-out value is q itself)
template <bool WITH_FP, bool WITH_T, bool WITH_B>
__device__ __forceinline__ void emit_int_q(const ob2_slot_desc &d, int v0, int v1, Walk &w, const WalkConst &c) {
    const long long q0 = v0, q1 = v1;
    st_i64x2(reinterpret_cast<long long *>(w.col), q0, q1);
    if (WITH_T) {
        if (d.t_op == OB2_OP_TO_REAL) {
            st_f64x2(reinterpret_cast<double *>(w.tcol), __int2double_rn(v0), __int2double_rn(v1));
        } else {
            st_i64x2(reinterpret_cast<long long *>(w.tcol), q0, q1);
        }
        w.tcol += c.stride8;
    }
    if (WITH_B && c.split) st_i64x2(reinterpret_cast<long long *>(w.col + c.bdelta), q0, q1);
    if (WITH_FP) {
        fp_add(w.acc0, d, (uint64_t)q0);
        fp_add(w.acc1, d, (uint64_t)q1);
    }
    w.col += c.stride8;
}

template <bool WITH_FP, bool WITH_T, bool WITH_B>
__device__ __forceinline__ void emit_cat(const ob2_slot_desc &d, int c0, int c1, Walk &w, const WalkConst &c) {
    st_i32x2(reinterpret_cast<int *>(w.col - c.i0 * 4), c0, c1);
    int *bcol = reinterpret_cast<int *>(w.col - c.i0 * 4 + c.bdelta);
    w.col += c.stride4;
    int b0 = c0, b1 = c1;  // WITH_B: reverse() of the transformed value, evaluated on what is stored
    if (WITH_T) {
        if (d.t_op != OB2_OP_ONEHOT) {  // Enumerate / Identity: the index both ways
            st_i32x2(reinterpret_cast<int *>(w.tcol - c.i0 * 4), c0, c1);
            w.tcol += c.stride4;
        } else if (d.t_width == 1) {
            const double t0 = (double)c0, t1 = (double)c1;
            st_f64x2(reinterpret_cast<double *>(w.tcol), t0, t1);
            w.tcol += c.stride8;
            if (WITH_B) {  // k == 2: (v > 0.5); k == 1: always category 0
                b0 = (d.n_cats == 2 && t0 > 0.5) ? 1 : 0;
                b1 = (d.n_cats == 2 && t1 > 0.5) ? 1 : 0;
            }
        } else {
            // column j is at tcol + j * stride8; the launcher guarantees stride8 < 2^32,
            // so each address is ONE 32x32+64 multiply-add (IMAD.WIDE.U32)
            const uint32_t s8 = (uint32_t)c.stride8, k = (uint32_t)d.t_width;
            // WITH_B: reverse_kernel's argmax (first maximum) over these k values is the position
            // of the single 1.0 stored below, i.e. the index itself: the Philox draw is always in
            // [0, k).  Tracking the maximum while storing (tried: 6 more instructions per column
            // pair) costs the kernel 4.5 % for the same bytes, so the identity is used; the staged
            // kernels and tests/test_gpu_suggest_fused.py hold it to account.
#pragma unroll 4
            for (uint32_t j = 0; j < k; ++j) {
                const unsigned h0 = c0 == (int)j ? 0x3FF00000u : 0u, h1 = c1 == (int)j ? 0x3FF00000u : 0u;
                st_onehot_hi(reinterpret_cast<double *>(w.tcol + (uint64_t)j * s8), h0, h1);
            }
            w.tcol += (uint64_t)k * s8;
        }
    }
    if (WITH_B && c.split) st_i32x2(bcol, b0, b1);
    if (WITH_FP) {
        fp_add(w.acc0, d, (uint64_t)(int64_t)b0);
        fp_add(w.acc1, d, (uint64_t)(int64_t)b1);
    }
}

// One slot through the general prior switch, from a pair of double uniforms (injected uniforms,
// and in Philox mode the slots no specialised run takes: further priors, Integer over a real prior,
// constants, categoricals whose thresholds do not fit 32 bits).
template <bool PHILOX, bool WITH_FP, bool WITH_T, bool WITH_B, bool EXT, int ROUNDS>
__device__ __forceinline__ void generic_slot(const ob2_slot_desc *s_slots, const ob2_slot_plan *s_plan, const double *s_cdf,
                                             int s, double u0, double u1, Walk &w, const WalkConst &c) {
    const ob2_slot_desc &d = s_slots[s];
    if (d.kind == OB2_KIND_CATEGORICAL) {
        const double *cdf = s_cdf + d.cdf_off;
        const bool uni = (d.flags & OB2_SLOT_CAT_UNIFORM) != 0;
        const int c0 = draw_cat_fast(cdf, d.n_cats, u0, uni);
        const int c1 = draw_cat_fast(cdf, d.n_cats, u1, uni);
        emit_cat<WITH_FP, WITH_T, WITH_B>(d, c0, c1, w, c);
        if (!PHILOX) {  // an injected u >= 1 can run off the table; a Philox u cannot
            w.all0 &= c0 < d.n_cats;
            w.all1 &= c1 < d.n_cats;
        }
    } else {
        double x0, x1;
        if (!PHILOX && c.base_normal && d.dist == OB2_DIST_NORM) {
            // the variate itself arrives (what RandomState.standard_normal hands scipy's norm._rvs)
            x0 = __dadd_rn(__dmul_rn(u0, d.scale), d.loc);
            x1 = __dadd_rn(__dmul_rn(u1, d.scale), d.loc);
        } else {
            x0 = draw_real<EXT>(d, u0);
            x1 = draw_real<EXT>(d, u1);
        }
        if (d.kind == OB2_KIND_INTEGER) {
            x0 = floor(x0);  // _Discrete.sample, space.py:294
            x1 = floor(x1);
            emit_int_f<WITH_FP, WITH_T, WITH_B>(d, x0, x1, w, c);
        } else {
            emit_real<WITH_FP, WITH_T, WITH_B>(d, x0, x1, w, c);
        }
    }
}

// The Philox block `blk` of both candidates (attempt 0).
template <int ROUNDS>
__device__ __forceinline__ void new_block(uint32_t blk, Walk &w, const WalkConst &c) {
    w.r0 = philox4x32<ROUNDS>((uint32_t)c.idx0, (uint32_t)(c.idx0 >> 32), blk, 0u, c.k0, c.k1, c.pm0, c.pm1);
    w.r1 = philox4x32<ROUNDS>((uint32_t)c.idx1, (uint32_t)(c.idx1 >> 32), blk, 0u, c.k0, c.k1, c.pm0, c.pm1);
}

// ---- a run of Real slots of one prior off 53-bit uniforms: words (0, 1) then (2, 3) of a block
template <int CLS, bool WITH_FP, bool WITH_T, bool WITH_B, bool EXT, int ROUNDS>
__device__ __forceinline__ void real_slot(const ob2_slot_desc *s_slots, const ob2_slot_plan *s_plan, int e, double u0,
                                          double u1, Walk &w, const WalkConst &c) {
    const ob2_slot_desc &d = s_slots[e];
    double b0, b1;
    if (CLS == OB2_CLS_UNIFORM) {
        b0 = u0;
        b1 = u1;
    } else if (CLS == OB2_CLS_RECIP) {
        b0 = ob2_exp_bounded<WITH_T>(__dadd_rn(d.p0, __dmul_rn(u0, d.p1)));
        b1 = ob2_exp_bounded<WITH_T>(__dadd_rn(d.p0, __dmul_rn(u1, d.p1)));
    } else {
        b0 = ob2_ndtri<WITH_T>(u0);
        b1 = ob2_ndtri<WITH_T>(u1);
    }
    // base(u) * scale + loc, rounded multiply then rounded add (scipy rv_generic.rvs)
    emit_real<WITH_FP, WITH_T, WITH_B>(d, __dadd_rn(__dmul_rn(b0, d.scale), d.loc), __dadd_rn(__dmul_rn(b1, d.scale), d.loc),
                                       w, c);
}

template <int CLS, bool WITH_FP, bool WITH_T, bool WITH_B, bool EXT, int ROUNDS>
__device__ __forceinline__ void run_real(const ob2_slot_desc *s_slots, const ob2_slot_plan *s_plan, int s, int end,
                                         Walk &w, const WalkConst &c) {
    if (WITH_T) {
        // store-heavy variants: ONE copy of the slot body, its words selected by the slot's position in the
        // block (a warp-uniform predicate) -- the pairwise form below costs these kernels registers
        // (suggest 4.25 -> 4.41 ms) where it gains the plain sampler 2 %
#pragma unroll 1
        for (; s < end; ++s) {
            const bool first = s_plan[s].word == 0;
            if (first) new_block<ROUNDS>(s_plan[s].block, w, c);
            const double u0 = first ? u53(w.r0.x, w.r0.y) : u53(w.r0.z, w.r0.w);
            const double u1 = first ? u53(w.r1.x, w.r1.y) : u53(w.r1.z, w.r1.w);
            real_slot<CLS, WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>(s_slots, s_plan, s, u0, u1, w, c);
        }
        return;
    }
    if (s < end && s_plan[s].word != 0) {  // the run begins in the second half of a block another class opened
        real_slot<CLS, WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>(s_slots, s_plan, s, u53(w.r0.z, w.r0.w), u53(w.r1.z, w.r1.w), w, c);
        ++s;
    }
    // from here on every block belongs to this run: slot pairs, words (0, 1) then (2, 3)
    if (CLS == OB2_CLS_NORM) {
        // the long body: ONE Philox site and a predicated second slot (the kernel's code has to stay within
        // reach of the instruction cache: mixed64 sample 1.66 -> 1.62 ms, with fingerprint 2.42 -> 2.32)
#pragma unroll 1
        for (; s < end; s += 2) {
            new_block<ROUNDS>(s_plan[s].block, w, c);
            real_slot<CLS, WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>(s_slots, s_plan, s, u53(w.r0.x, w.r0.y), u53(w.r1.x, w.r1.y), w, c);
            if (s + 1 < end)
                real_slot<CLS, WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>(s_slots, s_plan, s + 1, u53(w.r0.z, w.r0.w), u53(w.r1.z, w.r1.w), w, c);
        }
    } else {
        // the short bodies: an unconditional pair per iteration (real16 loses 5 % to the predicate) and a remainder
#pragma unroll 1
        for (; s + 2 <= end; s += 2) {
            new_block<ROUNDS>(s_plan[s].block, w, c);
            real_slot<CLS, WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>(s_slots, s_plan, s, u53(w.r0.x, w.r0.y), u53(w.r1.x, w.r1.y), w, c);
            real_slot<CLS, WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>(s_slots, s_plan, s + 1, u53(w.r0.z, w.r0.w), u53(w.r1.z, w.r1.w), w, c);
        }
        if (s < end) {
            new_block<ROUNDS>(s_plan[s].block, w, c);
            real_slot<CLS, WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>(s_slots, s_plan, s, u53(w.r0.x, w.r0.y), u53(w.r1.x, w.r1.y), w, c);
        }
    }
}

// ---- a run of slots that take one 32-bit word each: four share a block; integer arithmetic only,
// in bounds by construction (no bounds bit, no redraw)
template <int CLS, bool WITH_FP, bool WITH_T, bool WITH_B>
__device__ __forceinline__ void word_slot(const ob2_slot_desc *s_slots, const ob2_slot_plan *s_plan, const uint32_t *s_thr,
                                          int e, uint32_t w0, uint32_t w1, Walk &w, const WalkConst &c) {
    const ob2_slot_desc &d = s_slots[e];
    if (CLS == OB2_CLS_INT_MULHI) {
        const uint32_t range = s_plan[e].irange;
        const int loc = s_plan[e].iloc;
        emit_int_q<WITH_FP, WITH_T, WITH_B>(d, loc + (int)__umulhi(w0, range), loc + (int)__umulhi(w1, range), w, c);
    } else if (CLS == OB2_CLS_CAT_MULHI) {
        const uint32_t k = (uint32_t)d.n_cats;
        emit_cat<WITH_FP, WITH_T, WITH_B>(d, (int)__umulhi(w0, k), (int)__umulhi(w1, k), w, c);
    } else {
        const uint32_t *thr = s_thr + s_plan[e].thr_off;
        const bool uni = (d.flags & OB2_SLOT_CAT_UNIFORM) != 0;
        emit_cat<WITH_FP, WITH_T, WITH_B>(d, draw_cat_u32(thr, d.n_cats, w0, uni), draw_cat_u32(thr, d.n_cats, w1, uni), w, c);
    }
}

template <int CLS, bool WITH_FP, bool WITH_T, bool WITH_B, int ROUNDS>
__device__ __forceinline__ void run_words(const ob2_slot_desc *s_slots, const ob2_slot_plan *s_plan, int s, int end,
                                          Walk &w, const WalkConst &c, const uint32_t *s_thr) {
    while (s < end) {
        const uint32_t word = s_plan[s].word;
        if (word == 0) new_block<ROUNDS>(s_plan[s].block, w, c);
        if (word == 0 && s + 4 <= end) {  // the next three slots own words 1, 2, 3
            word_slot<CLS, WITH_FP, WITH_T, WITH_B>(s_slots, s_plan, s_thr, s, w.r0.x, w.r1.x, w, c);
            word_slot<CLS, WITH_FP, WITH_T, WITH_B>(s_slots, s_plan, s_thr, s + 1, w.r0.y, w.r1.y, w, c);
            word_slot<CLS, WITH_FP, WITH_T, WITH_B>(s_slots, s_plan, s_thr, s + 2, w.r0.z, w.r1.z, w, c);
            word_slot<CLS, WITH_FP, WITH_T, WITH_B>(s_slots, s_plan, s_thr, s + 3, w.r0.w, w.r1.w, w, c);
            s += 4;
        } else {
            word_slot<CLS, WITH_FP, WITH_T, WITH_B>(s_slots, s_plan, s_thr, s, pick_word(w.r0, word), pick_word(w.r1, word), w, c);
            ++s;
        }
    }
}

// pm0 / pm1 are the Philox multipliers, passed as kernel parameters so that the
// multiplies read them straight from the constant bank.
//
// The walk is run-specialised (Philox mode): ob2_space_create groups consecutive slots of one
// draw class into runs (ob2_internal.cuh), and each class has its own tight loop here -- no
// per-slot switch on kind / prior, a Philox block shared by the 2 (53-bit) or 4 (32-bit) slots
// that consume it, integer-only arithmetic for the classes that allow it.  The kernel is issue
// bound on its own (profiles/): every instruction not spent is time.
//
// WITH_B (Random.suggest behind PrimaryAlgo, in one pass): also evaluate reverse() on the
// transformed values while they are still in registers and store the handed-out row to
// `back` = cols + bdelta (same column layout); the fingerprint is then that of the `back`
// row, exactly what reverse_kernel<true> computes from the materialised tcols.
//
// sample_body is the whole per-thread job.  The tables are either staged in shared memory
// (sample_kernel) or passed BY VALUE as a kernel parameter (sample_kernel_p): the parameter
// lives in the constant bank, so every descriptor field is read with a uniform load into a
// uniform register and the class / op branches are known to be warp-uniform.
template <bool PHILOX, bool WITH_FP, bool WITH_T, bool WITH_B, bool EXT, int ROUNDS>
__device__ __forceinline__ void sample_body(const ob2_slot_desc *s_slots, const ob2_slot_plan *s_plan,
                                            const ob2_run *s_runs, int n_runs, const double *s_cdf,
                                            const uint32_t *s_thr, int n_slots, int64_t n, int64_t n_stride,
                                            int64_t stride4, int64_t stride8, uint32_t k0, uint32_t k1, uint32_t pm0,
                                            uint32_t pm1, uint64_t base_index, const uint64_t *__restrict__ index,
                                            const double *__restrict__ U, char *__restrict__ cols,
                                            char *__restrict__ tcols, int64_t bdelta, uint8_t *__restrict__ mask,
                                            uint64_t *__restrict__ fp, uint32_t *__restrict__ n_failed) {
    const int64_t i0 = ((int64_t)blockIdx.x * OB2_BLOCK + threadIdx.x) * OB2_CPT;
    if (i0 >= n) return;
    const bool has1 = i0 + 1 < n;
    uint64_t idx0 = base_index + (uint64_t)i0, idx1 = idx0 + 1;
    if (PHILOX && index != nullptr) {  // explicit counters (replay / planted duplicates)
        idx0 = __ldg(index + i0);
        idx1 = has1 ? __ldg(index + i0 + 1) : idx0;
    }
    if (!has1) idx1 = idx0;  // the padding candidate replays its neighbour: always well defined

    WalkConst c;
    c.i0 = i0;
    c.n_stride = n_stride;
    c.stride4 = stride4;  // 4 * n_stride / 8 * n_stride arrive as kernel parameters: the pointer bumps
    c.stride8 = stride8;  // take them straight from the constant bank
    c.bdelta = bdelta;
    // back == cols (bdelta == 0): the caller keeps only the handed-out rows; the second store of every
    // value is skipped (reverse(transform(x)) == x bit for bit except through a lossy Quantize)
    c.split = WITH_B && bdelta != 0;
    c.has1 = has1;
    c.idx0 = idx0;
    c.idx1 = idx1;
    c.k0 = k0;
    c.k1 = k1;
    c.pm0 = pm0;
    c.pm1 = pm1;
    c.base_normal = !PHILOX && (base_index & OB2_BASE_NORMAL_VARIATE) != 0;  // injected mode: base_index carries the flags
    Walk w;
    w.acc0 = w.acc1 = FpAcc{0, 0};
    w.all0 = w.all1 = true;
    w.col = cols + i0 * 8;
    w.tcol = WITH_T ? tcols + i0 * 8 : nullptr;
    w.r0 = w.r1 = Philox4{0, 0, 0, 0};

    if (!PHILOX) {
        const double *ucol = U + i0;
        for (int s = 0; s < n_slots; ++s) {
            double u0, u1;
            ld_f64x2(ucol, u0, u1);
            ucol += n_stride;
            generic_slot<false, WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>(s_slots, s_plan, s_cdf, s, u0, u1, w, c);
        }
    } else {
        for (int r = 0; r < n_runs; ++r) {
            const int first = s_runs[r].first, end = first + s_runs[r].count;
#define OB2_RUN_ARGS s_slots, s_plan, first, end, w, c
            switch (s_runs[r].cls) {
                case OB2_CLS_UNIFORM:
                    run_real<OB2_CLS_UNIFORM, WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>(OB2_RUN_ARGS);
                    break;
                case OB2_CLS_RECIP:
                    run_real<OB2_CLS_RECIP, WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>(OB2_RUN_ARGS);
                    break;
                case OB2_CLS_NORM:
                    run_real<OB2_CLS_NORM, WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>(OB2_RUN_ARGS);
                    break;
                case OB2_CLS_INT_MULHI:
                    run_words<OB2_CLS_INT_MULHI, WITH_FP, WITH_T, WITH_B, ROUNDS>(OB2_RUN_ARGS, s_thr);
                    break;
                case OB2_CLS_CAT_MULHI:
                    run_words<OB2_CLS_CAT_MULHI, WITH_FP, WITH_T, WITH_B, ROUNDS>(OB2_RUN_ARGS, s_thr);
                    break;
                case OB2_CLS_CAT_TABLE:
                    run_words<OB2_CLS_CAT_TABLE, WITH_FP, WITH_T, WITH_B, ROUNDS>(OB2_RUN_ARGS, s_thr);
                    break;
                default:
                    for (int s = first; s < end; ++s) {
                        double u0, u1;
                        if (s_plan[s].word == 0) {
                            new_block<ROUNDS>(s_plan[s].block, w, c);
                            u0 = u53(w.r0.x, w.r0.y);
                            u1 = u53(w.r1.x, w.r1.y);
                        } else {
                            u0 = u53(w.r0.z, w.r0.w);
                            u1 = u53(w.r1.z, w.r1.w);
                        }
                        generic_slot<true, WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>(s_slots, s_plan, s_cdf, s, u0, u1, w, c);
                    }
                    break;
            }
#undef OB2_RUN_ARGS
        }
    }
    FpAcc acc0 = w.acc0, acc1 = w.acc1;
    bool all0 = w.all0, all1 = w.all1 || !has1;
    if (PHILOX && !(all0 & all1)) {  // cold: some dimension of this pair missed its bounds on the first attempt
        if (!all0) {
            const Fixed fx = fixup_candidate<WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>(s_slots, s_plan, n_slots, idx0, i0, n_stride, cols,
                                                                                   tcols, bdelta, k0, k1, pm0, pm1, acc0, n_failed);
            acc0 = fx.acc;
            all0 = fx.ok;
        }
        if (!all1) {
            const Fixed fx = fixup_candidate<WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>(s_slots, s_plan, n_slots, idx1, i0 + 1, n_stride,
                                                                                   cols, tcols, bdelta, k0, k1, pm0, pm1, acc1, n_failed);
            acc1 = fx.acc;
            all1 = fx.ok;
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

// shared-memory layout of the staged tables: [slots][plan][runs][cdf][thr]
struct StagedTables {
    ob2_slot_desc *slots;
    ob2_slot_plan *plan;
    ob2_run *runs;
    double *cdf;
    uint32_t *thr;
};
__device__ __forceinline__ StagedTables carve_tables(unsigned char *smem, int n_slots, int cdf_len) {
    StagedTables t;
    t.slots = reinterpret_cast<ob2_slot_desc *>(smem);
    t.plan = reinterpret_cast<ob2_slot_plan *>(t.slots + n_slots);
    t.runs = reinterpret_cast<ob2_run *>(t.plan + n_slots);
    t.cdf = reinterpret_cast<double *>(t.runs + n_slots);
    t.thr = reinterpret_cast<uint32_t *>(t.cdf + cdf_len);
    return t;
}

template <bool PHILOX, bool WITH_FP, bool WITH_T, bool WITH_B, bool EXT, int ROUNDS>
__global__ void __launch_bounds__(OB2_BLOCK, WITH_B ? OB2_SUGGEST_CTAS : 4)
sample_kernel(const ob2_slot_desc *__restrict__ g_slots, const ob2_slot_plan *__restrict__ g_plan,
              const ob2_run *__restrict__ g_runs, int n_runs, int n_slots, const double *__restrict__ g_cdf,
              const uint32_t *__restrict__ g_thr, int cdf_len, int64_t n, int64_t n_stride, int64_t stride4,
              int64_t stride8, uint32_t k0, uint32_t k1, uint32_t pm0, uint32_t pm1, uint64_t base_index,
              const uint64_t *__restrict__ index, const double *__restrict__ U, char *__restrict__ cols,
              char *__restrict__ tcols, int64_t bdelta, uint8_t *__restrict__ mask, uint64_t *__restrict__ fp,
              uint32_t *__restrict__ n_failed) {
    extern __shared__ __align__(16) unsigned char smem[];
    const StagedTables t = carve_tables(smem, n_slots, cdf_len);
    // descriptors (128 B) and plan entries (16 B) are 16-byte multiples: copy as uint4
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(g_slots);
        uint4 *dst = reinterpret_cast<uint4 *>(t.slots);
        for (int i = threadIdx.x; i < n_slots * 8; i += blockDim.x) dst[i] = __ldg(src + i);
        const uint4 *psrc = reinterpret_cast<const uint4 *>(g_plan);
        uint4 *pdst = reinterpret_cast<uint4 *>(t.plan);
        for (int i = threadIdx.x; i < n_slots; i += blockDim.x) pdst[i] = __ldg(psrc + i);
        const uint2 *rsrc = reinterpret_cast<const uint2 *>(g_runs);
        uint2 *rdst = reinterpret_cast<uint2 *>(t.runs);
        for (int i = threadIdx.x; i < n_runs; i += blockDim.x) rdst[i] = __ldg(rsrc + i);
        for (int i = threadIdx.x; i < cdf_len; i += blockDim.x) {
            t.cdf[i] = __ldg(g_cdf + i);
            t.thr[i] = __ldg(g_thr + i);
        }
        __syncthreads();
    }
    sample_body<PHILOX, WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>(t.slots, t.plan, t.runs, n_runs, t.cdf, t.thr, n_slots, n, n_stride,
                                                              stride4, stride8, k0, k1, pm0, pm1, base_index, index, U, cols,
                                                              tcols, bdelta, mask, fp, n_failed);
}

// The same kernel for spaces of up to OB2_PARAM_SLOTS slots: descriptors, plan and runs travel in the
// launch parameters (constant bank); only the categorical tables are staged in shared memory.
struct ob2_slot_table {
    ob2_slot_desc s[OB2_PARAM_SLOTS];
    ob2_slot_plan p[OB2_PARAM_SLOTS];
    ob2_run r[OB2_PARAM_SLOTS];
};
static_assert(sizeof(ob2_slot_table) <= 30 * 1024, "the table must fit the 32 KB launch parameter space");

template <bool PHILOX, bool WITH_FP, bool WITH_T, bool WITH_B, bool EXT, int ROUNDS>
__global__ void __launch_bounds__(OB2_BLOCK, WITH_B ? OB2_SUGGEST_CTAS : 4)
sample_kernel_p(const __grid_constant__ ob2_slot_table tbl, int n_runs, int n_slots, const double *__restrict__ g_cdf,
                const uint32_t *__restrict__ g_thr, int cdf_len, int64_t n, int64_t n_stride, int64_t stride4,
                int64_t stride8, uint32_t k0, uint32_t k1, uint32_t pm0, uint32_t pm1, uint64_t base_index,
                const uint64_t *__restrict__ index, const double *__restrict__ U, char *__restrict__ cols,
                char *__restrict__ tcols, int64_t bdelta, uint8_t *__restrict__ mask, uint64_t *__restrict__ fp,
                uint32_t *__restrict__ n_failed) {
    extern __shared__ __align__(16) unsigned char smem[];
    double *s_cdf = reinterpret_cast<double *>(smem);
    uint32_t *s_thr = reinterpret_cast<uint32_t *>(s_cdf + cdf_len);
    for (int i = threadIdx.x; i < cdf_len; i += blockDim.x) {
        s_cdf[i] = g_cdf[i];
        s_thr[i] = g_thr[i];
    }
    __syncthreads();
    sample_body<PHILOX, WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>(tbl.s, tbl.p, tbl.r, n_runs, s_cdf, s_thr, n_slots, n, n_stride,
                                                              stride4, stride8, k0, k1, pm0, pm1, base_index, index, U, cols,
                                                              tcols, bdelta, mask, fp, n_failed);
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

// Which instantiations exist (each is a large kernel; the build time is proportional to their number):
//   Philox mode, tables in the launch parameters (spaces of up to OB2_PARAM_SLOTS slots): 5 output
//     combinations x {10 rounds, 7 rounds, 10 rounds with the further priors};
//   Philox mode, tables staged in shared memory (larger spaces): the 5 combinations x {common, further priors};
//   injected uniforms (parity mode): shared-memory staging, with / without the fingerprint, x the same two.
// EXT = the space uses one of the further priors: their inverse CDFs sit behind an out-of-line call, and a
// call anywhere in the walk costs the hot loops their register allocation, so the common kernels have none.
template <bool WITH_FP, bool WITH_T, bool WITH_B, bool EXT, int ROUNDS>
int launch_param(const SampleArgs &a, unsigned blocks) {
    const ob2_space *sp = a.sp;
    const int64_t bdelta = WITH_B ? (const char *)a.back - (const char *)a.cols : 0;
    auto kern_p = sample_kernel_p<true, WITH_FP, WITH_T, WITH_B, EXT, ROUNDS>;
    static thread_local ob2_slot_table tbl;  // (29 KB: not on the stack)
    memcpy(tbl.s, sp->h_slots, (size_t)sp->n_slots * sizeof(ob2_slot_desc));
    memcpy(tbl.p, sp->h_plan, (size_t)sp->n_slots * sizeof(ob2_slot_plan));
    memcpy(tbl.r, sp->h_runs, (size_t)sp->n_runs * sizeof(ob2_run));
    const size_t smem_p = (size_t)(sp->cdf_len > 0 ? sp->cdf_len : 1) * (sizeof(double) + sizeof(uint32_t));
    int rc_p = prepare_smem(kern_p, smem_p);
    if (rc_p != OB2_OK) return rc_p;
    kern_p<<<blocks, OB2_BLOCK, smem_p, static_cast<cudaStream_t>(a.stream)>>>(
        tbl, sp->n_runs, sp->n_slots, sp->d_cdf, sp->d_thr, sp->cdf_len, a.n, a.n_stride, 4 * a.n_stride, 8 * a.n_stride,
        (uint32_t)a.seed, (uint32_t)(a.seed >> 32), 0xD2511F53u, 0xCD9E8D57u, a.base_index, a.index, a.U, (char *)a.cols,
        (char *)a.tcols, bdelta, a.mask, a.fp, a.n_failed);
    return ob2_check_cuda(cudaGetLastError(), "sample_kernel_p launch");
}

template <bool PHILOX, bool WITH_FP, bool WITH_T, bool WITH_B, bool EXT>
int launch_smem(const SampleArgs &a, unsigned blocks) {
    const ob2_space *sp = a.sp;
    const int64_t bdelta = WITH_B ? (const char *)a.back - (const char *)a.cols : 0;
    auto kern = sample_kernel<PHILOX, WITH_FP, WITH_T, WITH_B, EXT, 10>;
    int rc = prepare_smem(kern, sp->smem_bytes);
    if (rc != OB2_OK) return rc;
    kern<<<blocks, OB2_BLOCK, sp->smem_bytes, static_cast<cudaStream_t>(a.stream)>>>(
        sp->d_slots, sp->d_plan, sp->d_runs, sp->n_runs, sp->n_slots, sp->d_cdf, sp->d_thr, sp->cdf_len, a.n, a.n_stride,
        4 * a.n_stride, 8 * a.n_stride, (uint32_t)a.seed, (uint32_t)(a.seed >> 32), 0xD2511F53u, 0xCD9E8D57u, a.base_index,
        a.index, a.U, (char *)a.cols, (char *)a.tcols, bdelta, a.mask, a.fp, a.n_failed);
    return ob2_check_cuda(cudaGetLastError(), "sample_kernel launch");
}

template <bool WITH_FP, bool WITH_T, bool WITH_B>
int launch_philox(const SampleArgs &a, unsigned blocks) {
    const ob2_space *sp = a.sp;
    const bool ext = sp->has_extended_prior;  // a slot uses one of the further priors (draw_extended)
    if (sp->philox_rounds != 10 && (ext || sp->n_slots > OB2_PARAM_SLOTS)) {
        ob2_set_error("sample: %d-round Philox is built for spaces of up to %d slots over the uniform / loguniform / "
                      "normal priors only", sp->philox_rounds, (int)OB2_PARAM_SLOTS);
        return OB2_EINVAL;
    }
    if (sp->n_slots <= OB2_PARAM_SLOTS && sp->h_slots != nullptr) {
        // Tables in the launch parameters: descriptor / plan / run fields become uniform loads from the
        // constant bank and the class branches are provably warp-uniform.
        if (ext) return launch_param<WITH_FP, WITH_T, WITH_B, true, 10>(a, blocks);
        if (sp->philox_rounds == 7) return launch_param<WITH_FP, WITH_T, WITH_B, false, 7>(a, blocks);
        return launch_param<WITH_FP, WITH_T, WITH_B, false, 10>(a, blocks);
    }
    return ext ? launch_smem<true, WITH_FP, WITH_T, WITH_B, true>(a, blocks)
               : launch_smem<true, WITH_FP, WITH_T, WITH_B, false>(a, blocks);
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
    if (!PHILOX) {  // parity mode: ob2_sample_from_uniform (no transformed / handed-out outputs)
        if (a.sp->has_extended_prior)
            return with_fp ? launch_smem<false, true, false, false, true>(a, b) : launch_smem<false, false, false, false, true>(a, b);
        return with_fp ? launch_smem<false, true, false, false, false>(a, b) : launch_smem<false, false, false, false, false>(a, b);
    }
    if (a.back != nullptr) return launch_philox<true, true, true>(a, b);  // ob2_suggest: everything in the one pass
    if (with_fp && with_t) return launch_philox<true, true, false>(a, b);
    if (with_fp) return launch_philox<true, false, false>(a, b);
    if (with_t) return launch_philox<false, true, false>(a, b);
    return launch_philox<false, false, false>(a, b);
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

extern "C" int ob2_sample_from_base(const ob2_space *space, int64_t n, int64_t n_stride, const double *B, uint32_t flags,
                                    void *cols, uint8_t *ok, uint64_t *fp, void *stream) {
    if (flags & ~(uint32_t)OB2_BASE_NORMAL_VARIATE) {
        ob2_set_error("ob2_sample_from_base: unknown flag bits 0x%x", flags);
        return OB2_EINVAL;
    }
    // the injected-mode kernels have no Philox counter: the base_index argument carries the flags
    const SampleArgs a{space, n, n_stride, 0, flags, nullptr, B, cols, nullptr, nullptr, ok, fp, nullptr, stream};
    return launch_sample<false>(a);
}

// Space.sample(n, seed) for a handful of points on the caller's own random stream (see ob2_sample_from_base):
// numbers in, kernel, columns and row verdicts out, one synchronisation.
extern "C" int ob2_sample_from_base_host(ob2_space *space, int64_t n, const double *host_base, uint32_t flags,
                                         void *host_cols, uint8_t *host_ok, void *stream) {
    if (space == nullptr || host_base == nullptr || host_cols == nullptr || host_ok == nullptr) {
        ob2_set_error("ob2_sample_from_base_host: null argument");
        return OB2_EINVAL;
    }
    if (n < 0 || n > OB2_HOST_SAMPLE_MAX) {
        ob2_set_error("ob2_sample_from_base_host: n = %lld outside [0, %d]; use ob2_sample_from_base for bulk draws",
                      (long long)n, (int)OB2_HOST_SAMPLE_MAX);
        return OB2_ETOOBIG;
    }
    if (n == 0) return OB2_OK;
    const int64_t n_stride = (n + 31) / 32 * 32;
    int rc = ob2_check_batch(space, n, n_stride);
    if (rc != OB2_OK) return rc;
    std::lock_guard<std::mutex> lock(*space->stage_mutex);
    rc = ensure_stage(space);
    if (rc != OB2_OK) return rc;
    if (space->d_base == nullptr) {
        const size_t bytes = (size_t)space->n_slots * sizeof(double) * OB2_HOST_SAMPLE_MAX;
        OB2_CUDA(cudaMalloc(&space->d_base, bytes));
        rc = ob2_check_cuda(cudaMallocHost(&space->h_base, bytes), "cudaMallocHost(base)");
        if (rc != OB2_OK) {
            cudaFree(space->d_base);
            space->d_base = nullptr;
            return rc;
        }
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char *d = static_cast<char *>(space->d_stage), *h = static_cast<char *>(space->h_stage);
    const size_t used = (size_t)space->row_bytes * n_stride;
    const size_t base_bytes = (size_t)space->n_slots * sizeof(double) * n_stride;
    memcpy(space->h_base, host_base, base_bytes);
    OB2_CUDA(cudaMemcpyAsync(space->d_base, space->h_base, base_bytes, cudaMemcpyHostToDevice, st));
    rc = ob2_sample_from_base(space, n, n_stride, static_cast<const double *>(space->d_base), flags, d,
                              reinterpret_cast<uint8_t *>(d + used), nullptr, st);
    if (rc != OB2_OK) return rc;
    OB2_CUDA(cudaMemcpyAsync(h, d, used + (size_t)n_stride, cudaMemcpyDeviceToHost, st));  // the verdicts sit behind the columns
    OB2_CUDA(cudaStreamSynchronize(st));
    memcpy(host_cols, h, used);
    memcpy(host_ok, h + used, (size_t)n);
    return OB2_OK;
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
