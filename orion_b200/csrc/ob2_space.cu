// Space handle: validates the slot table, uploads it (and the categorical cdf
// tables) once, and keeps the totals the launch wrappers need.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <new>

#include "ob2_internal.cuh"

static_assert(sizeof(ob2_slot_desc) == 128, "ob2_slot_desc must stay 128 bytes (staged as uint4)");

static thread_local char g_err[512] = "";

void ob2_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int ob2_check_cuda(cudaError_t err, const char *what) {
    if (err == cudaSuccess) return OB2_OK;
    ob2_set_error("CUDA error %d (%s) in %s", (int)err, cudaGetErrorString(err), what);
    return OB2_ECUDA;
}

extern "C" const char *ob2_last_error(void) { return g_err; }
extern "C" int ob2_version(void) { return OB2_VERSION; }

static int elem_bytes_orig(const ob2_slot_desc &d) { return d.kind == OB2_KIND_CATEGORICAL ? 4 : 8; }

extern "C" int ob2_space_create(const ob2_slot_desc *slots, int32_t n_slots, const double *cdf,
                                int32_t cdf_len, ob2_space **out) {
    if (out == nullptr || slots == nullptr || n_slots <= 0 || cdf_len < 0 || (cdf_len > 0 && cdf == nullptr)) {
        ob2_set_error("ob2_space_create: bad arguments (n_slots=%d cdf_len=%d)", n_slots, cdf_len);
        return OB2_EINVAL;
    }
    *out = nullptr;
    int64_t units = 0, t_units = 0;
    int32_t t_cols = 0;
    bool has_onehot = false;
    int unsamplable = -1;
    bool extended = false;
    for (int s = 0; s < n_slots; ++s) {
        const ob2_slot_desc &d = slots[s];
        if (d.kind < OB2_KIND_REAL || d.kind > OB2_KIND_CATEGORICAL) {
            ob2_set_error("slot %d: unknown kind %d", s, d.kind);
            return OB2_EINVAL;
        }
        if (d.kind != OB2_KIND_CATEGORICAL && (d.dist < OB2_DIST_UNIFORM || d.dist > OB2_DIST_LAST)) {
            ob2_set_error("slot %d: prior %d has no device inverse CDF", s, d.dist);
            return OB2_EINVAL;
        }
        if (d.kind == OB2_KIND_CATEGORICAL &&
            (d.n_cats <= 0 || d.cdf_off < 0 || (int64_t)d.cdf_off + d.n_cats > cdf_len)) {
            ob2_set_error("slot %d: categorical cdf [%d, %d) outside the table of %d", s, d.cdf_off,
                          d.cdf_off + d.n_cats, cdf_len);
            return OB2_EINVAL;
        }
        if (d.kind != OB2_KIND_CATEGORICAL && d.dist == OB2_DIST_NONE && unsamplable < 0) unsamplable = s;
        if (d.kind != OB2_KIND_CATEGORICAL && d.dist > OB2_DIST_NONE) extended = true;
        if (d.dim_first < 0 || d.dim_first > s) {
            ob2_set_error("slot %d: dim_first %d out of order", s, d.dim_first);
            return OB2_EINVAL;
        }
        if (d.col_units != units || d.t_units != t_units) {
            ob2_set_error("slot %d: col_units/t_units (%d, %d) do not match the running sums (%lld, %lld)", s,
                          d.col_units, d.t_units, (long long)units, (long long)t_units);
            return OB2_EINVAL;
        }
        int t_elem = 8, width = 1;
        switch (d.t_op) {
            case OB2_OP_IDENTITY:
                t_elem = elem_bytes_orig(d);
                break;
            case OB2_OP_QUANTIZE:
                if (d.kind != OB2_KIND_REAL) {
                    ob2_set_error("slot %d: Quantize applies to Real slots", s);
                    return OB2_EINVAL;
                }
                break;
            case OB2_OP_TO_REAL:
                if (d.kind != OB2_KIND_INTEGER) {
                    ob2_set_error("slot %d: Reverse(Quantize) applies to Integer slots", s);
                    return OB2_EINVAL;
                }
                break;
            case OB2_OP_ENUMERATE:
                if (d.kind != OB2_KIND_CATEGORICAL) {
                    ob2_set_error("slot %d: Enumerate applies to Categorical slots", s);
                    return OB2_EINVAL;
                }
                t_elem = 4;
                break;
            case OB2_OP_ONEHOT:
                if (d.kind != OB2_KIND_CATEGORICAL) {
                    ob2_set_error("slot %d: OneHotEncode applies to Categorical slots", s);
                    return OB2_EINVAL;
                }
                width = d.n_cats <= 2 ? 1 : d.n_cats;
                has_onehot = true;
                break;
            default:
                ob2_set_error("slot %d: unknown transform op %d", s, d.t_op);
                return OB2_EINVAL;
        }
        if (d.t_width != width) {
            ob2_set_error("slot %d: t_width %d, expected %d", s, d.t_width, width);
            return OB2_EINVAL;
        }
        units += elem_bytes_orig(d);
        t_units += (int64_t)t_elem * width;
        t_cols += width;
        if (units > INT32_MAX || t_units > INT32_MAX) {
            ob2_set_error("space too wide");
            return OB2_ETOOBIG;
        }
    }
    const size_t smem = (size_t)n_slots * sizeof(ob2_slot_desc) + (size_t)cdf_len * sizeof(double);
    if (smem > 200 * 1024) {
        ob2_set_error("space needs %zu bytes of descriptor staging; the limit is 200 KiB", smem);
        return OB2_ETOOBIG;
    }
    ob2_space *sp = new (std::nothrow) ob2_space();
    if (sp == nullptr) {
        ob2_set_error("out of host memory");
        return OB2_ENOMEM;
    }
    sp->n_slots = n_slots;
    sp->cdf_len = cdf_len;
    sp->row_bytes = units;
    sp->trow_bytes = t_units;
    sp->t_cols = t_cols;
    sp->smem_bytes = smem;
    sp->has_onehot = has_onehot;
    sp->unsamplable_slot = unsamplable;
    sp->has_extended_prior = extended;
    sp->d_slots = nullptr;
    sp->d_cdf = nullptr;
    int rc = ob2_check_cuda(cudaGetDevice(&sp->device), "cudaGetDevice");
    if (rc == OB2_OK) rc = ob2_check_cuda(cudaMalloc(&sp->d_slots, (size_t)n_slots * sizeof(ob2_slot_desc)), "cudaMalloc(slots)");
    if (rc == OB2_OK)
        rc = ob2_check_cuda(cudaMemcpy(sp->d_slots, slots, (size_t)n_slots * sizeof(ob2_slot_desc), cudaMemcpyHostToDevice),
                            "cudaMemcpy(slots)");
    // always allocate at least one element so kernels get a valid pointer
    if (rc == OB2_OK) rc = ob2_check_cuda(cudaMalloc(&sp->d_cdf, (size_t)(cdf_len > 0 ? cdf_len : 1) * sizeof(double)), "cudaMalloc(cdf)");
    if (rc == OB2_OK && cdf_len > 0)
        rc = ob2_check_cuda(cudaMemcpy(sp->d_cdf, cdf, (size_t)cdf_len * sizeof(double), cudaMemcpyHostToDevice), "cudaMemcpy(cdf)");
    if (rc != OB2_OK) {
        ob2_space_destroy(sp);
        return rc;
    }
    *out = sp;
    return OB2_OK;
}

extern "C" void ob2_space_destroy(ob2_space *sp) {
    if (sp == nullptr) return;
    if (sp->d_slots) cudaFree(sp->d_slots);
    if (sp->d_cdf) cudaFree(sp->d_cdf);
    delete sp;
}

extern "C" int64_t ob2_space_row_bytes(const ob2_space *sp) { return sp ? sp->row_bytes : -1; }
extern "C" int64_t ob2_space_trow_bytes(const ob2_space *sp) { return sp ? sp->trow_bytes : -1; }
