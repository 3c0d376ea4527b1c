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

// Element size of a slot in the wire layout: the narrowest integer type that holds every value
// of the slot's interval (Integer: signed, [low, high) inside [-2^(8w-1), 2^(8w-1)); Categorical:
// the index, unsigned for w < 4).  Real slots stay 8-byte doubles.
static int elem_bytes_wire(const ob2_slot_desc &d) {
    if (d.kind == OB2_KIND_REAL) return 8;
    if (d.kind == OB2_KIND_CATEGORICAL) return d.n_cats <= 256 ? 1 : d.n_cats <= 65536 ? 2 : 4;
    for (int w = 1; w <= 4; w *= 2) {
        const double lim = (double)(1ll << (8 * w - 1));
        if (d.low >= -lim && d.high <= lim) return w;  // (false for NaN / infinite bounds)
    }
    return 8;
}

// The sampler's draw plan (rules: ob2_internal.cuh).  `staged` is the library's copy of the table;
// `thr` receives the u32 thresholds of every categorical slot (cdf_len entries).
static int compile_plan(ob2_slot_desc *staged, int n_slots, const double *cdf, ob2_slot_plan *plan, ob2_run *runs,
                        uint32_t *thr) {
    int block = 0, word = 0, n_runs = 0;
    for (int s = 0; s < n_slots; ++s) {
        ob2_slot_desc &d = staged[s];
        ob2_slot_plan &p = plan[s];
        p.irange = 0;
        p.iloc = 0;
        p.thr_off = (uint32_t)d.cdf_off;
        int cls = OB2_CLS_GENERIC;
        if (d.kind == OB2_KIND_CATEGORICAL) {
            bool fits = true, exact = true;
            for (int j = 0; j + 1 < d.n_cats; ++j) {
                const double v = ceil(cdf[d.cdf_off + j] * 4294967296.0);
                if (!(v >= 0.0) || v > 4294967295.0) {
                    fits = false;
                    break;
                }
                thr[d.cdf_off + j] = (uint32_t)v;
                const uint64_t even = ((((uint64_t)(j + 1)) << 32) + (uint64_t)d.n_cats - 1) / (uint64_t)d.n_cats;
                exact = exact && even == (uint64_t)v;
            }
            if (d.n_cats >= 1) thr[d.cdf_off + d.n_cats - 1] = 0xFFFFFFFFu;  // (never read: cdf = 1 > u)
            if (fits) cls = exact ? OB2_CLS_CAT_MULHI : OB2_CLS_CAT_TABLE;
        } else if (d.kind == OB2_KIND_INTEGER) {
            if (d.dist == OB2_DIST_UNIFORM && d.scale >= 1.0 && d.scale <= 1048576.0 && d.scale == floor(d.scale) &&
                d.loc == floor(d.loc) && fabs(d.loc) <= 1048576.0 && d.low <= d.loc && d.high >= d.loc + d.scale) {
                cls = OB2_CLS_INT_MULHI;
                p.irange = (uint32_t)d.scale;
                p.iloc = (int32_t)d.loc;
            }
        } else if (d.dist == OB2_DIST_UNIFORM) {
            cls = OB2_CLS_UNIFORM;
        } else if (d.dist == OB2_DIST_NORM) {
            cls = OB2_CLS_NORM;
        }
        if (d.kind != OB2_KIND_CATEGORICAL && d.dist == OB2_DIST_RECIPROCAL && fabs(d.p0) + fabs(d.p1) < 700.0) {
            d.flags |= OB2_SLOT_FAST_EXP;
            if (d.kind == OB2_KIND_REAL) cls = OB2_CLS_RECIP;
        }
        const int need = (cls == OB2_CLS_INT_MULHI || cls == OB2_CLS_CAT_MULHI || cls == OB2_CLS_CAT_TABLE) ? 1 : 2;
        if (need == 2 && (word & 1)) ++word;
        if (word + need > 4) {
            ++block;
            word = 0;
        }
        if (block > 0xFFFF) return -1;
        p.block = (uint16_t)block;
        p.word = (uint8_t)word;
        p.cls = (uint8_t)cls;
        word += need;
        if (n_runs > 0 && runs[n_runs - 1].cls == (uint32_t)cls && runs[n_runs - 1].count < 0xFFFF) {
            ++runs[n_runs - 1].count;
        } else {
            runs[n_runs].first = (uint16_t)s;
            runs[n_runs].count = 1;
            runs[n_runs].cls = (uint32_t)cls;
            ++n_runs;
        }
    }
    return n_runs;
}

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
        // a dimension is a run of slots of one kind sharing dim_first, closed by OB2_SLOT_DIM_LAST
        // (the whole-dimension redraw walks [dim_first, last] and relies on exactly that)
        const bool opens = s == 0 || (slots[s - 1].flags & OB2_SLOT_DIM_LAST) != 0;
        if (opens ? d.dim_first != s : (d.dim_first != slots[s - 1].dim_first || d.kind != slots[s - 1].kind)) {
            ob2_set_error("slot %d: dim_first %d / kind %d inconsistent with the dimension it belongs to", s,
                          d.dim_first, d.kind);
            return OB2_EINVAL;
        }
        if (s == n_slots - 1 && !(d.flags & OB2_SLOT_DIM_LAST)) {
            ob2_set_error("slot %d: the last slot must close its dimension (OB2_SLOT_DIM_LAST)", s);
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
    // staged per block by the sampler: descriptors, draw plan, runs, cdf tables and their u32 thresholds
    const size_t smem = (size_t)n_slots * (sizeof(ob2_slot_desc) + sizeof(ob2_slot_plan) + sizeof(ob2_run)) +
                        (size_t)cdf_len * (sizeof(double) + sizeof(uint32_t));
    if (smem > 227 * 1024) {  // the most one thread block may allocate on sm_100
        ob2_set_error("space needs %zu bytes of descriptor staging; the limit is 227 KiB", smem);
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
    sp->h_plan = sp->d_plan = nullptr;
    sp->h_runs = sp->d_runs = nullptr;
    sp->d_thr = nullptr;
    sp->n_runs = 0;
    sp->philox_rounds = 10;
    sp->d_stage = sp->h_stage = nullptr;
    sp->d_base = sp->h_base = nullptr;
    sp->stage_mutex = new (std::nothrow) std::mutex();
    // the device copy of the table carries the wire layout in `reserved`: (offset units << 4) | element bytes
    ob2_slot_desc *staged = new (std::nothrow) ob2_slot_desc[n_slots];
    sp->wire_units = new (std::nothrow) int32_t[n_slots];
    sp->wire_bytes = new (std::nothrow) int32_t[n_slots];
    if (staged == nullptr || sp->wire_units == nullptr || sp->wire_bytes == nullptr || sp->stage_mutex == nullptr) {
        delete[] staged;
        ob2_space_destroy(sp);
        ob2_set_error("out of host memory");
        return OB2_ENOMEM;
    }
    int64_t wire = 0;
    for (int s = 0; s < n_slots; ++s) {
        staged[s] = slots[s];
        const int w = elem_bytes_wire(slots[s]);
        sp->wire_units[s] = (int32_t)wire;
        sp->wire_bytes[s] = w;
        staged[s].reserved = (int32_t)((wire << 4) | w);  // wire <= row_bytes < 2^27 (smem limit on n_slots)
        wire += w;
    }
    sp->wire_row_bytes = wire;
    sp->h_slots = staged;  // kept: small spaces travel in the sampler's launch parameters
    // the sampler's draw plan (sets OB2_SLOT_FAST_EXP in the staged flags: before the upload)
    sp->h_plan = new (std::nothrow) ob2_slot_plan[n_slots];
    sp->h_runs = new (std::nothrow) ob2_run[n_slots];
    uint32_t *h_thr = new (std::nothrow) uint32_t[cdf_len > 0 ? cdf_len : 1]();
    if (sp->h_plan == nullptr || sp->h_runs == nullptr || h_thr == nullptr) {
        delete[] h_thr;
        ob2_space_destroy(sp);
        ob2_set_error("out of host memory");
        return OB2_ENOMEM;
    }
    sp->n_runs = compile_plan(staged, n_slots, cdf, sp->h_plan, sp->h_runs, h_thr);
    if (sp->n_runs < 0) {
        delete[] h_thr;
        ob2_space_destroy(sp);
        ob2_set_error("space needs more than 65536 Philox blocks per candidate");
        return OB2_ETOOBIG;
    }
    int rc = ob2_check_cuda(cudaGetDevice(&sp->device), "cudaGetDevice");
    if (rc == OB2_OK) rc = ob2_check_cuda(cudaMalloc(&sp->d_plan, (size_t)n_slots * sizeof(ob2_slot_plan)), "cudaMalloc(plan)");
    if (rc == OB2_OK)
        rc = ob2_check_cuda(cudaMemcpy(sp->d_plan, sp->h_plan, (size_t)n_slots * sizeof(ob2_slot_plan), cudaMemcpyHostToDevice), "cudaMemcpy(plan)");
    if (rc == OB2_OK) rc = ob2_check_cuda(cudaMalloc(&sp->d_runs, (size_t)n_slots * sizeof(ob2_run)), "cudaMalloc(runs)");
    if (rc == OB2_OK)
        rc = ob2_check_cuda(cudaMemcpy(sp->d_runs, sp->h_runs, (size_t)sp->n_runs * sizeof(ob2_run), cudaMemcpyHostToDevice), "cudaMemcpy(runs)");
    if (rc == OB2_OK) rc = ob2_check_cuda(cudaMalloc(&sp->d_thr, (size_t)(cdf_len > 0 ? cdf_len : 1) * sizeof(uint32_t)), "cudaMalloc(thr)");
    if (rc == OB2_OK)
        rc = ob2_check_cuda(cudaMemcpy(sp->d_thr, h_thr, (size_t)(cdf_len > 0 ? cdf_len : 1) * sizeof(uint32_t), cudaMemcpyHostToDevice), "cudaMemcpy(thr)");
    delete[] h_thr;
    if (rc == OB2_OK) rc = ob2_check_cuda(cudaMalloc(&sp->d_slots, (size_t)n_slots * sizeof(ob2_slot_desc)), "cudaMalloc(slots)");
    if (rc == OB2_OK)
        rc = ob2_check_cuda(cudaMemcpy(sp->d_slots, staged, (size_t)n_slots * sizeof(ob2_slot_desc), cudaMemcpyHostToDevice),
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
    if (sp->d_plan) cudaFree(sp->d_plan);
    if (sp->d_runs) cudaFree(sp->d_runs);
    if (sp->d_thr) cudaFree(sp->d_thr);
    delete[] sp->h_plan;
    delete[] sp->h_runs;
    if (sp->d_stage) cudaFree(sp->d_stage);
    if (sp->h_stage) cudaFreeHost(sp->h_stage);
    if (sp->d_base) cudaFree(sp->d_base);
    if (sp->h_base) cudaFreeHost(sp->h_base);
    delete sp->stage_mutex;
    delete[] sp->h_slots;
    delete[] sp->wire_units;
    delete[] sp->wire_bytes;
    delete sp;
}

extern "C" int ob2_space_set_philox_rounds(ob2_space *sp, int32_t rounds) {
    if (sp == nullptr || (rounds != 7 && rounds != 10)) {
        ob2_set_error("ob2_space_set_philox_rounds: rounds must be 7 or 10");
        return OB2_EINVAL;
    }
    sp->philox_rounds = rounds;
    return OB2_OK;
}

extern "C" int64_t ob2_space_row_bytes(const ob2_space *sp) { return sp ? sp->row_bytes : -1; }
extern "C" int64_t ob2_space_trow_bytes(const ob2_space *sp) { return sp ? sp->trow_bytes : -1; }
extern "C" int64_t ob2_space_wire_row_bytes(const ob2_space *sp) { return sp ? sp->wire_row_bytes : -1; }

extern "C" int ob2_space_wire_layout(const ob2_space *sp, int32_t *units, int32_t *elem_bytes) {
    if (sp == nullptr || units == nullptr || elem_bytes == nullptr) {
        ob2_set_error("ob2_space_wire_layout: null argument");
        return OB2_EINVAL;
    }
    memcpy(units, sp->wire_units, (size_t)sp->n_slots * sizeof(int32_t));
    memcpy(elem_bytes, sp->wire_bytes, (size_t)sp->n_slots * sizeof(int32_t));
    return OB2_OK;
}
