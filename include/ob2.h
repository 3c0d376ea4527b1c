/* ob2.h -- C ABI of libob2: B200-native kernels for Orion's candidate-generation
 * hot path (sample -> contains -> transform/reverse -> fingerprint -> dedup).
 *
 * The reference (DatCorno/orion) is pure Python and has NO FFI of its own; the
 * boundary a maintainer binds is therefore new, and each entry point below
 * cites the reference function it stands in for (paths relative to the
 * reference checkout).  INTEGRATION.md shows the ctypes stub.
 *
 * Conventions
 *   - every function returns 0 on success, a negative OB2_E* code otherwise;
 *     ob2_last_error() gives the message for the calling thread.
 *   - all data pointers are DEVICE pointers owned by the caller; the library
 *     never allocates caller-visible memory.  Handles own only their small
 *     descriptor tables.  The exceptions are the three *_host entry points
 *     (small calls with HOST buffers in and out, synchronous): they stage
 *     through a block the space handle allocates on first use.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default
 *     stream).  Calls are asynchronous on that stream unless stated.
 *   - column sets: one contiguous block per batch; the column of scalar slot s
 *     starts at byte offset  col_units[s] * n_stride  where col_units is the
 *     running sum of element sizes (8 for Real/Integer, 4 for Categorical) and
 *     n_stride >= n is a multiple of 32 (so every column is 128-byte aligned).
 *     Per-candidate arrays (mask, keep, fingerprints, U) are n_stride long.
 *   - alignment: column blocks (cols, tcols, back, out_cols), fingerprint arrays and U must be
 *     16-byte aligned (they are accessed 128 bits at a time), mask / ok arrays 2-byte aligned,
 *     index arrays 8-byte, counters 4-byte.  Anything else is refused with OB2_EINVAL -- a
 *     misaligned vector access on the device would poison the CUDA context instead.
 */
#ifndef OB2_H_
#define OB2_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OB2_VERSION 1

enum { OB2_OK = 0, OB2_EINVAL = -1, OB2_ECUDA = -2, OB2_ETOOBIG = -3, OB2_ENOMEM = -4 };

/* dimension kinds: Real / Integer / Categorical, src/orion/algo/space.py:199,322,350 */
enum { OB2_KIND_REAL = 0, OB2_KIND_INTEGER = 1, OB2_KIND_CATEGORICAL = 2 };

/* priors with a device inverse CDF (scipy names the reference's DSL maps to,
 * src/orion/core/io/space_builder.py:124-154).  CONST = scale == 0 (rvs returns loc).
 * NONE = any other scipy prior: the space can be checked, transformed, hashed
 * and deduplicated (only its bounds matter), but ob2_sample* refuse it. */
enum { OB2_DIST_UNIFORM = 0, OB2_DIST_RECIPROCAL = 1, OB2_DIST_NORM = 2, OB2_DIST_CONST = 3, OB2_DIST_NONE = 4,
       /* further scipy priors with a closed-form _ppf (any name DimensionBuilder._build accepts,
        * space_builder.py:156-186); p0 / p1 carry the shape parameter and its reciprocal */
       OB2_DIST_EXPON = 5, OB2_DIST_LOGNORM = 6, OB2_DIST_LOGISTIC = 7, OB2_DIST_LAPLACE = 8, OB2_DIST_GUMBEL_R = 9,
       OB2_DIST_WEIBULL_MIN = 10, OB2_DIST_PARETO = 11, OB2_DIST_RAYLEIGH = 12, OB2_DIST_HALFNORM = 13,
       OB2_DIST_POWERLAW = 14,
       /* the priors of the reference's own DSL examples (space_builder.py:84,98-99: `alpha(.9)`,
        * `randint(-3, 5)`, `poisson(mu=3)`).  ALPHA: p0 = a, p1 = ndtr(a).  RANDINT: p0 = low,
        * p1 = high - low, scipy's _ppf verbatim.  POISSON: p0 = mu (<= 500), p1 = exp(-mu); the
        * smallest k with cdf(k) >= u by summing the pmf. */
       OB2_DIST_ALPHA = 15, OB2_DIST_RANDINT = 16, OB2_DIST_POISSON = 17,
       /* GAMMA (also erlang): p0 = a, p1 = lgamma(a).  BETA: p0 = a, p1 = b, hk[3] = bit pattern of
        * the double lbeta(a, b).  Inverse of the regularised incomplete gamma / beta function by
        * Halley steps from a closed-form first guess (scipy: special.gammaincinv / betaincinv). */
       OB2_DIST_GAMMA = 18, OB2_DIST_BETA = 19,
       /* CAUCHY: tan(pi (u - 1/2)).  CHI2: p0 = df / 2, p1 = lgamma(df / 2), twice the gamma
        * inverse.  T: p0 = df, p1 = lbeta(df / 2, 1 / 2), through the beta inverse.  TRUNCNORM (a, b):
        * hk[3] > 0: ndtri(p0 + u p1) with p0 = Phi(a); hk[3] < 0 (a >= 0, scipy's right-hand
        * case): -ndtri(p0 + (1 - u) p1) with p0 = Phi(-b); p1 = Phi(b) - Phi(a) either way. */
       OB2_DIST_CAUCHY = 20, OB2_DIST_CHI2 = 21, OB2_DIST_T = 22, OB2_DIST_TRUNCNORM = 23,
       /* GEOM: p0 = p, p1 = log1p(-p), scipy's _ppf verbatim.  BINOM: p0 = n, p1 = p, hk[3] = bits of
        * (1 - p)^n (must not underflow); the smallest k with cdf(k) >= u by summing the pmf. */
       OB2_DIST_GEOM = 24, OB2_DIST_BINOM = 25,
       /* more closed forms, each scipy's own _ppf formula (scipy/stats/_continuous_distns.py); p0 / p1 = the
        * shape parameters in scipy's order unless said otherwise.  The gamma-based ones carry p1 = lgamma(p0):
        * CHI (also maxwell = chi(3)): p0 = df / 2, sqrt(2 gammaincinv);  INVGAMMA: p0 = a, 1 / gammainccinv;
        * NAKAGAMI: p0 = nu, sqrt(gammaincinv / nu);  DGAMMA: p0 = a, signed;  GENNORM: p0 = 1 / beta, p1 =
        * lgamma(1 / beta), hk[3] = bits of beta.  RDIST (also semicircular = rdist(3)): p0 = c / 2, hk[3] = bits of
        * lbeta(c / 2, c / 2), 2 betaincinv - 1.  FISK = BURR with d = 1. */
       OB2_DIST_GUMBEL_L = 26, OB2_DIST_GENPARETO = 27, OB2_DIST_LOMAX = 28, OB2_DIST_BURR = 29, OB2_DIST_BURR12 = 30,
       OB2_DIST_INVWEIBULL = 31, OB2_DIST_TRUNCEXPON = 32, OB2_DIST_BRADFORD = 33, OB2_DIST_ARCSINE = 34,
       OB2_DIST_TRIANG = 35, OB2_DIST_LOGLAPLACE = 36, OB2_DIST_EXPONPOW = 37, OB2_DIST_GOMPERTZ = 38,
       OB2_DIST_GENLOGISTIC = 39, OB2_DIST_HYPSECANT = 40, OB2_DIST_HALFCAUCHY = 41, OB2_DIST_HALFLOGISTIC = 42,
       OB2_DIST_CHI = 43, OB2_DIST_INVGAMMA = 44, OB2_DIST_NAKAGAMI = 45, OB2_DIST_GIBRAT = 46,
       OB2_DIST_GENEXTREME = 47, OB2_DIST_KAPPA3 = 48, OB2_DIST_DWEIBULL = 49, OB2_DIST_ANGLIT = 50,
       OB2_DIST_EXPONWEIB = 51, OB2_DIST_MIELKE = 52, OB2_DIST_GENHALFLOGISTIC = 53, OB2_DIST_TUKEYLAMBDA = 54,
       OB2_DIST_LEVY = 55, OB2_DIST_LEVY_L = 56, OB2_DIST_WEIBULL_MAX = 57, OB2_DIST_RDIST = 58, OB2_DIST_DGAMMA = 59,
       OB2_DIST_GENNORM = 60, OB2_DIST_MOYAL = 61, OB2_DIST_WRAPCAUCHY = 62,
       /* F: p0 = dfn, p1 = dfd, hk[3] = bits of lbeta(dfn / 2, dfd / 2); through the beta inverse, from whichever
        * tail is the small one.  LOGGAMMA: p0 = c, p1 = lgamma(c), hk[3] = bits of lgamma(c + 1) (scipy's
        * one-term form where gammaincinv underflows).  NBINOM: p0 = n, p1 = p, hk[3] = bits of p^n (must not
        * underflow); HYPERGEOM: p0 = M, p1 = n, hk[3] = bits of N: the smallest k with cdf(k) >= u by summing
        * the pmf from the lower end of the support.  VONMISES: p0 = kappa (< 50), p1 = 1 / (2 pi i0e(kappa));
        * safeguarded Newton steps on the cdf series scipy itself sums (Hill's algorithm 518). */
       OB2_DIST_F = 63, OB2_DIST_LOGGAMMA = 64, OB2_DIST_NBINOM = 65, OB2_DIST_HYPERGEOM = 66, OB2_DIST_VONMISES = 67,
       /* a third batch, again scipy's own _ppf formulas with the shape parameters in scipy's order (a third one in
        * hk[3]).  GENGAMMA: p0 = a, p1 = lgamma(a), hk[3] = bits of c.  PLANCK / DLAPLACE / BOLTZMANN: scipy's
        * ceil-then-correct forms, operation for operation (bernoulli = BINOM with n = 1). */
       OB2_DIST_JOHNSONSB = 68, OB2_DIST_JOHNSONSU = 69, OB2_DIST_FATIGUELIFE = 70, OB2_DIST_POWERNORM = 71,
       OB2_DIST_POWERLOGNORM = 72, OB2_DIST_GENGAMMA = 73, OB2_DIST_KAPPA4 = 74, OB2_DIST_LAPLACE_ASYMMETRIC = 75,
       OB2_DIST_TRAPEZOID = 76, OB2_DIST_TRUNCPARETO = 77, OB2_DIST_TRUNCWEIBULL_MIN = 78, OB2_DIST_SKEWCAUCHY = 79,
       OB2_DIST_PLANCK = 80, OB2_DIST_DLAPLACE = 81, OB2_DIST_BOLTZMANN = 82,
       OB2_DIST_LAST = 82 };

/* per-slot transformer (build_required_space table) */
enum {
    OB2_OP_IDENTITY = 0,  /* copy, dtype kept                                     */
    OB2_OP_QUANTIZE = 1,  /* Quantize: f64 -> i64 (round half-even, or floor)     */
    OB2_OP_TO_REAL = 2,   /* Reverse(Quantize): i64 -> f64                        */
    OB2_OP_ENUMERATE = 3, /* Enumerate: category index i32 -> i32                 */
    OB2_OP_ONEHOT = 4     /* Compose[Enumerate, OneHotEncode(k)]: i32 -> k x f64  */
};
#define OB2_TFLAG_FLOOR 1 /* Quantize rounds with floor instead of rint */
#define OB2_SLOT_DIM_LAST 1
#define OB2_SLOT_CAT_UNIFORM 2 /* categorical with equal probabilities (a hint: enables the guess-and-walk draw) */

/* One scalar slot of the space (a dimension with shape contributes prod(shape)
 * slots, C order).  Plain data, identical layout on host and device. */
typedef struct ob2_slot_desc {
    int32_t kind;      /* OB2_KIND_*                                              */
    int32_t dist;      /* OB2_DIST_* (unused for categorical)                     */
    int32_t dim_first; /* index of the first slot of this slot's dimension        */
    int32_t flags;     /* OB2_SLOT_DIM_LAST on the last slot of a dimension       */
    int32_t n_cats;    /* categorical: number of categories                       */
    int32_t cdf_off;   /* categorical: offset of its cdf in the cdf table         */
    int32_t col_units; /* byte offset of this column per unit of n_stride         */
    int32_t t_op;      /* OB2_OP_*                                                */
    int32_t t_units;   /* same as col_units, in the transformed column set        */
    int32_t t_flags;   /* OB2_TFLAG_*                                             */
    int32_t t_width;   /* number of transformed columns this slot expands to      */
    int32_t reserved;  /* ignored on input (the library keeps the wire layout here) */
    double p0, p1;     /* reciprocal: log(a), log(b) - log(a)                     */
    double loc, scale; /* x = base(u) * scale + loc  (scipy rv_generic.rvs)       */
    double low, high;  /* accepted iff low <= x < high (space.py:160-164)         */
    uint64_t hk[4];    /* fingerprint: xor key, 2 x packed odd multipliers; hk[3]: bits of a third prior parameter (BETA, TRUNCNORM) */
} ob2_slot_desc;

typedef struct ob2_space ob2_space;
typedef struct ob2_set ob2_set;

const char *ob2_last_error(void);
int ob2_version(void);

/* Compile a Space into device tables.  Replaces the per-call Python walk over
 * `Space.values()` (src/orion/algo/space.py:483) and the per-scalar
 * `interval()` recomputation (space.py:160).  `cdf` holds the categorical cdf
 * tables (cumsum(p) / cumsum(p)[-1], as numpy's legacy RandomState.choice
 * builds them, space.py:400).  Synchronous; uses the current device. */
int ob2_space_create(const ob2_slot_desc *slots, int32_t n_slots, const double *cdf,
                     int32_t cdf_len, ob2_space **out);
void ob2_space_destroy(ob2_space *space);
/* Philox round count of the sampler for this space: 10 (default, Philox4x32-10 as published) or 7
 * (Philox4x32-7, the smallest variant that passes BigCrush; ~5 % fewer instructions per
 * candidate).  Different rounds = different draws for the same seed.  No counterpart in the
 * reference (numpy's MT19937, src/orion/algo/space.py:482). */
int ob2_space_set_philox_rounds(ob2_space *space, int32_t rounds);
int64_t ob2_space_row_bytes(const ob2_space *space);  /* bytes per candidate, original set   */
int64_t ob2_space_trow_bytes(const ob2_space *space); /* bytes per candidate, transformed set */
/* The "wire" layout used to hand candidates to the host (ob2_compact_wire): same columnar
 * scheme, but every Integer / Categorical column is stored in the narrowest integer type
 * that holds the slot's whole interval (Integer: signed 1/2/4/8 bytes from [low, high);
 * Categorical index: unsigned 1/2 bytes, else int32); Real columns stay float64.  Lossless
 * for every row that is in the space.  units[s] * stride = byte offset of slot s's column,
 * elem_bytes[s] its element size; both are host arrays of n_slots entries. */
int64_t ob2_space_wire_row_bytes(const ob2_space *space);
int ob2_space_wire_layout(const ob2_space *space, int32_t *units, int32_t *elem_bytes);

/* Space.sample(n, seed) -- src/orion/algo/space.py:456-484, with Real.sample's
 * <=4-attempt rejection (space.py:256-276), _Discrete.sample's floor
 * (space.py:281-294) and Categorical.sample (space.py:392-402), as called by
 * Random.suggest (src/orion/algo/random.py:24-33).  Uniforms come from
 * Philox4x32: key = seed, counter (index_lo, index_hi, block, attempt) with
 * candidate index = base_index + i, or index[i] when `index` (device u64[n],
 * nullable) is given -- replaying chosen candidates from their counters.  Walking
 * the slots in order, a slot takes the next free words of the current block: two
 * (starting at an even word) for a 53-bit uniform ((w0 >> 5) 2^26 + (w1 >> 6)) / 2^53,
 * or ONE for a 32-bit uniform w / 2^32 where that is all the slot needs -- a
 * Categorical slot whose cdf thresholds ceil(cdf[j] 2^32) fit 32 bits, and an Integer
 * slot with a uniform prior, integral loc / scale, 1 <= scale <= 2^20, |loc| <= 2^20 and
 * no narrower user bounds (there floor(u scale + loc) is exact in integers).  A space
 * of Real slots only uses block = slot / 2.
 * `mask[i]` (nullable) is the fused
 * `point in space` of PrimaryAlgo.suggest (src/orion/core/worker/primary_algo.py:53-54);
 * `fp` (nullable, [n_stride][2] u64) receives the row fingerprints;
 * `n_failed` (device u32[n_slots], nullable, zeroed by the caller) counts, at
 * the index of a dimension's first slot, the candidates for which that
 * dimension missed its bounds 4 times (the reference's ValueError). */
int ob2_sample(const ob2_space *space, int64_t n, int64_t n_stride, uint64_t seed,
               uint64_t base_index, const uint64_t *index, void *cols, uint8_t *mask,
               uint64_t *fp, uint32_t *n_failed, void *stream);

/* TransformedSpace.sample = transform(original.sample(...)) in ONE kernel: the
 * same draw as ob2_sample, and each value is also written through its slot's
 * transformer into `tcols` (the transformed column set) while it is still in
 * registers, so the sampled columns are never re-read.  Same arguments as
 * ob2_sample plus `tcols`.  (Reference: TODO at
 * src/orion/core/worker/primary_algo.py:54-55) */
int ob2_sample_transform(const ob2_space *space, int64_t n, int64_t n_stride, uint64_t seed,
                         uint64_t base_index, const uint64_t *index, void *cols, void *tcols,
                         uint8_t *mask, uint64_t *fp, uint32_t *n_failed, void *stream);

/* Random.suggest behind PrimaryAlgo in ONE kernel (src/orion/algo/random.py:24-33 called
 * through src/orion/core/worker/primary_algo.py:44-56): the draw and transform of
 * ob2_sample_transform, plus reverse() of every transformed value while it is still in
 * registers -- written to `back` (the handed-out rows; same layout as `cols`, 16-byte
 * aligned; `back` == `cols` is allowed and means "keep only the handed-out rows": one store
 * per value instead of two) -- plus the fingerprint of the `back` rows into `fp` (required).
 * Results are bit-identical to ob2_sample_transform followed by ob2_reverse_fingerprint;
 * what is saved is re-reading the transformed columns (the widest stream of the path).
 * Algorithms that compute in the transformed space keep using the separate kernels. */
int ob2_suggest(const ob2_space *space, int64_t n, int64_t n_stride, uint64_t seed,
                uint64_t base_index, const uint64_t *index, void *cols, void *tcols, void *back,
                uint8_t *mask, uint64_t *fp, uint32_t *n_failed, void *stream);

/* Space.sample(n) for the small n the reference's own callers ask for (Producer.produce ->
 * suggest(pool_size), src/orion/core/worker/producer.py:45; DimensionBuilder's warm-up draw,
 * src/orion/core/io/space_builder.py:205): the same draw as ob2_sample, delivered into HOST
 * memory.  host_cols receives the column block with n_stride = n rounded up to 32
 * (ob2_space_row_bytes * n_stride bytes); host_n_failed the n_slots rejection counters.
 * Synchronous.  The first call on a space allocates a staging block (device + pinned host,
 * row_bytes * OB2_HOST_SAMPLE_MAX bytes each) that lives until ob2_space_destroy; concurrent
 * calls on one space serialise. */
#define OB2_HOST_SAMPLE_MAX 4096
int ob2_sample_host(ob2_space *space, int64_t n, uint64_t seed, uint64_t base_index, void *host_cols,
                    uint32_t *host_n_failed, void *stream);

/* ob2_sample_host plus the producer's dedup: the drawn rows are
 * fingerprinted and deduplicated against `set` (ob2_set_dedup semantics: kept rows join the
 * set); host_keep receives n verdicts.  One kernel sequence, one copy, one synchronisation. */
int ob2_sample_dedup_host(ob2_space *space, ob2_set *set, int64_t n, uint64_t seed, uint64_t base_index,
                          void *host_cols, uint8_t *host_keep, uint32_t *host_n_failed, void *stream);

/* Space.__contains__ for a handful of HOST points (PrimaryAlgo.suggest / observe / score /
 * judge check every point, src/orion/core/worker/primary_algo.py:53-54,64-66,78-84):
 * host_cols is a column block with n_stride = n rounded up to 32 (as ob2_contains takes it on
 * the device), host_mask receives n verdicts.  Same staging block and limits as ob2_sample_host. */
int ob2_contains_host(ob2_space *space, int64_t n, const void *host_cols, int32_t int_as_f64,
                      uint8_t *host_mask, void *stream);

/* Parity mode of the same kernel: uniforms are read from U[S][n_stride] (one
 * attempt, no redraw); ok[i] = every Real/Integer slot landed in bounds. */
int ob2_sample_from_uniform(const ob2_space *space, int64_t n, int64_t n_stride,
                            const double *U, void *cols, uint8_t *ok, uint64_t *fp,
                            void *stream);

/* The same with the caller's OWN random stream, so that a seeded Space.sample(n, seed) returns the
 * reference's values: B[S][n_stride] holds, per slot, what numpy's RandomState hands scipy for one
 * draw (src/orion/algo/space.py:133-136 -> rv_generic.rvs): a uniform of random_sample() for uniform /
 * reciprocal / Categorical slots (RandomState.choice, space.py:402-403) and, with
 * OB2_BASE_NORMAL_VARIATE in flags, a variate of standard_normal() for norm slots (the kernel then only
 * applies scale, loc, floor and the bounds check).  One attempt, no redraw: ok[i] tells the host which
 * rows the reference would have drawn again (Real.sample, space.py:262-274). */
#define OB2_BASE_NORMAL_VARIATE 1u
int ob2_sample_from_base(const ob2_space *space, int64_t n, int64_t n_stride,
                         const double *B, uint32_t flags, void *cols, uint8_t *ok, uint64_t *fp,
                         void *stream);

/* The same for a handful of points with HOST buffers (n <= OB2_HOST_SAMPLE_MAX; the seeded
 * Space.sample(n, seed) / Dimension.sample calls of the reference's users and tests): host_base is
 * [S][n_stride] with n_stride = n rounded up to 32, host_cols a column block of that stride, host_ok
 * n verdict bytes.  One copy in, one kernel, one copy out, one synchronisation. */
int ob2_sample_from_base_host(ob2_space *space, int64_t n, const double *host_base, uint32_t flags,
                              void *host_cols, uint8_t *host_ok, void *stream);

/* Space.__contains__ over N points -- src/orion/algo/space.py:521-545 over
 * Dimension.__contains__ :150-164, Integer.__contains__ :335-347 and
 * Categorical.__contains__ :415-426 (device: 0 <= index < n_cats).  With
 * int_as_f64 != 0 Integer columns hold float64 and `mod(x, 1) == 0` is checked. */
int ob2_contains(const ob2_space *space, int64_t n, int64_t n_stride, const void *cols,
                 int32_t int_as_f64, uint8_t *mask, void *stream);

/* TransformedSpace.transform / .reverse -- absent from the reference snapshot
 * (TODOs at src/orion/core/worker/primary_algo.py:13-14,42,55); contract in
 * SURVEY.md Appendix B.1/B.2. */
int ob2_transform(const ob2_space *space, int64_t n, int64_t n_stride, const void *cols,
                  void *tcols, void *stream);
int ob2_reverse(const ob2_space *space, int64_t n, int64_t n_stride, const void *tcols,
                void *cols, void *stream);
/* ob2_reverse that also fingerprints the rows it writes (fp[n_stride][2]), i.e.
 * the candidates as they are handed out.  The reverse is bandwidth bound with
 * idle issue slots, so this is where the dedup fingerprint is cheapest; under
 * the 'real' / 'numerical' / None requirements reverse(transform(x)) == x and it
 * equals the fingerprint ob2_sample would have produced. */
int ob2_reverse_fingerprint(const ob2_space *space, int64_t n, int64_t n_stride,
                            const void *tcols, void *cols, uint64_t *fp, void *stream);

/* 128-bit row fingerprint of existing columns (history ingestion); the trial
 * identity the producer dedups on.  No counterpart in the snapshot
 * (src/orion/core/worker/producer.py:42-52 registers blindly); SURVEY B.3. */
int ob2_fingerprint(const ob2_space *space, int64_t n, int64_t n_stride, const void *cols,
                    uint64_t *fp, void *stream);

/* History hash set over caller-owned memory: n_slots (power of two, <= 2^31) slots of
 * OB2_SET_SLOT_BYTES, preceded by an OB2_SET_HEADER_BYTES header.  create
 * zeroes the memory on `stream`. */
#define OB2_SET_SLOT_BYTES 32
#define OB2_SET_HEADER_BYTES 256
int ob2_set_create(void *mem, uint64_t n_slots, void *stream, ob2_set **out);
void ob2_set_destroy(ob2_set *set);
/* insert fingerprints (history load / observe); duplicates collapse */
int ob2_set_insert(ob2_set *set, int64_t n, const uint64_t *fp, void *stream);
/* keep[i] = fp[i] not in the set before this call AND no entry of this call
 * with a smaller order key has the same fingerprint (order = `order[i]`, or i
 * when NULL).  Entries with fp.lo == 0 are padding (keep = 0).  Kept
 * fingerprints stay in the set.  `scratch` is n u32 of caller scratch. */
int ob2_set_dedup(ob2_set *set, int64_t n, const uint64_t *fp, const uint32_t *order,
                  uint32_t *scratch, uint8_t *keep, void *stream);

/* How ob2_set_dedup (and the exchange's inbox dedup) lays out its first pass.  0 (default): one thread per
 * entry -- the fastest when the kernel has the GPU to itself (0.91 ms per 1.25e7 entries).  ctas > 0: a fixed
 * grid of that many CTAs walks the entries instead.  Alone that is slower (148 CTAs: 1.7 ms, 296: 1.2 ms), but a
 * small resident grid SHARES the SMs with a kernel running on another stream, where the big grid takes turns
 * with it: next to the issue-bound sampler of BASELINE.json's config 5 (ob2_sample with the fused fingerprint,
 * 2.30 ms) the step drops from 3.14 ms (the sum) to 2.48 ms with 148-222 CTAs.  Next to a kernel that already
 * saturates HBM (ob2_suggest's stores) there is nothing to share and the default stays best.  The verdicts do
 * not depend on the setting.  The environment variable OB2_DEDUP_GRID sets the initial value of new sets. */
int ob2_set_set_dedup_grid(ob2_set *set, int32_t ctas);
/* Same verdicts as ob2_set_dedup, for callers that cannot allocate device
 * memory (no counterpart in the reference, which has no device code at all).
 * Who that is: a plugin inside a process whose framework already owns all of
 * HBM through its own pool, so a cudaMalloc of its own would fail or fragment
 * it; a producer loop that must not stall, because cudaMalloc/cudaFree
 * synchronise the whole device; a minimal binding that only ever holds the
 * input and output buffers it was given.  The u32-per-entry scratch that
 * ob2_set_dedup takes from the caller is instead a fixed array inside the
 * library: OB2_STATIC_DEDUP_MAX * 4 bytes (4 MiB) that a GPU carries from the
 * moment the CUDA module with the hash-set kernels is loaded on it until the
 * process exits.  Under CUDA's default lazy loading that moment is the first
 * hash-set call on the GPU -- ob2_set_create, which also looks the array up --
 * so the 4 MiB are claimed at set-up time, not by a dedup call.  The call
 * itself allocates nothing (no cudaMalloc/cudaFree, no allocating library
 * routine, no module load once ob2_set_create has run on that device).
 *   - n > OB2_STATIC_DEDUP_MAX is refused with OB2_ETOOBIG: split the batch
 *     into consecutive calls (a row kept by an earlier call beats its copies
 *     in later ones), or use ob2_set_dedup.
 *   - the array's device address is looked up once per device
 *     (cudaGetSymbolAddress, with that device current) by ob2_set_create; a set
 *     created on another device than the one the call runs on is refused, so
 *     the look-up has always happened.  Create the set before any stream
 *     capture begins.
 *   - all calls share the one array per GPU, and it keeps whatever the last
 *     call wrote.  Nothing in it needs to start at zero (pass 1 builds the
 *     list pass 2 reads, from empty, every call).  Two calls that overlap on the GPU -- different
 *     streams, or different sets -- race on it: keep them ordered. */
#define OB2_STATIC_DEDUP_MAX (1 << 20)
int ob2_set_dedup_static(ob2_set *set, int64_t n, const uint64_t *fp, const uint32_t *order,
                         uint8_t *keep, void *stream);
int ob2_set_probe(const ob2_set *set, int64_t n, const uint64_t *fp, uint8_t *found,
                  void *stream);
/* Table growth: move every fingerprint of `src` into `dst`, a freshly created
 * (larger) set; they count as history for later dedup calls.  Synchronises
 * `stream`.  The reference rebuilds algorithm state by re-observing stored
 * trials (src/orion/core/worker/experiment.py:230-247); this avoids that. */
int ob2_set_rehash(const ob2_set *src, ob2_set *dst, void *stream);
/* Snapshot (checkpoint / resume): write the stored fingerprints, in no
 * particular order, to out_fp[max_out][2]; *n_out (device u64) receives their
 * number (entries beyond max_out are counted but not written). */
int ob2_set_export(const ob2_set *set, uint64_t max_out, uint64_t *out_fp, uint64_t *n_out,
                   void *stream);
/* number of distinct fingerprints stored; synchronises `stream` */
int ob2_set_count(const ob2_set *set, uint64_t *count, void *stream);

/* Order-preserving gather of the rows with keep[i] != 0 into a dense column
 * block of the same layout with stride out_stride; *n_kept (device u64)
 * receives their number.  The device half of Producer.produce's
 * suggest -> tuple_to_trial -> register_trials hand-over
 * (src/orion/core/worker/producer.py:45-51, src/orion/core/utils/format_trials.py:28-42):
 * only surviving candidates are copied to the host and turned into trials.
 * `transformed` != 0 compacts a transformed column set.  `scratch` is
 * ob2_compact_scratch_bytes(n) bytes of caller scratch. */
int64_t ob2_compact_scratch_bytes(int64_t n);
int ob2_compact(const ob2_space *space, int64_t n, int64_t n_stride, const void *cols,
                const uint8_t *keep, int32_t transformed, int64_t out_stride, void *out_cols,
                void *scratch, uint64_t *n_kept, void *stream);
/* The same gather of an ORIGINAL column set, written in the wire layout above: what crosses
 * PCIe on the way to tuple_to_trial (src/orion/core/utils/format_trials.py:28-42) is the
 * survivors' values at their narrowest lossless width (mixed64: 288 instead of 448 bytes
 * per candidate).  Rows that are not in the space (an "Improbable bounds" failure the host
 * raises on anyway) may be truncated. */
int ob2_compact_wire(const ob2_space *space, int64_t n, int64_t n_stride, const void *cols,
                     const uint8_t *keep, int64_t out_stride, void *out_cols, void *scratch,
                     uint64_t *n_kept, void *stream);

/* ... and their way to the host: copy the first n_rows of every column of a device column block
 * (layout 0 = original, 1 = transformed, 2 = wire; stride src_stride) into a HOST block of the same
 * layout with stride dst_stride, asynchronously on `stream` (pin the host block for a true async
 * copy).  Runs of adjacent equal-width columns go as one 2-D copy each, so a compacted batch of
 * k survivors crosses PCIe as k * row_bytes bytes in 2-3 transfers -- what
 * Experiment.register_trials then turns into trial documents
 * (src/orion/core/worker/experiment.py:216-228). */
int ob2_columns_to_host(const ob2_space *space, int32_t layout, int64_t n_rows, int64_t src_stride,
                        const void *src_cols, int64_t dst_stride, void *dst_cols, void *stream);

/* Multi-GPU exchange helpers: owner rank of a fingerprint is
 * fp.hi % n_ranks.  bucket writes (fp, global order key) into
 * out_fp[n_ranks][cap][2] / out_key[n_ranks][cap], zero-pads each bucket's
 * tail, and sets counts[r]; overflow beyond cap is reported through
 * counts[r] > cap.  scatter writes verdict bytes back by local index. */
int ob2_bucket_by_owner(int64_t n, const uint64_t *fp, uint32_t key_base, int32_t n_ranks,
                        int64_t cap, uint64_t *out_fp, uint32_t *out_key, uint32_t *out_src,
                        uint32_t *counts, void *stream);
int ob2_scatter_verdict(int64_t n_entries, const uint32_t *src_index, const uint8_t *verdict,
                        const uint64_t *fp, uint8_t *keep, void *stream);

/* Peer-memory exchange: the same sharded dedup with the transfers done by the
 * kernels themselves over NVLink (no counterpart in the reference).
 * Every rank creates an exchange (it owns an inbox region in its HBM -- the one
 * place where the library allocates, because CUDA IPC shares whole
 * allocations), publishes its 64-byte IPC handle, and opens everybody else's.
 * Per batch, on one stream:
 *     ob2_xchg_send     bucket by owner, storing (fp, order key) straight into
 *                       the owners' inboxes
 *     <barrier>         any stream-ordered collective (e.g. a 1-element all-reduce)
 *     ob2_xchg_dedup    dedup my inbox against my shard, store the verdicts
 *                       straight into the sources' verdict inboxes
 *                       (insert_only bit 0: history load, no verdicts; bit 1: lanes arrive padded)
 *     <barrier>
 *     ob2_xchg_collect  scatter the verdicts I received into keep[0..n)
 * `cap` = capacity of one (source, owner) lane, rounded up by the library to a multiple of 16
 * (ob2_xchg_cap returns the value in force); ob2_xchg_send_counts reads back how many entries
 * went to each owner (a count above cap means dropped entries; synchronises), and
 * ob2_xchg_flag_overflow raises *flag (device u32, never cleared by the library) to 1 on the
 * stream if any lane of the last send overflowed -- so that the ranks can agree on it with a
 * collective they run anyway, without a host round trip per batch. */
typedef struct ob2_xchg ob2_xchg;
int ob2_xchg_create(int32_t rank, int32_t world, int64_t cap, ob2_xchg **out);
void ob2_xchg_destroy(ob2_xchg *x);
int ob2_xchg_ipc_handle(const ob2_xchg *x, void *handle64);
int ob2_xchg_open_peers(ob2_xchg *x, const void *handles /* [world][64] */);
int ob2_xchg_set_peers(ob2_xchg *x, void *const *regions /* same-process ranks */);
void *ob2_xchg_region(const ob2_xchg *x);
int ob2_xchg_send(ob2_xchg *x, int64_t n, const uint64_t *fp, uint32_t key_base, void *stream);
int ob2_xchg_dedup(ob2_xchg *x, ob2_set *set, int32_t insert_only, void *stream);
int ob2_xchg_collect(ob2_xchg *x, int64_t n, uint8_t *keep, void *stream);
int ob2_xchg_send_counts(const ob2_xchg *x, uint32_t *counts_host, void *stream);
/* The same two phases with the bytes moved by the DMA engines (cudaMemcpyAsync into / out of the
 * peers' regions) instead of remote stores and loads issued by kernels: ob2_xchg_send_copy buckets
 * into local staging and ships whole zero-padded lanes (call ob2_xchg_dedup with bit 1 of
 * `insert_only` set: no padding pass needed); ob2_xchg_collect_copy pulls the verdict lanes and
 * scatters them.  Same inboxes, same verdicts; the transfers do not occupy SMs. */
int ob2_xchg_send_copy(ob2_xchg *x, int64_t n, const uint64_t *fp, uint32_t key_base, void *stream);
int ob2_xchg_collect_copy(ob2_xchg *x, int64_t n, uint8_t *keep, void *stream);
int64_t ob2_xchg_cap(const ob2_xchg *x);
int ob2_xchg_flag_overflow(const ob2_xchg *x, uint32_t *flag, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* OB2_H_ */
