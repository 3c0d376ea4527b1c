// K3: row fingerprints and the history hash set.
//
// The set is an open-addressing table of 32-byte slots -- one DRAM sector per
// probe -- keyed by the 128-bit fingerprint:
//     { u64 lo, u64 hi, u64 stamp, u32 born, u32 winner }
// lo == 0 marks an empty slot (fingerprints never have lo == 0).  Keys are
// claimed with one 128-bit compare-and-swap, so a slot is either empty or holds
// a complete key.  Deterministic first-wins inside a batch: every dedup call
// gets a new epoch.  What bounds this kernel is the number of L2 transactions
// per entry, not DRAM bytes (benchmarks/micro/random_access.cu: 1.25e7 random
// 16-byte reads take 0.29 ms, read + CAS 0.62, + one store 0.84, + the second
// store and the atomicMax of the first protocol 1.12), so pass 1 spends as few
// as it can.  It finds-or-inserts each fingerprint: a key that was there before
// the call is a duplicate (final, two reads, no atomic); a key it inserts is kept
// provisionally and tagged {born = epoch, winner = entry} with ONE 8-byte store;
// an entry that finds a key inserted by ANOTHER entry of the same call (`born` ==
// epoch) bids for it with atomicMax on the slot's stamp ((epoch << 32) | ~order).
// Pass 2 revisits only those in-batch duplicates: the best bid takes the slot
// from the inserter if its order key is smaller than the inserter's (whose
// provisional keep is then withdrawn).  No sort, no per-batch clearing.
#include <mutex>
#include <new>

#include "ob2_internal.cuh"

namespace {

struct __align__(32) SetSlot {
    uint64_t lo, hi;
    unsigned long long stamp;
    uint32_t born;    // 0 = never completed an insert; epoch of the dedup call that inserted it; or OB2_BORN_HISTORY
    uint32_t winner;  // batch position of the entry whose CAS inserted the key (meaningful within its epoch)
};
static_assert(sizeof(SetSlot) == OB2_SET_SLOT_BYTES, "slot size");

struct SetHeader {
    unsigned long long count;  // distinct fingerprints stored
    uint32_t overflow;         // a probe sequence wrapped the whole table
    uint32_t n_conflicts;      // entries of the running dedup call that met a key inserted by the same call
};

}  // namespace

struct ob2_set {
    int device;
    SetHeader *header;
    SetSlot *slots;
    uint64_t n_slots;
    uint64_t upper_bound;  // host-side bound on entries inserted so far
    uint32_t epoch;
    int32_t dedup_grid;  // 0: one thread per entry; > 0: that many CTAs walk the entries (ob2_set_set_dedup_grid)
};

namespace {

// 128-bit CAS on the key; returns the previous (lo, hi).
__device__ __forceinline__ void cas128(SetSlot *slot, uint64_t cmp_lo, uint64_t cmp_hi, uint64_t new_lo,
                                       uint64_t new_hi, uint64_t &old_lo, uint64_t &old_hi) {
    asm volatile(
        "{\n"
        ".reg .b128 cmp, val, old;\n"
        "mov.b128 cmp, {%2, %3};\n"
        "mov.b128 val, {%4, %5};\n"
        "atom.relaxed.gpu.global.cas.b128 old, [%6], cmp, val;\n"
        "mov.b128 {%0, %1}, old;\n"
        "}\n"
        : "=l"(old_lo), "=l"(old_hi)
        : "l"(cmp_lo), "l"(cmp_hi), "l"(new_lo), "l"(new_hi), "l"(slot)
        : "memory");
}

__device__ __forceinline__ void ld_key(const SetSlot *slot, uint64_t &lo, uint64_t &hi) {
    // L2-coherent read: keys only ever go empty -> set, so a stale "empty" is
    // resolved by the CAS and a complete key never changes again
    asm volatile("ld.global.cg.v2.u64 {%0, %1}, [%2];" : "=l"(lo), "=l"(hi) : "l"(slot));
}

enum { FOUND = 0, INSERTED = 1, OVERFLOW = 2 };

// Find the slot holding (lo, hi), inserting it if absent.  Read first, CAS only into a slot seen
// empty: a blind CAS (one transaction per insert on an EMPTY table, benchmarks/micro/random_access.cu)
// was tried and doubled the kernel's time on a table that is a quarter full -- every probe of an
// occupied slot, the ~0.5 extra per entry that linear probing costs at this load plus the 10 %
// history duplicates, became an atomic on a line that then has to be written back.
__device__ __forceinline__ int find_or_insert(SetSlot *slots, uint64_t mask, uint64_t lo, uint64_t hi, uint64_t &pos) {
    pos = lo & mask;
    for (uint64_t step = 0; step <= mask; ++step) {
        SetSlot *slot = slots + pos;
        uint64_t klo, khi;
        ld_key(slot, klo, khi);
        if (klo == 0) {
            cas128(slot, 0, 0, lo, hi, klo, khi);
            if (klo == 0 && khi == 0) return INSERTED;
            if (klo == lo && khi == hi) return FOUND;
        } else if (klo == lo) {
            if (khi == hi) return FOUND;
            // same low word, different high word: true 64-bit collision or a torn
            // read of a key being written -- re-read atomically to tell
            cas128(slot, lo, hi, lo, hi, klo, khi);
            if (klo == lo && khi == hi) return FOUND;
        }
        pos = (pos + 1) & mask;
    }
    return OVERFLOW;
}

// Read-only lookup.
__device__ __forceinline__ bool find(const SetSlot *slots, uint64_t mask, uint64_t lo, uint64_t hi) {
    uint64_t pos = lo & mask;
    for (uint64_t step = 0; step <= mask; ++step) {
        uint64_t klo, khi;
        ld_key(slots + pos, klo, khi);
        if (klo == 0) return false;
        if (klo == lo && khi == hi) return true;
        pos = (pos + 1) & mask;
    }
    return false;
}

// Slot index of a key that is known to be stored.
__device__ __forceinline__ uint64_t locate(const SetSlot *slots, uint64_t mask, uint64_t lo, uint64_t hi) {
    uint64_t pos = lo & mask;
    for (uint64_t step = 0; step <= mask; ++step) {
        uint64_t klo, khi;
        ld_key(slots + pos, klo, khi);
        if ((klo == lo && khi == hi) || klo == 0) break;
        pos = (pos + 1) & mask;
    }
    return pos;
}

__device__ __forceinline__ void warp_count_add(unsigned long long *counter, bool flag) {
    const unsigned ballot = __ballot_sync(__activemask(), flag);
    if (flag && (threadIdx.x & 31) == (__ffs(ballot) - 1)) atomicAdd(counter, (unsigned long long)__popc(ballot));
}

#define OB2_BORN_HISTORY 0xFFFFFFFFu  // inserted by ob2_set_insert (not by a dedup call)
#define OB2_NO_CONFLICT 0xFFFFFFFFu   // scratch marker: verdict already final after pass 1

// Which entry a thread owns.  Flat batches: the global thread index, valid below n.  An exchange
// inbox (lane_cnt != nullptr): blockIdx.y is the source lane of `lane_cap` entries of which the
// first lane_cnt[y] are real -- the tails are never read, so nobody has to blank them.
__device__ __forceinline__ bool entry_index(int64_t n, const uint32_t *__restrict__ lane_cnt, int64_t lane_cap, int64_t &i) {
    const int64_t local = (int64_t)blockIdx.x * OB2_BLOCK + threadIdx.x;
    if (lane_cnt == nullptr) {
        i = local;
        return i < n;
    }
    i = (int64_t)blockIdx.y * lane_cap + local;
    const int64_t used = (int64_t)__ldg(lane_cnt + blockIdx.y);
    return local < (used < lane_cap ? used : lane_cap);
}

__global__ void __launch_bounds__(OB2_BLOCK)
set_insert_kernel(SetHeader *hdr, SetSlot *slots, uint64_t mask, int64_t n, const uint64_t *__restrict__ fp,
                  const uint32_t *__restrict__ lane_cnt, int64_t lane_cap) {
    int64_t i;
    const bool valid = entry_index(n, lane_cnt, lane_cap, i);
    bool inserted = false;
    if (valid) {
        const ulonglong2 key = __ldg(reinterpret_cast<const ulonglong2 *>(fp) + i);
        if (key.x != 0) {
            uint64_t pos;
            const int r = find_or_insert(slots, mask, key.x, key.y, pos);
            inserted = r == INSERTED;
            if (inserted) slots[pos].born = OB2_BORN_HISTORY;  // every completed insert has born != 0
            if (r == OVERFLOW) hdr->overflow = 1;
        }
    }
    warp_count_add(&hdr->count, inserted);
}

__device__ __forceinline__ unsigned long long dedup_tag(uint32_t epoch, uint32_t ord) {
    return ((unsigned long long)epoch << 32) | (unsigned long long)(0xFFFFFFFFu - ord);  // max tag = smallest order key
}

// Pass 1 settles almost every entry with ONE probe (scratch: the list of entry indices left to pass 2):
//   - key absent  -> insert it (128-bit CAS), keep = 1 for now, remember who won the slot;
//   - key present since an earlier call (born != 0 and != epoch) -> keep = 0, no atomic at all;
//   - key present and inserted during THIS call (born == epoch, or 0 = the winner's store is
//     still in flight) -> a duplicate inside the batch: compete for the slot's stamp and leave
//     the verdict to pass 2 (the entry's index is appended to the list in `scratch`).
// Only the last class -- the rows that really are duplicated inside the batch -- touches the
// table a second time.
__device__ __forceinline__ void dedup_pass1_entry(SetHeader *hdr, SetSlot *slots, uint64_t mask, int64_t i, bool valid,
                                                  const uint64_t *__restrict__ fp, const uint32_t *__restrict__ order,
                                                  uint32_t epoch, uint32_t *__restrict__ scratch,
                                                  uint8_t *__restrict__ keep) {
    bool inserted = false, conflict = false;
    if (valid) {
        uint8_t verdict = 0;
        const ulonglong2 key = __ldg(reinterpret_cast<const ulonglong2 *>(fp) + i);
        if (key.x != 0) {  // lo == 0 is padding
            uint64_t pos;
            const int r = find_or_insert(slots, mask, key.x, key.y, pos);
            SetSlot *slot = slots + pos;
            if (r == INSERTED) {
                inserted = true;
                // {born, winner} in one 8-byte store; the inserter places no bid of its own (pass 2
                // compares the best bid with the inserter's order key)
                *reinterpret_cast<uint2 *>(&slot->born) = make_uint2(epoch, (uint32_t)i);
                verdict = 1;
            } else if (r == FOUND) {
                uint32_t born;
                asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(born) : "l"(&slot->born));
                if (born == 0 || born == epoch) {
                    const uint32_t ord = order ? __ldg(order + i) : (uint32_t)i;
                    atomicMax(&slot->stamp, dedup_tag(epoch, ord));
                    conflict = true;
                }
            } else {
                hdr->overflow = 1;
            }
        }
        keep[i] = verdict;
    }
    // the ~1 % that met a key of this same call go on a list for pass 2 (scratch = their entry indices)
    const unsigned bidders = __ballot_sync(0xFFFFFFFFu, conflict);
    if (bidders) {
        const unsigned lane = threadIdx.x & 31, leader = __ffs(bidders) - 1;
        uint32_t base = 0;
        if (lane == leader) base = atomicAdd(&hdr->n_conflicts, (uint32_t)__popc(bidders));
        base = __shfl_sync(0xFFFFFFFFu, base, leader);
        if (conflict) scratch[base + __popc(bidders & ((1u << lane) - 1))] = (uint32_t)i;
    }
    warp_count_add(&hdr->count, inserted);
}

__global__ void __launch_bounds__(OB2_BLOCK)
set_dedup_pass1(SetHeader *hdr, SetSlot *slots, uint64_t mask, int64_t n, const uint64_t *__restrict__ fp,
                const uint32_t *__restrict__ order, uint32_t epoch, uint32_t *__restrict__ scratch,
                uint8_t *__restrict__ keep, const uint32_t *__restrict__ lane_cnt, int64_t lane_cap) {
    int64_t i;
    const bool valid = entry_index(n, lane_cnt, lane_cap, i);
    dedup_pass1_entry(hdr, slots, mask, i, valid, fp, order, epoch, scratch, keep);
}

// The same over a small fixed grid that walks the entries (ob2_set_set_dedup_grid): few CTAs, resident for the
// whole pass.  A grid of one thread per entry takes every SM slot a kernel of another stream frees (the two
// take turns, their times add up); a walking grid of one or two CTAs per SM leaves the rest to that kernel.
__global__ void __launch_bounds__(OB2_BLOCK)
set_dedup_pass1_walk(SetHeader *hdr, SetSlot *slots, uint64_t mask, int64_t n, const uint64_t *__restrict__ fp,
                     const uint32_t *__restrict__ order, uint32_t epoch, uint32_t *__restrict__ scratch,
                     uint8_t *__restrict__ keep, const uint32_t *__restrict__ lane_cnt, int64_t lane_cap) {
    if (lane_cnt == nullptr) {
        const int64_t stride = (int64_t)gridDim.x * OB2_BLOCK;
        for (int64_t base = (int64_t)blockIdx.x * OB2_BLOCK; base < n; base += stride) {
            const int64_t i = base + threadIdx.x;
            dedup_pass1_entry(hdr, slots, mask, i, i < n, fp, order, epoch, scratch, keep);
        }
        return;
    }
    // an exchange inbox: n / lane_cap lanes, the first lane_cnt[lane] entries of each are real
    const int64_t per_lane = (lane_cap + OB2_BLOCK - 1) / OB2_BLOCK, tiles = per_lane * (n / lane_cap);
    for (int64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
        const int64_t lane = t / per_lane, first = (t - lane * per_lane) * OB2_BLOCK;
        int64_t used = (int64_t)__ldg(lane_cnt + lane);
        used = used < lane_cap ? used : lane_cap;
        if (first >= used) continue;  // (the whole CTA takes this branch together)
        const int64_t local = first + threadIdx.x;
        dedup_pass1_entry(hdr, slots, mask, lane * lane_cap + local, local < used, fp, order, epoch, scratch, keep);
    }
}

// Pass 2: only the listed entries -- those that met a key inserted by this same call.  `stamp` holds
// the best bid among those FINDERS (largest tag = smallest order key; a slot inserted in this call
// started with stamp 0).  If that bid beats the inserter's own order key, the bidder takes the slot and
// the inserter's provisional keep = 1 is withdrawn (every loser writes the same 0, so the race is
// benign); otherwise the inserter stays and every finder loses.  A small fixed grid walks the list.
__global__ void __launch_bounds__(OB2_BLOCK)
set_dedup_pass2(const SetHeader *__restrict__ hdr, const SetSlot *__restrict__ slots, uint64_t mask,
                const uint64_t *__restrict__ fp, const uint32_t *__restrict__ order, uint32_t epoch,
                const uint32_t *__restrict__ scratch, uint8_t *__restrict__ keep) {
    const uint32_t n_listed = hdr->n_conflicts;
    for (uint32_t j = blockIdx.x * OB2_BLOCK + threadIdx.x; j < n_listed; j += gridDim.x * OB2_BLOCK) {
        const uint32_t i = scratch[j];
        const ulonglong2 key = __ldg(reinterpret_cast<const ulonglong2 *>(fp) + i);
        const SetSlot *slot = slots + locate(slots, mask, key.x, key.y);
        const unsigned long long stamp = slot->stamp;
        const uint32_t w = slot->winner;
        const uint32_t ord_w = order ? __ldg(order + w) : w;
        if (stamp > dedup_tag(epoch, ord_w)) {
            keep[w] = 0;
            const uint32_t ord = order ? __ldg(order + i) : i;
            if (stamp == dedup_tag(epoch, ord)) keep[i] = 1;
        }
    }
}

__global__ void __launch_bounds__(OB2_BLOCK)
set_probe_kernel(const SetSlot *__restrict__ slots, uint64_t mask, int64_t n, const uint64_t *__restrict__ fp,
                 uint8_t *__restrict__ found) {
    const int64_t i = (int64_t)blockIdx.x * OB2_BLOCK + threadIdx.x;
    if (i >= n) return;
    const ulonglong2 key = __ldg(reinterpret_cast<const ulonglong2 *>(fp) + i);
    found[i] = (key.x != 0 && find(slots, mask, key.x, key.y)) ? 1 : 0;
}

// Fingerprint of existing columns (same arithmetic as the fused path in the sampler).
__global__ void __launch_bounds__(OB2_BLOCK)
fingerprint_kernel(const ob2_slot_desc *__restrict__ g_slots, int n_slots, int64_t n, int64_t n_stride,
                   const char *__restrict__ cols, uint64_t *__restrict__ fp) {
    extern __shared__ __align__(16) unsigned char smem[];
    ob2_slot_desc *s_slots = reinterpret_cast<ob2_slot_desc *>(smem);
    ob2_stage_tables(g_slots, n_slots, nullptr, 0, s_slots, nullptr);
    const int64_t i0 = ((int64_t)blockIdx.x * OB2_BLOCK + threadIdx.x) * OB2_CPT;
    if (i0 >= n) return;
    FpAcc acc0{0, 0}, acc1{0, 0};
    for (int s = 0; s < n_slots; ++s) {
        const ob2_slot_desc &d = s_slots[s];
        const char *col = cols + (int64_t)d.col_units * n_stride;
        uint64_t w0, w1;
        if (d.kind == OB2_KIND_CATEGORICAL) {
            int c0, c1;
            ld_i32x2(reinterpret_cast<const int *>(col) + i0, c0, c1);
            w0 = (uint64_t)(int64_t)c0;
            w1 = (uint64_t)(int64_t)c1;
        } else if (d.kind == OB2_KIND_INTEGER) {
            long long q0, q1;
            ld_i64x2(reinterpret_cast<const long long *>(col) + i0, q0, q1);
            w0 = (uint64_t)q0;
            w1 = (uint64_t)q1;
        } else {
            double x0, x1;
            ld_f64x2(reinterpret_cast<const double *>(col) + i0, x0, x1);
            w0 = canon_f64(x0);
            w1 = canon_f64(x1);
        }
        fp_add(acc0, d, w0);
        fp_add(acc1, d, w1);
    }
    uint64_t lo0, hi0, lo1, hi1;
    fp_finish(acc0, n_slots, lo0, hi0);
    fp_finish(acc1, n_slots, lo1, hi1);
    if (i0 + 1 >= n) lo1 = hi1 = 0;
    st_u64x4(fp + 2 * i0, lo0, hi0, lo1, hi1);
}

// ---- multi-GPU exchange helpers
#define OB2_MAX_RANKS 64

__global__ void __launch_bounds__(OB2_BLOCK)
bucket_kernel(int64_t n, const uint64_t *__restrict__ fp, uint32_t key_base, int n_ranks, int64_t cap,
              uint64_t *__restrict__ out_fp, uint32_t *__restrict__ out_key, uint32_t *__restrict__ out_src,
              uint32_t *__restrict__ counts) {
    __shared__ uint32_t s_cnt[OB2_MAX_RANKS];
    __shared__ uint32_t s_base[OB2_MAX_RANKS];
    if (threadIdx.x < n_ranks) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    const int64_t i = (int64_t)blockIdx.x * OB2_BLOCK + threadIdx.x;
    int owner = -1;
    uint32_t rank_in_block = 0;
    ulonglong2 key = make_ulonglong2(0, 0);
    if (i < n) {
        key = __ldg(reinterpret_cast<const ulonglong2 *>(fp) + i);
        if (key.x != 0) {
            owner = ob2_owner_of(key.y, (uint32_t)n_ranks);
            rank_in_block = atomicAdd(&s_cnt[owner], 1u);
        }
    }
    __syncthreads();
    if (threadIdx.x < n_ranks) s_base[threadIdx.x] = s_cnt[threadIdx.x] ? atomicAdd(&counts[threadIdx.x], s_cnt[threadIdx.x]) : 0;
    __syncthreads();
    if (owner >= 0) {
        const int64_t pos = (int64_t)s_base[owner] + rank_in_block;
        if (pos < cap) {
            const int64_t o = (int64_t)owner * cap + pos;
            reinterpret_cast<ulonglong2 *>(out_fp)[o] = key;
            out_key[o] = key_base + (uint32_t)i;
            out_src[o] = (uint32_t)i;
        }
    }
}

__global__ void __launch_bounds__(OB2_BLOCK)
bucket_pad_kernel(int n_ranks, int64_t cap, const uint32_t *__restrict__ counts, uint64_t *__restrict__ out_fp) {
    const int r = blockIdx.y;
    const int64_t used = min((int64_t)counts[r], cap);
    for (int64_t p = used + (int64_t)blockIdx.x * OB2_BLOCK + threadIdx.x; p < cap; p += (int64_t)gridDim.x * OB2_BLOCK) {
        reinterpret_cast<ulonglong2 *>(out_fp)[(int64_t)r * cap + p] = make_ulonglong2(0, 0);
    }
}

__global__ void __launch_bounds__(OB2_BLOCK)
scatter_verdict_kernel(int64_t n_entries, const uint32_t *__restrict__ src_index, const uint8_t *__restrict__ verdict,
                       const uint64_t *__restrict__ fp, uint8_t *__restrict__ keep) {
    const int64_t o = (int64_t)blockIdx.x * OB2_BLOCK + threadIdx.x;
    if (o >= n_entries) return;
    if (__ldg(fp + 2 * o) == 0) return;  // padding
    keep[src_index[o]] = verdict[o];
}

int grid_for(int64_t n, unsigned *blocks) {
    const int64_t b = (n + OB2_BLOCK - 1) / OB2_BLOCK;
    if (b > 0x7FFFFFFFll) {
        ob2_set_error("batch too large for one launch");
        return OB2_ETOOBIG;
    }
    *blocks = (unsigned)b;
    return OB2_OK;
}

int check_set_args(const ob2_set *set, int64_t n, const void *fp) {
    if (set == nullptr || n < 0 || (n > 0 && fp == nullptr)) {
        ob2_set_error("hash set: bad arguments");
        return OB2_EINVAL;
    }
    if (n >= 0xFFFFFFFFll) {
        ob2_set_error("hash set: at most 2^32 - 2 entries per call");
        return OB2_ETOOBIG;
    }
    if ((reinterpret_cast<uintptr_t>(fp) & 15) != 0) {
        ob2_set_error("hash set: the fingerprint array must be 16-byte aligned (it is read as 128-bit words)");
        return OB2_EINVAL;
    }
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev != set->device) {
        ob2_set_error("current device %d is not the device %d this set lives on", dev, set->device);
        return OB2_EINVAL;
    }
    return OB2_OK;
}

// Refuse to push the load factor past 7/8: linear probing degrades sharply and a
// full table would spin.  The host bound counts every submitted entry; when it
// trips, fall back to the exact device count (one synchronising read).
int reserve(ob2_set *set, int64_t n, cudaStream_t st) {
    const uint64_t limit = set->n_slots - set->n_slots / 8;
    if (set->upper_bound + (uint64_t)n > limit) {
        unsigned long long exact = 0;
        OB2_CUDA(cudaMemcpyAsync(&exact, &set->header->count, sizeof(exact), cudaMemcpyDeviceToHost, st));
        OB2_CUDA(cudaStreamSynchronize(st));
        set->upper_bound = exact;
        if (exact + (uint64_t)n > limit) {
            ob2_set_error("hash set full: %llu stored + %lld incoming exceeds 7/8 of %llu slots", exact, (long long)n,
                          (unsigned long long)set->n_slots);
            return OB2_ENOMEM;
        }
    }
    set->upper_bound += (uint64_t)n;
    return OB2_OK;
}

}  // namespace

extern "C" int ob2_fingerprint(const ob2_space *space, int64_t n, int64_t n_stride, const void *cols, uint64_t *fp,
                               void *stream) {
    int rc = ob2_check_batch(space, n, n_stride);
    if (rc != OB2_OK) return rc;
    if (cols == nullptr || fp == nullptr) {
        ob2_set_error("fingerprint: null buffer");
        return OB2_EINVAL;
    }
    OB2_ALIGNED(cols, 16, "fingerprint");
    OB2_ALIGNED(fp, 16, "fingerprint");
    if (n == 0) return OB2_OK;
    unsigned blocks;
    rc = grid_for((n + OB2_CPT - 1) / OB2_CPT, &blocks);
    if (rc != OB2_OK) return rc;
    const size_t smem = (size_t)space->n_slots * sizeof(ob2_slot_desc);
    if (smem > 48 * 1024)
        OB2_CUDA(cudaFuncSetAttribute(fingerprint_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    fingerprint_kernel<<<blocks, OB2_BLOCK, smem, static_cast<cudaStream_t>(stream)>>>(
        space->d_slots, space->n_slots, n, n_stride, (const char *)cols, fp);
    return ob2_check_cuda(cudaGetLastError(), "fingerprint_kernel launch");
}

// ---------------------------------------------------------------------------
// ob2_set_dedup_static: the same two-pass dedup for callers that must not
// allocate device memory.
//
// ob2_set_dedup needs n u32 of scratch (the slot index pass 1 hands to pass 2)
// and leaves it to the caller to provide it.  Some callers cannot: a plugin
// running inside a host process whose framework has already claimed the whole
// GPU for its own memory pool (a cudaMalloc from the plugin then fails, or
// fragments the pool); a producer loop that may not stall, since cudaMalloc and
// cudaFree synchronise the device; a thin binding (cgo/JNI/ctypes) that only
// ever owns the input and output buffers it was handed.  For them the scratch
// lives in the library image itself, in the file-scope array below: once the
// CUDA module that holds the hash-set kernels is loaded on a GPU (with CUDA's
// default lazy loading: at the first hash-set call on that GPU, which is
// ob2_set_create), that GPU holds OB2_STATIC_DEDUP_MAX * 4 bytes (4 MiB) of it
// until the process exits, whether or not this entry point is used.
//
// On the host, `g_static_dedup_scratch` is only a handle for the symbol, not a
// device address; the address for the current device is looked up once per
// device with cudaGetSymbolAddress and cached.  ob2_set_create does that
// look-up (it runs with the set's device current, outside any capture, and is
// where a caller expects set-up cost), so that by the time a dedup call is
// made the module is resident and the call itself neither loads nor allocates
// anything; ob2_set_dedup_static repeats it only if it was never done.
static __device__ __align__(256) uint32_t g_static_dedup_scratch[OB2_STATIC_DEDUP_MAX];

namespace {

constexpr int kMaxDevices = 64;
uint32_t *g_static_scratch_ptr[kMaxDevices];  // zero-initialised: not looked up yet
std::mutex g_static_scratch_mutex;

// Device address of the static scratch on the CURRENT device (set-up on first use).
int static_scratch_for_current_device(uint32_t **out) {
    int dev = -1;
    OB2_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) {
        ob2_set_error("ob2_set_dedup_static: device ordinal %d outside [0, %d)", dev, kMaxDevices);
        return OB2_EINVAL;
    }
    std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
    if (g_static_scratch_ptr[dev] == nullptr) {
        void *ptr = nullptr;
        OB2_CUDA(cudaGetSymbolAddress(&ptr, g_static_dedup_scratch));
        g_static_scratch_ptr[dev] = static_cast<uint32_t *>(ptr);
    }
    *out = g_static_scratch_ptr[dev];
    return OB2_OK;
}

}  // namespace

extern "C" int ob2_set_create(void *mem, uint64_t n_slots, void *stream, ob2_set **out) {
    if (out == nullptr || mem == nullptr || n_slots < 64 || (n_slots & (n_slots - 1)) != 0 || n_slots > (1ull << 31)) {
        ob2_set_error("ob2_set_create: n_slots must be a power of two in [64, 2^31] (got %llu)", (unsigned long long)n_slots);
        return OB2_EINVAL;
    }
    if ((reinterpret_cast<uintptr_t>(mem) & 255) != 0) {
        ob2_set_error("ob2_set_create: memory must be 256-byte aligned");
        return OB2_EINVAL;
    }
    *out = nullptr;
    ob2_set *set = new (std::nothrow) ob2_set();
    if (set == nullptr) {
        ob2_set_error("out of host memory");
        return OB2_ENOMEM;
    }
    set->header = reinterpret_cast<SetHeader *>(mem);
    set->slots = reinterpret_cast<SetSlot *>(reinterpret_cast<char *>(mem) + OB2_SET_HEADER_BYTES);
    set->n_slots = n_slots;
    set->upper_bound = 0;
    set->epoch = 0;
    set->dedup_grid = getenv("OB2_DEDUP_GRID") ? atoi(getenv("OB2_DEDUP_GRID")) : 0;
    int rc = ob2_check_cuda(cudaGetDevice(&set->device), "cudaGetDevice");
    if (rc == OB2_OK)
        rc = ob2_check_cuda(cudaMemsetAsync(mem, 0, OB2_SET_HEADER_BYTES + n_slots * OB2_SET_SLOT_BYTES,
                                            static_cast<cudaStream_t>(stream)),
                            "cudaMemsetAsync(table)");
    if (rc != OB2_OK) {
        delete set;
        return rc;
    }
    // Set-up for ob2_set_dedup_static on this device: loads the module that carries its
    // built-in scratch and caches the array's address.  Best effort -- a failure here is
    // reported by ob2_set_dedup_static itself if that entry point is ever used.
    uint32_t *unused = nullptr;
    (void)static_scratch_for_current_device(&unused);
    *out = set;
    return OB2_OK;
}

extern "C" void ob2_set_destroy(ob2_set *set) { delete set; }

extern "C" int ob2_set_insert(ob2_set *set, int64_t n, const uint64_t *fp, void *stream) {
    int rc = check_set_args(set, n, fp);
    if (rc != OB2_OK) return rc;
    if (n == 0) return OB2_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    rc = reserve(set, n, st);
    if (rc != OB2_OK) return rc;
    unsigned blocks;
    rc = grid_for(n, &blocks);
    if (rc != OB2_OK) return rc;
    set_insert_kernel<<<blocks, OB2_BLOCK, 0, st>>>(set->header, set->slots, set->n_slots - 1, n, fp, nullptr, 0);
    return ob2_check_cuda(cudaGetLastError(), "set_insert_kernel launch");
}

namespace {

// The two passes on `st`, after `n_reserve` incoming entries have been cleared against the
// 7/8 load limit (n_reserve <= n: an exchange inbox is launched over its padded lanes but
// only its real entries can be inserted).
int dedup_launch(ob2_set *set, int64_t n, int64_t n_reserve, const uint64_t *fp, const uint32_t *order,
                 uint32_t *scratch, uint8_t *keep, cudaStream_t st, const char *who,
                 const uint32_t *lane_cnt = nullptr, int64_t lane_cap = 0) {
    int rc = reserve(set, n_reserve, st);
    if (rc != OB2_OK) return rc;
    unsigned blocks;
    rc = grid_for(n, &blocks);
    if (rc != OB2_OK) return rc;
    dim3 grid1(blocks);
    if (lane_cnt != nullptr) grid1 = dim3((unsigned)((lane_cap + OB2_BLOCK - 1) / OB2_BLOCK), (unsigned)(n / lane_cap));
    if (set->epoch == 0xFFFFFFFFu) {
        ob2_set_error("%s: epoch counter exhausted", who);
        return OB2_ETOOBIG;
    }
    const uint32_t epoch = ++set->epoch;
    OB2_CUDA(cudaMemsetAsync(&set->header->n_conflicts, 0, sizeof(uint32_t), st));
    if (set->dedup_grid > 0 && blocks > (unsigned)set->dedup_grid) {
        set_dedup_pass1_walk<<<(unsigned)set->dedup_grid, OB2_BLOCK, 0, st>>>(set->header, set->slots, set->n_slots - 1, n, fp,
                                                                              order, epoch, scratch, keep, lane_cnt, lane_cap);
    } else {
        set_dedup_pass1<<<grid1, OB2_BLOCK, 0, st>>>(set->header, set->slots, set->n_slots - 1, n, fp, order, epoch, scratch,
                                                     keep, lane_cnt, lane_cap);
    }
    rc = ob2_check_cuda(cudaGetLastError(), "set_dedup_pass1 launch");
    if (rc != OB2_OK) return rc;
    const unsigned blocks2 = blocks < 592u ? blocks : 592u;  // the list is ~1 % of the batch: 4 CTAs per SM walk it
    set_dedup_pass2<<<blocks2, OB2_BLOCK, 0, st>>>(set->header, set->slots, set->n_slots - 1, fp, order, epoch, scratch, keep);
    return ob2_check_cuda(cudaGetLastError(), "set_dedup_pass2 launch");
}

}  // namespace

extern "C" int ob2_set_dedup(ob2_set *set, int64_t n, const uint64_t *fp, const uint32_t *order, uint32_t *scratch,
                             uint8_t *keep, void *stream) {
    int rc = check_set_args(set, n, fp);
    if (rc != OB2_OK) return rc;
    if (n == 0) return OB2_OK;
    if (scratch == nullptr || keep == nullptr) {
        ob2_set_error("ob2_set_dedup: null scratch/keep");
        return OB2_EINVAL;
    }
    return dedup_launch(set, n, n, fp, order, scratch, keep, static_cast<cudaStream_t>(stream), "ob2_set_dedup");
}

extern "C" int ob2_set_set_dedup_grid(ob2_set *set, int32_t ctas) {
    if (set == nullptr || ctas < 0 || ctas > 65535) {
        ob2_set_error("ob2_set_set_dedup_grid: null set or a CTA count outside [0, 65535]");
        return OB2_EINVAL;
    }
    set->dedup_grid = ctas;
    return OB2_OK;
}

// library-internal (ob2_xchg.cu): an inbox of `n` padded lanes holding `n_real` entries
bool ob2_set_would_trip(const ob2_set *set, int64_t n) {
    return set->upper_bound + (uint64_t)n > set->n_slots - set->n_slots / 8;
}

int ob2_set_dedup_padded(ob2_set *set, int64_t n, int64_t n_real, const uint64_t *fp, const uint32_t *order,
                         uint32_t *scratch, uint8_t *keep, void *stream, const uint32_t *lane_cnt, int64_t lane_cap) {
    int rc = check_set_args(set, n, fp);
    if (rc != OB2_OK) return rc;
    if (n == 0) return OB2_OK;
    return dedup_launch(set, n, n_real, fp, order, scratch, keep, static_cast<cudaStream_t>(stream), "ob2_xchg_dedup",
                        lane_cnt, lane_cap);
}

int ob2_set_insert_padded(ob2_set *set, int64_t n, int64_t n_real, const uint64_t *fp, void *stream,
                          const uint32_t *lane_cnt, int64_t lane_cap) {
    int rc = check_set_args(set, n, fp);
    if (rc != OB2_OK) return rc;
    if (n == 0) return OB2_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    rc = reserve(set, n_real, st);
    if (rc != OB2_OK) return rc;
    unsigned blocks;
    rc = grid_for(n, &blocks);
    if (rc != OB2_OK) return rc;
    dim3 grid(blocks);
    if (lane_cnt != nullptr) grid = dim3((unsigned)((lane_cap + OB2_BLOCK - 1) / OB2_BLOCK), (unsigned)(n / lane_cap));
    set_insert_kernel<<<grid, OB2_BLOCK, 0, st>>>(set->header, set->slots, set->n_slots - 1, n, fp, lane_cnt, lane_cap);
    return ob2_check_cuda(cudaGetLastError(), "set_insert_kernel launch");
}

extern "C" int ob2_set_dedup_static(ob2_set *set, int64_t n, const uint64_t *fp, const uint32_t *order,
                                    uint8_t *keep, void *stream) {
    int rc = check_set_args(set, n, fp);
    if (rc != OB2_OK) return rc;
    if (n > (int64_t)OB2_STATIC_DEDUP_MAX) {
        ob2_set_error("ob2_set_dedup_static: n = %lld exceeds the built-in scratch of %d entries; split the batch "
                      "or use ob2_set_dedup with caller-provided scratch",
                      (long long)n, (int)OB2_STATIC_DEDUP_MAX);
        return OB2_ETOOBIG;
    }
    if (n == 0) return OB2_OK;
    if (keep == nullptr) {
        ob2_set_error("ob2_set_dedup_static: null keep");
        return OB2_EINVAL;
    }
    int dev = -1;
    OB2_CUDA(cudaGetDevice(&dev));
    if (dev != set->device) {
        ob2_set_error("ob2_set_dedup_static: current device %d is not the set's device %d", dev, set->device);
        return OB2_EINVAL;
    }
    uint32_t *scratch = nullptr;
    rc = static_scratch_for_current_device(&scratch);
    if (rc != OB2_OK) return rc;
    // Nothing in the scratch has to start at zero: pass 1 appends the entries that need a second
    // look to a list that starts empty on every call (its length lives in the set's header, which
    // the call resets), and pass 2 reads exactly that many -- whatever an earlier call left
    // behind in the array is never consumed.
    return dedup_launch(set, n, n, fp, order, scratch, keep, static_cast<cudaStream_t>(stream), "ob2_set_dedup_static");
}

extern "C" int ob2_set_probe(const ob2_set *set, int64_t n, const uint64_t *fp, uint8_t *found, void *stream) {
    int rc = check_set_args(set, n, fp);
    if (rc != OB2_OK) return rc;
    if (n == 0) return OB2_OK;
    if (found == nullptr) {
        ob2_set_error("ob2_set_probe: null output");
        return OB2_EINVAL;
    }
    unsigned blocks;
    rc = grid_for(n, &blocks);
    if (rc != OB2_OK) return rc;
    set_probe_kernel<<<blocks, OB2_BLOCK, 0, static_cast<cudaStream_t>(stream)>>>(set->slots, set->n_slots - 1, n, fp, found);
    return ob2_check_cuda(cudaGetLastError(), "set_probe_kernel launch");
}

namespace {

// table growth: every stored key moves to the new table as history
__global__ void __launch_bounds__(OB2_BLOCK)
set_rehash_kernel(const SetSlot *__restrict__ old_slots, uint64_t old_n, SetHeader *hdr, SetSlot *slots, uint64_t mask) {
    const uint64_t i = (uint64_t)blockIdx.x * OB2_BLOCK + threadIdx.x;
    bool inserted = false;
    if (i < old_n) {
        const ulonglong2 key = *reinterpret_cast<const ulonglong2 *>(old_slots + i);
        if (key.x != 0) {
            uint64_t pos;
            const int r = find_or_insert(slots, mask, key.x, key.y, pos);
            inserted = r == INSERTED;
            if (inserted) slots[pos].born = OB2_BORN_HISTORY;
            if (r == OVERFLOW) hdr->overflow = 1;
        }
    }
    warp_count_add(&hdr->count, inserted);
}

// snapshot: the stored fingerprints, in no particular order
__global__ void __launch_bounds__(OB2_BLOCK)
set_export_kernel(const SetSlot *__restrict__ slots, uint64_t n_slots, uint64_t max_out, uint64_t *__restrict__ out,
                  unsigned long long *__restrict__ n_out) {
    const uint64_t i = (uint64_t)blockIdx.x * OB2_BLOCK + threadIdx.x;
    ulonglong2 key = make_ulonglong2(0, 0);
    if (i < n_slots) key = *reinterpret_cast<const ulonglong2 *>(slots + i);
    const bool live = key.x != 0;
    const unsigned ballot = __ballot_sync(0xFFFFFFFFu, live);
    if (ballot == 0) return;
    const unsigned lane = threadIdx.x & 31, leader = __ffs(ballot) - 1;
    unsigned long long base = 0;
    if (lane == leader) base = atomicAdd(n_out, (unsigned long long)__popc(ballot));
    base = __shfl_sync(0xFFFFFFFFu, base, leader);
    if (live) {
        const unsigned long long at = base + __popc(ballot & ((1u << lane) - 1));
        if (at < max_out) reinterpret_cast<ulonglong2 *>(out)[at] = key;
    }
}

}  // namespace

extern "C" int ob2_set_rehash(const ob2_set *src, ob2_set *dst, void *stream) {
    if (src == nullptr || dst == nullptr || src == dst) {
        ob2_set_error("ob2_set_rehash: bad arguments");
        return OB2_EINVAL;
    }
    if (dst->upper_bound != 0) {
        ob2_set_error("ob2_set_rehash: the destination set must be freshly created");
        return OB2_EINVAL;
    }
    unsigned blocks;
    int rc = grid_for((int64_t)src->n_slots, &blocks);
    if (rc != OB2_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    set_rehash_kernel<<<blocks, OB2_BLOCK, 0, st>>>(src->slots, src->n_slots, dst->header, dst->slots, dst->n_slots - 1);
    rc = ob2_check_cuda(cudaGetLastError(), "set_rehash_kernel launch");
    if (rc != OB2_OK) return rc;
    unsigned long long moved = 0;
    OB2_CUDA(cudaMemcpyAsync(&moved, &dst->header->count, sizeof(moved), cudaMemcpyDeviceToHost, st));
    OB2_CUDA(cudaStreamSynchronize(st));
    dst->upper_bound = moved;
    dst->epoch = src->epoch;  // later dedup calls keep distinguishing "this call" from everything before
    return OB2_OK;
}

extern "C" int ob2_set_export(const ob2_set *set, uint64_t max_out, uint64_t *out_fp, uint64_t *n_out, void *stream) {
    if (set == nullptr || n_out == nullptr || (max_out > 0 && out_fp == nullptr)) {
        ob2_set_error("ob2_set_export: bad arguments");
        return OB2_EINVAL;
    }
    unsigned blocks;
    int rc = grid_for((int64_t)set->n_slots, &blocks);
    if (rc != OB2_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    OB2_CUDA(cudaMemsetAsync(n_out, 0, sizeof(uint64_t), st));
    set_export_kernel<<<blocks, OB2_BLOCK, 0, st>>>(set->slots, set->n_slots, max_out, out_fp,
                                                    reinterpret_cast<unsigned long long *>(n_out));
    return ob2_check_cuda(cudaGetLastError(), "set_export_kernel launch");
}

extern "C" int ob2_set_count(const ob2_set *set, uint64_t *count, void *stream) {
    if (set == nullptr || count == nullptr) {
        ob2_set_error("ob2_set_count: bad arguments");
        return OB2_EINVAL;
    }
    SetHeader h;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    OB2_CUDA(cudaMemcpyAsync(&h, set->header, sizeof(h), cudaMemcpyDeviceToHost, st));
    OB2_CUDA(cudaStreamSynchronize(st));
    if (h.overflow) {
        ob2_set_error("hash set overflowed: a probe sequence wrapped the whole table");
        return OB2_ENOMEM;
    }
    *count = h.count;
    return OB2_OK;
}

extern "C" int ob2_bucket_by_owner(int64_t n, const uint64_t *fp, uint32_t key_base, int32_t n_ranks, int64_t cap,
                                   uint64_t *out_fp, uint32_t *out_key, uint32_t *out_src, uint32_t *counts,
                                   void *stream) {
    if (n < 0 || n_ranks < 1 || n_ranks > OB2_MAX_RANKS || cap < 1 || fp == nullptr || out_fp == nullptr ||
        out_key == nullptr || out_src == nullptr || counts == nullptr) {
        ob2_set_error("ob2_bucket_by_owner: bad arguments (n=%lld ranks=%d cap=%lld)", (long long)n, n_ranks, (long long)cap);
        return OB2_EINVAL;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    OB2_CUDA(cudaMemsetAsync(counts, 0, sizeof(uint32_t) * n_ranks, st));
    if (n > 0) {
        unsigned blocks;
        int rc = grid_for(n, &blocks);
        if (rc != OB2_OK) return rc;
        bucket_kernel<<<blocks, OB2_BLOCK, 0, st>>>(n, fp, key_base, n_ranks, cap, out_fp, out_key, out_src, counts);
        rc = ob2_check_cuda(cudaGetLastError(), "bucket_kernel launch");
        if (rc != OB2_OK) return rc;
    }
    const int64_t pad_blocks = (cap + OB2_BLOCK - 1) / OB2_BLOCK;
    dim3 grid((unsigned)(pad_blocks < 256 ? pad_blocks : 256), (unsigned)n_ranks);
    bucket_pad_kernel<<<grid, OB2_BLOCK, 0, st>>>(n_ranks, cap, counts, out_fp);
    return ob2_check_cuda(cudaGetLastError(), "bucket_pad_kernel launch");
}

extern "C" int ob2_scatter_verdict(int64_t n_entries, const uint32_t *src_index, const uint8_t *verdict,
                                   const uint64_t *fp, uint8_t *keep, void *stream) {
    if (n_entries < 0 || src_index == nullptr || verdict == nullptr || fp == nullptr || keep == nullptr) {
        ob2_set_error("ob2_scatter_verdict: bad arguments");
        return OB2_EINVAL;
    }
    if (n_entries == 0) return OB2_OK;
    unsigned blocks;
    int rc = grid_for(n_entries, &blocks);
    if (rc != OB2_OK) return rc;
    scatter_verdict_kernel<<<blocks, OB2_BLOCK, 0, static_cast<cudaStream_t>(stream)>>>(n_entries, src_index, verdict, fp, keep);
    return ob2_check_cuda(cudaGetLastError(), "scatter_verdict_kernel launch");
}
