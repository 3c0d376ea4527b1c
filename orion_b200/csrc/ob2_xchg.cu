// Peer-memory exchange for the fingerprint-sharded history.
//
// With NCCL the dedup exchange is: bucket -> pad -> all_to_all(fp) ->
// all_to_all(key) -> dedup -> all_to_all(verdict) -> scatter.  Here every rank
// owns an INBOX in its own HBM, mapped into the other ranks through CUDA IPC,
// and the kernels do the transfers themselves over NVLink / NVSwitch:
//
//   send     the bucketing kernel stores each (fingerprint, order key) straight
//            into the owner's inbox -- the "all-to-all" happens store by store
//            while the kernel is still bucketing;
//   dedup    the owner dedups its inbox against its shard (ob2_set_dedup); the
//            verdicts stay in its region, lane by source;
//   collect  the source pulls its verdict lanes out of the owners' regions
//            (128-bit loads over NVLink) and scatters them into its keep mask.
//
// The caller orders the three phases across ranks with two stream-ordered
// barriers per batch (any tiny collective on the same stream): one after send,
// one after dedup.  No staging buffers, no zero-padding of bucket tails anywhere
// (counts travel instead, and the owner's set kernels read only what arrived),
// three collectives fewer per batch.
#include <cstdlib>
#include <cstring>
#include <new>

#include "ob2_internal.cuh"

#define OB2_XCHG_MAX_RANKS 64

namespace {

struct Region {           // layout of one rank's exchange region (same on every rank)
    size_t fp_off, key_off, cnt_off, verdict_off, total;
};

Region region_layout(int world, int64_t cap) {
    Region r;
    size_t off = 0;
    r.fp_off = off;
    off += (size_t)world * cap * 16;
    r.key_off = off;
    off += (size_t)world * cap * 4;
    r.verdict_off = off;
    off += ((size_t)world * cap + 255) / 256 * 256;
    r.cnt_off = off;
    off += 256;
    r.total = off;
    return r;
}

struct PeerTable {  // device-resident: base address of every rank's region
    char *base[OB2_XCHG_MAX_RANKS];
};

}  // namespace

struct ob2_xchg {
    int device, rank, world;
    int64_t cap;
    Region lay;
    char *region;          // my inbox region (peer-visible)
    char *peer[OB2_XCHG_MAX_RANKS];
    bool opened[OB2_XCHG_MAX_RANKS];
    PeerTable *d_peers;    // device copy of the table
    // local-only buffers
    uint32_t *out_src;     // [world][cap] local index of what I sent to each owner
    uint32_t *send_cnt;    // [world]
    uint32_t *scratch;     // [world * cap] dedup scratch
    // copy-engine variant: local staging of the bucketed batch and of the pulled verdicts
    uint64_t *stage_fp;    // [world][cap][2]
    uint32_t *stage_key;   // [world][cap]
    uint8_t *stage_verdict;  // [world * cap]
    bool peers_ready;
};

namespace {

constexpr int kSendPerThread = 4;
constexpr int kSendTile = OB2_BLOCK * kSendPerThread;  // 1024 fingerprints per tile
#ifndef OB2_SEND_GRID
#define OB2_SEND_GRID 296   // 2 blocks per SM
#endif
constexpr int kSendGrid = OB2_SEND_GRID;

// Bucket-and-push.  A block takes a tile of 1024 fingerprints, counting-sorts it by owner in
// shared memory, reserves one contiguous run per owner in that owner's inbox lane, and then
// writes the runs out with consecutive threads on consecutive addresses: every owner receives
// ~2 KB (8 ranks) of contiguous 16-byte stores per block instead of scattered single entries,
// which is what NVLink wants.
__global__ void __launch_bounds__(OB2_BLOCK)
xchg_send_kernel(int64_t n, const uint64_t *__restrict__ fp, uint32_t key_base, int rank, int world, int64_t cap,
                 const PeerTable *__restrict__ peers, size_t fp_off, size_t key_off, uint32_t *__restrict__ out_src,
                 uint32_t *__restrict__ send_cnt) {
    __shared__ uint32_t s_cnt[OB2_XCHG_MAX_RANKS];   // entries of this tile per owner
    __shared__ uint32_t s_off[OB2_XCHG_MAX_RANKS];   // where the owner's run starts in the staging arrays
    __shared__ uint32_t s_base[OB2_XCHG_MAX_RANKS];  // where it starts in the owner's inbox lane
    __shared__ ulonglong2 s_fp[kSendTile];
    __shared__ uint32_t s_idx[kSendTile];
    const int tid = threadIdx.x;
    // A capped grid walks the tiles: the kernel is NVLink bound (it saturates the links from a
    // fraction of the SMs), and it runs beside the next batch's sampling kernel -- flooding
    // every SM with its blocks would only take issue slots from that one.
    const int64_t n_tiles = (n + kSendTile - 1) / kSendTile;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    if (tid < world) s_cnt[tid] = 0;
    __syncthreads();
    const int64_t tile0 = tile * kSendTile;
    ulonglong2 key[kSendPerThread];
    int owner[kSendPerThread];
    uint32_t local_rank[kSendPerThread];
#pragma unroll
    for (int q = 0; q < kSendPerThread; ++q) {
        const int64_t i = tile0 + q * OB2_BLOCK + tid;
        owner[q] = -1;
        local_rank[q] = 0;
        key[q] = make_ulonglong2(0, 0);
        if (i < n) {
            key[q] = __ldg(reinterpret_cast<const ulonglong2 *>(fp) + i);
            if (key[q].x != 0) {
                owner[q] = ob2_owner_of(key[q].y, (uint32_t)world);
                local_rank[q] = atomicAdd(&s_cnt[owner[q]], 1u);
            }
        }
    }
    __syncthreads();
    if (tid == 0) {
        uint32_t off = 0;
        for (int o = 0; o < world; ++o) {
            s_off[o] = off;
            off += s_cnt[o];
        }
    }
    if (tid < world) s_base[tid] = s_cnt[tid] ? atomicAdd(&send_cnt[tid], s_cnt[tid]) : 0;
    __syncthreads();
#pragma unroll
    for (int q = 0; q < kSendPerThread; ++q) {
        if (owner[q] >= 0) {
            const uint32_t slot = s_off[owner[q]] + local_rank[q];
            s_fp[slot] = key[q];
            s_idx[slot] = (uint32_t)(q * OB2_BLOCK + tid);
        }
    }
    __syncthreads();
    const uint32_t total = s_off[world - 1] + s_cnt[world - 1];
    for (uint32_t t = tid; t < total; t += OB2_BLOCK) {
        int o = 0;
        while (o + 1 < world && t >= s_off[o + 1]) ++o;           // which owner's run holds staged entry t
        const int64_t pos = (int64_t)s_base[o] + (t - s_off[o]);
        if (pos < cap) {
            char *dst = peers->base[o];                            // the OWNER's region: stores over NVLink
            const int64_t lane = (int64_t)rank * cap + pos;        // my lane of its inbox
            const uint32_t local = (uint32_t)(tile0 + s_idx[t]);
            reinterpret_cast<ulonglong2 *>(dst + fp_off)[lane] = s_fp[t];
            reinterpret_cast<uint32_t *>(dst + key_off)[lane] = key_base + local;
            out_src[(int64_t)o * cap + pos] = local;
        }
    }
    __syncthreads();  // the staging arrays are reused by the next tile
    }
}

// tell every owner how many entries I put in my lane of its inbox
__global__ void xchg_counts_kernel(int rank, int world, const PeerTable *__restrict__ peers, size_t cnt_off,
                                   const uint32_t *__restrict__ send_cnt) {
    const int o = threadIdx.x;
    if (o < world) reinterpret_cast<uint32_t *>(peers->base[o] + cnt_off)[rank] = send_cnt[o];
}

// source side: PULL the verdicts of what I sent to owner `blockIdx.y` out of that owner's
// region (16 verdicts per 128-bit load over NVLink) and scatter them into my keep mask
__global__ void __launch_bounds__(OB2_BLOCK)
xchg_collect_kernel(int rank, int64_t cap, const uint32_t *__restrict__ send_cnt, const uint32_t *__restrict__ out_src,
                    const PeerTable *__restrict__ peers, size_t verdict_off, uint8_t *__restrict__ keep) {
    const int owner = blockIdx.y;
    const int64_t used = min((int64_t)send_cnt[owner], cap);
    const uint8_t *src = reinterpret_cast<const uint8_t *>(peers->base[owner] + verdict_off) + (int64_t)rank * cap;
    const uint32_t *idx = out_src + (int64_t)owner * cap;
    for (int64_t p0 = ((int64_t)blockIdx.x * OB2_BLOCK + threadIdx.x) * 16; p0 < used; p0 += (int64_t)gridDim.x * OB2_BLOCK * 16) {
        const uint4 v = *reinterpret_cast<const uint4 *>(src + p0);   // cap is a multiple of 16: aligned
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int q = 0; q < 16; ++q)
            if (p0 + q < used) keep[idx[p0 + q]] = (uint8_t)((w[q >> 2] >> (8 * (q & 3))) & 0xFFu);
    }
}

// sticky overflow flag: any (source = me, owner) lane that received more than it can hold
__global__ void xchg_flag_overflow_kernel(int world, int64_t cap, const uint32_t *__restrict__ send_cnt,
                                          uint32_t *__restrict__ flag) {
    const int o = threadIdx.x;
    if (o < world && (int64_t)send_cnt[o] > cap) atomicMax(flag, 1u);
}

int upload_peers(ob2_xchg *x) {
    PeerTable t;
    for (int r = 0; r < OB2_XCHG_MAX_RANKS; ++r) t.base[r] = r < x->world ? x->peer[r] : nullptr;
    OB2_CUDA(cudaMemcpy(x->d_peers, &t, sizeof(t), cudaMemcpyHostToDevice));
    x->peers_ready = true;
    return OB2_OK;
}

unsigned lane_blocks(int64_t cap) {
    const int64_t b = (cap + OB2_BLOCK - 1) / OB2_BLOCK;
    return (unsigned)(b < 1 ? 1 : (b > 1024 ? 1024 : b));
}

int check_ready(const ob2_xchg *x, const char *who) {
    if (x == nullptr || !x->peers_ready) {
        ob2_set_error("%s: exchange has no peer table yet (call ob2_xchg_open_peers / ob2_xchg_set_peers)", who);
        return OB2_EINVAL;
    }
    return OB2_OK;
}

}  // namespace

extern "C" int ob2_xchg_create(int32_t rank, int32_t world, int64_t cap, ob2_xchg **out) {
    if (out == nullptr || world < 1 || world > OB2_XCHG_MAX_RANKS || rank < 0 || rank >= world || cap < 1 ||
        (int64_t)world * cap >= 0xFFFFFFFFll) {
        ob2_set_error("ob2_xchg_create: bad arguments (rank=%d world=%d cap=%lld)", rank, world, (long long)cap);
        return OB2_EINVAL;
    }
    *out = nullptr;
    ob2_xchg *x = new (std::nothrow) ob2_xchg();
    if (x == nullptr) {
        ob2_set_error("out of host memory");
        return OB2_ENOMEM;
    }
    x->rank = rank;
    x->world = world;
    cap = (cap + 15) / 16 * 16;  // verdict lanes are read back 16 bytes at a time
    x->cap = cap;
    x->lay = region_layout(world, cap);
    x->peers_ready = false;
    for (int r = 0; r < OB2_XCHG_MAX_RANKS; ++r) {
        x->peer[r] = nullptr;
        x->opened[r] = false;
    }
    int rc = ob2_check_cuda(cudaGetDevice(&x->device), "cudaGetDevice");
    const size_t lanes = (size_t)world * cap;
    if (rc == OB2_OK) rc = ob2_check_cuda(cudaMalloc(&x->region, x->lay.total), "cudaMalloc(exchange region)");
    if (rc == OB2_OK) rc = ob2_check_cuda(cudaMemset(x->region, 0, x->lay.total), "cudaMemset(exchange region)");
    if (rc == OB2_OK) rc = ob2_check_cuda(cudaMalloc(&x->d_peers, sizeof(PeerTable)), "cudaMalloc(peer table)");
    if (rc == OB2_OK) rc = ob2_check_cuda(cudaMalloc(&x->out_src, lanes * 4), "cudaMalloc(out_src)");
    if (rc == OB2_OK) rc = ob2_check_cuda(cudaMalloc(&x->send_cnt, 256), "cudaMalloc(send_cnt)");
    if (rc == OB2_OK) rc = ob2_check_cuda(cudaMalloc(&x->scratch, lanes * 4), "cudaMalloc(scratch)");
    if (rc == OB2_OK) rc = ob2_check_cuda(cudaMalloc(&x->stage_fp, lanes * 16), "cudaMalloc(stage_fp)");
    if (rc == OB2_OK) rc = ob2_check_cuda(cudaMalloc(&x->stage_key, lanes * 4), "cudaMalloc(stage_key)");
    if (rc == OB2_OK) rc = ob2_check_cuda(cudaMalloc(&x->stage_verdict, (lanes + 255) / 256 * 256), "cudaMalloc(stage_verdict)");
    if (rc != OB2_OK) {
        ob2_xchg_destroy(x);
        return rc;
    }
    x->peer[rank] = x->region;
    if (world == 1) rc = upload_peers(x);
    if (rc != OB2_OK) {
        ob2_xchg_destroy(x);
        return rc;
    }
    *out = x;
    return OB2_OK;
}

extern "C" void ob2_xchg_destroy(ob2_xchg *x) {
    if (x == nullptr) return;
    for (int r = 0; r < x->world; ++r)
        if (x->opened[r] && x->peer[r] != nullptr) cudaIpcCloseMemHandle(x->peer[r]);
    if (x->region) cudaFree(x->region);
    if (x->d_peers) cudaFree(x->d_peers);
    if (x->out_src) cudaFree(x->out_src);
    if (x->send_cnt) cudaFree(x->send_cnt);
    if (x->scratch) cudaFree(x->scratch);
    if (x->stage_fp) cudaFree(x->stage_fp);
    if (x->stage_key) cudaFree(x->stage_key);
    if (x->stage_verdict) cudaFree(x->stage_verdict);
    delete x;
}

extern "C" int ob2_xchg_ipc_handle(const ob2_xchg *x, void *handle64) {
    if (x == nullptr || handle64 == nullptr) {
        ob2_set_error("ob2_xchg_ipc_handle: bad arguments");
        return OB2_EINVAL;
    }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaIpcMemHandle_t h;
    OB2_CUDA(cudaIpcGetMemHandle(&h, x->region));
    memcpy(handle64, &h, 64);
    return OB2_OK;
}

extern "C" int ob2_xchg_open_peers(ob2_xchg *x, const void *handles) {
    if (x == nullptr || handles == nullptr) {
        ob2_set_error("ob2_xchg_open_peers: bad arguments");
        return OB2_EINVAL;
    }
    for (int r = 0; r < x->world; ++r) {
        if (r == x->rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, static_cast<const char *>(handles) + 64 * (size_t)r, 64);
        void *p = nullptr;
        OB2_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        x->peer[r] = static_cast<char *>(p);
        x->opened[r] = true;
    }
    return upload_peers(x);
}

extern "C" int ob2_xchg_set_peers(ob2_xchg *x, void *const *regions) {
    // same-process variant (several ranks in one process, or tests on one GPU)
    if (x == nullptr || regions == nullptr) {
        ob2_set_error("ob2_xchg_set_peers: bad arguments");
        return OB2_EINVAL;
    }
    for (int r = 0; r < x->world; ++r)
        if (r != x->rank) x->peer[r] = static_cast<char *>(regions[r]);
    return upload_peers(x);
}

extern "C" void *ob2_xchg_region(const ob2_xchg *x) { return x ? x->region : nullptr; }

extern "C" int ob2_xchg_send(ob2_xchg *x, int64_t n, const uint64_t *fp, uint32_t key_base, void *stream) {
    int rc = check_ready(x, "ob2_xchg_send");
    if (rc != OB2_OK) return rc;
    if (n < 0 || (n > 0 && fp == nullptr) || n >= 0xFFFFFFFFll) {
        ob2_set_error("ob2_xchg_send: bad arguments");
        return OB2_EINVAL;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    OB2_CUDA(cudaMemsetAsync(x->send_cnt, 0, sizeof(uint32_t) * x->world, st));
    if (n > 0) {
        int64_t blocks = (n + kSendTile - 1) / kSendTile;
        static const int grid_cap = [] {  // OB2_SEND_GRID in the environment overrides the built-in cap (experiments)
            const char *e = getenv("OB2_SEND_GRID");
            const int v = e ? atoi(e) : 0;
            return v > 0 ? v : kSendGrid;
        }();
        if (blocks > grid_cap) blocks = grid_cap;
        xchg_send_kernel<<<(unsigned)blocks, OB2_BLOCK, 0, st>>>(n, fp, key_base, x->rank, x->world, x->cap, x->d_peers,
                                                                  x->lay.fp_off, x->lay.key_off, x->out_src, x->send_cnt);
        rc = ob2_check_cuda(cudaGetLastError(), "xchg_send_kernel launch");
        if (rc != OB2_OK) return rc;
    }
    xchg_counts_kernel<<<1, OB2_XCHG_MAX_RANKS, 0, st>>>(x->rank, x->world, x->d_peers, x->lay.cnt_off, x->send_cnt);
    return ob2_check_cuda(cudaGetLastError(), "xchg_counts_kernel launch");
}

extern "C" int ob2_xchg_dedup(ob2_xchg *x, ob2_set *set, int32_t insert_only, void *stream) {
    int rc = check_ready(x, "ob2_xchg_dedup");
    if (rc != OB2_OK) return rc;
    if (set == nullptr) {
        ob2_set_error("ob2_xchg_dedup: null set");
        return OB2_EINVAL;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    uint64_t *inbox_fp = reinterpret_cast<uint64_t *>(x->region + x->lay.fp_off);
    const uint32_t *inbox_key = reinterpret_cast<const uint32_t *>(x->region + x->lay.key_off);
    const uint32_t *cnt = reinterpret_cast<const uint32_t *>(x->region + x->lay.cnt_off);
    // Store-based sends leave the tails of the lanes unwritten: the set kernels are launched lane by lane
    // and read only the cnt[src] entries that arrived (no padding pass).  Bit 1 of insert_only: the senders
    // shipped whole zero-padded lanes (ob2_xchg_send_copy), a flat launch is fine.
    const uint32_t *lane_cnt = (insert_only & 2) ? nullptr : cnt;
    const int64_t lanes = (int64_t)x->world * x->cap;
    // The kernels run over the padded lanes, but only the entries really received can be
    // inserted: when the cheap host bound (which counts the padding) would trip the 7/8 load
    // limit, read the per-source counts and reserve against the real inbox population.
    int64_t real = lanes;
    if (ob2_set_would_trip(set, lanes)) {
        uint32_t h_cnt[OB2_XCHG_MAX_RANKS];
        OB2_CUDA(cudaMemcpyAsync(h_cnt, cnt, sizeof(uint32_t) * x->world, cudaMemcpyDeviceToHost, st));
        OB2_CUDA(cudaStreamSynchronize(st));
        real = 0;
        for (int r = 0; r < x->world; ++r) real += (int64_t)h_cnt[r] < x->cap ? (int64_t)h_cnt[r] : x->cap;
    }
    if (insert_only & 1) return ob2_set_insert_padded(set, lanes, real, inbox_fp, stream, lane_cnt, x->cap);
    // verdicts stay in MY region, lane by source; the sources pull them (ob2_xchg_collect)
    return ob2_set_dedup_padded(set, lanes, real, inbox_fp, inbox_key, x->scratch,
                                reinterpret_cast<uint8_t *>(x->region + x->lay.verdict_off), stream, lane_cnt, x->cap);
}

extern "C" int ob2_xchg_collect(ob2_xchg *x, int64_t n, uint8_t *keep, void *stream) {
    int rc = check_ready(x, "ob2_xchg_collect");
    if (rc != OB2_OK) return rc;
    if (n < 0 || keep == nullptr) {
        ob2_set_error("ob2_xchg_collect: bad arguments");
        return OB2_EINVAL;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    OB2_CUDA(cudaMemsetAsync(keep, 0, (size_t)n, st));     // padding entries and bucket overflow stay 0
    const dim3 grid(lane_blocks((x->cap + 15) / 16), (unsigned)x->world);
    xchg_collect_kernel<<<grid, OB2_BLOCK, 0, st>>>(x->rank, x->cap, x->send_cnt, x->out_src, x->d_peers,
                                                    x->lay.verdict_off, keep);
    return ob2_check_cuda(cudaGetLastError(), "xchg_collect_kernel launch");
}

// ---- copy-engine variant of send / collect: the same inboxes, but the bytes cross NVLink as
// cudaMemcpyAsync transfers (DMA engines, no SM involved) instead of remote stores from a kernel.
// The store-based send is NVLink bound yet holds SM slots -- and store-queue space -- that the
// next batch's store-bound sampling kernel wants; DMA transfers run entirely beside it.
extern "C" int ob2_xchg_send_copy(ob2_xchg *x, int64_t n, const uint64_t *fp, uint32_t key_base, void *stream) {
    int rc = check_ready(x, "ob2_xchg_send_copy");
    if (rc != OB2_OK) return rc;
    if (n < 0 || (n > 0 && fp == nullptr) || n >= 0xFFFFFFFFll) {
        ob2_set_error("ob2_xchg_send_copy: bad arguments");
        return OB2_EINVAL;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // bucket by owner into local staging, tails zero-padded (counts land in send_cnt)
    rc = ob2_bucket_by_owner(n, fp, key_base, x->world, x->cap, x->stage_fp, x->stage_key, x->out_src, x->send_cnt, stream);
    if (rc != OB2_OK) return rc;
    for (int o = 0; o < x->world; ++o) {
        char *dst = x->peer[o];
        const size_t lane = (size_t)x->rank * x->cap;
        OB2_CUDA(cudaMemcpyAsync(dst + x->lay.fp_off + lane * 16, x->stage_fp + (size_t)o * x->cap * 2, (size_t)x->cap * 16,
                                 cudaMemcpyDefault, st));
        OB2_CUDA(cudaMemcpyAsync(dst + x->lay.key_off + lane * 4, x->stage_key + (size_t)o * x->cap, (size_t)x->cap * 4,
                                 cudaMemcpyDefault, st));
    }
    xchg_counts_kernel<<<1, OB2_XCHG_MAX_RANKS, 0, st>>>(x->rank, x->world, x->d_peers, x->lay.cnt_off, x->send_cnt);
    return ob2_check_cuda(cudaGetLastError(), "xchg_counts_kernel launch");
}

extern "C" int ob2_xchg_collect_copy(ob2_xchg *x, int64_t n, uint8_t *keep, void *stream) {
    int rc = check_ready(x, "ob2_xchg_collect_copy");
    if (rc != OB2_OK) return rc;
    if (n < 0 || keep == nullptr) {
        ob2_set_error("ob2_xchg_collect_copy: bad arguments");
        return OB2_EINVAL;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    OB2_CUDA(cudaMemsetAsync(keep, 0, (size_t)n, st));  // padding entries and bucket overflow stay 0
    for (int o = 0; o < x->world; ++o)  // pull my verdict lane out of every owner's region
        OB2_CUDA(cudaMemcpyAsync(x->stage_verdict + (size_t)o * x->cap, x->peer[o] + x->lay.verdict_off + (size_t)x->rank * x->cap,
                                 (size_t)x->cap, cudaMemcpyDefault, st));
    return ob2_scatter_verdict((int64_t)x->world * x->cap, x->out_src, x->stage_verdict, x->stage_fp, keep, stream);
}

extern "C" int64_t ob2_xchg_cap(const ob2_xchg *x) { return x ? x->cap : 0; }

extern "C" int ob2_xchg_flag_overflow(const ob2_xchg *x, uint32_t *flag, void *stream) {
    if (x == nullptr || flag == nullptr) {
        ob2_set_error("ob2_xchg_flag_overflow: bad arguments");
        return OB2_EINVAL;
    }
    xchg_flag_overflow_kernel<<<1, OB2_XCHG_MAX_RANKS, 0, static_cast<cudaStream_t>(stream)>>>(x->world, x->cap, x->send_cnt, flag);
    return ob2_check_cuda(cudaGetLastError(), "xchg_flag_overflow_kernel launch");
}

extern "C" int ob2_xchg_send_counts(const ob2_xchg *x, uint32_t *counts_host, void *stream) {
    // synchronising read of how many entries went to each owner (overflow check: count > cap)
    if (x == nullptr || counts_host == nullptr) {
        ob2_set_error("ob2_xchg_send_counts: bad arguments");
        return OB2_EINVAL;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    OB2_CUDA(cudaMemcpyAsync(counts_host, x->send_cnt, sizeof(uint32_t) * x->world, cudaMemcpyDeviceToHost, st));
    return ob2_check_cuda(cudaStreamSynchronize(st), "cudaStreamSynchronize");
}
