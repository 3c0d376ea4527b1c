// Micro-benchmark: the random-access ceiling the history dedup runs against.
// 1.25e7 uniformly random probes into a table far larger than L2 -- what does the memory system
// allow when every probe is (a) one 16-byte read, (b) a read + a 128-bit CAS on the same slot
// (an insert), (c) ... + the stores / atomics the two dedup protocols add, (d) a blind CAS?
// The dedup kernel cannot be faster than the variant that does its memory operations and nothing else.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a random_access.cu -o random_access
//   ./random_access [log2_slots=31] [slot_bytes=32]
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t mix(uint64_t x) {
    x ^= x >> 33; x *= 0xFF51AFD7ED558CCDull; x ^= x >> 33; x *= 0xC4CEB9FE1A85EC53ull; x ^= x >> 33;
    return x;
}
__device__ __forceinline__ void cas128(void *slot, uint64_t cl, uint64_t ch, uint64_t nl, uint64_t nh, uint64_t &ol, uint64_t &oh) {
    asm volatile("{\n.reg .b128 c, v, o;\nmov.b128 c, {%2, %3};\nmov.b128 v, {%4, %5};\n"
                 "atom.relaxed.gpu.global.cas.b128 o, [%6], c, v;\nmov.b128 {%0, %1}, o;\n}\n"
                 : "=l"(ol), "=l"(oh) : "l"(cl), "l"(ch), "l"(nl), "l"(nh), "l"(slot) : "memory");
}

// mode bits: 1 = read key first, 2 = CAS the key in, 4 = store born/winner (8 B) after a successful CAS,
//            8 = atomicMax on the stamp, 16 = the r1 protocol's two separate 4-byte stores instead of one 8-byte store
template <int MODE>
__global__ void __launch_bounds__(256) probe(char *table, uint64_t mask, int slot_bytes, int64_t n, uint64_t salt,
                                             uint8_t *__restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const ulonglong2 key = make_ulonglong2(mix(i + salt) | 1, mix(~i + salt));
    char *slot = table + (key.x & mask) * (uint64_t)slot_bytes;
    uint64_t lo = 0, hi = 0;
    uint8_t v = 0;
    if (MODE & 1) asm volatile("ld.global.cg.v2.u64 {%0, %1}, [%2];" : "=l"(lo), "=l"(hi) : "l"(slot));
    if (MODE & 2) {
        if (!(MODE & 1) || lo == 0) {
            uint64_t ol, oh;
            cas128(slot, 0, 0, key.x, key.y, ol, oh);
            v = ol == 0;
        }
    } else {
        v = (uint8_t)(lo + hi);
    }
    if ((MODE & 4) && v) {      // dependent on the CAS result, as in the real protocol
        if (MODE & 16) {
            reinterpret_cast<uint32_t *>(slot)[7] = (uint32_t)i;
            reinterpret_cast<uint32_t *>(slot)[6] = (uint32_t)salt;
        } else {
            reinterpret_cast<uint64_t *>(slot)[3] = ((uint64_t)i << 32) | (uint32_t)salt;
        }
    }
    if ((MODE & 8) && v) atomicMax(reinterpret_cast<unsigned long long *>(slot) + 2, (unsigned long long)(salt << 32 | (uint32_t)~i));
    out[i] = v;
}

int main(int argc, char **argv) {
    const int log2_slots = argc > 1 ? atoi(argv[1]) : 31;
    const int slot_bytes = argc > 2 ? atoi(argv[2]) : 32;
    const int64_t n = 12500000;
    const uint64_t n_slots = 1ull << log2_slots;
    const size_t bytes = n_slots * (size_t)slot_bytes;
    char *table;
    uint8_t *out;
    if (cudaMalloc(&table, bytes) != cudaSuccess || cudaMalloc(&out, n) != cudaSuccess) {
        printf("{\"error\": \"cudaMalloc of %zu bytes failed\"}\n", bytes);
        return 1;
    }
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const unsigned blocks = (unsigned)((n + 255) / 256);
    const char *names[] = {"read16", "read16+cas128", "read16+cas128+store8", "read16+cas128+2xstore4+atomicMax64 (r1 insert)",
                           "blind cas128", "blind cas128+store8 (r2 insert)"};
    printf("{\"table_bytes\": %zu, \"slot_bytes\": %d, \"probes\": %lld, \"variants\": {", bytes, slot_bytes, (long long)n);
    for (int variant = 0; variant < 6; ++variant) {
        float best = 1e9f;
        for (int rep = 0; rep < 5; ++rep) {
            cudaMemset(table, 0, bytes);            // every probe finds an empty slot (and L2 is flushed)
            cudaDeviceSynchronize();
            const uint64_t salt = 1000 * variant + rep + 1;
            cudaEventRecord(e0);
            switch (variant) {
                case 0: probe<1><<<blocks, 256>>>(table, n_slots - 1, slot_bytes, n, salt, out); break;
                case 1: probe<1 | 2><<<blocks, 256>>>(table, n_slots - 1, slot_bytes, n, salt, out); break;
                case 2: probe<1 | 2 | 4><<<blocks, 256>>>(table, n_slots - 1, slot_bytes, n, salt, out); break;
                case 3: probe<1 | 2 | 4 | 8 | 16><<<blocks, 256>>>(table, n_slots - 1, slot_bytes, n, salt, out); break;
                case 4: probe<2><<<blocks, 256>>>(table, n_slots - 1, slot_bytes, n, salt, out); break;
                case 5: probe<2 | 4><<<blocks, 256>>>(table, n_slots - 1, slot_bytes, n, salt, out); break;
            }
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
        }
        printf("%s\"%s\": {\"ms\": %.4f, \"Gprobes_per_s\": %.2f}", variant ? ", " : "", names[variant], best, n / best / 1e6);
    }
    printf("}, \"cuda\": \"%s\"}\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
