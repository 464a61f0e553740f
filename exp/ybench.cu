// exp/ybench.cu -- microbenchmarks behind the sort architecture:
//   (a) 2^16-bin histogram of the key prefix with global reductions
//   (b) one-pass partition by the 16-bit prefix with global atomics
//   (c) per-bucket LSD radix sort of the remaining 24 bits in shared memory
// on N random 40-bit keys with the (1 - x) density of min(dir, rev).
// Not part of the product; build: nvcc -arch=sm_100a -O3 -o ybench ybench.cu
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                  \
    do {                                                                       \
        cudaError_t e = (x);                                                   \
        if (e != cudaSuccess) {                                                \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, \
                   __LINE__);                                                  \
            exit(1);                                                           \
        }                                                                      \
    } while (0)

__host__ __device__ inline uint64_t mix(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

constexpr int KEY_BITS = 40;
constexpr int FINE_BITS = 16;
constexpr int LOW_BITS = KEY_BITS - FINE_BITS;  // 24
constexpr uint32_t N_FINE = 1u << FINE_BITS;

__global__ void gen_keys(uint64_t* keys, uint64_t n, uint64_t seed, int dup) {
    uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
    if (i >= n) return;
    // dup: every key value appears ~dup times (like 17 near-identical genomes)
    uint64_t j = i % (n / dup + 1);
    uint64_t a = mix(j ^ seed) >> (64 - KEY_BITS);
    uint64_t b = mix(a + seed) >> (64 - KEY_BITS);
    keys[i] = a < b ? a : b;
}

__global__ void hist_red(const uint64_t* __restrict__ keys, uint64_t n,
                         uint32_t* __restrict__ hist) {
    uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
    for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < n;
            i += stride) {
        uint32_t f = uint32_t(keys[i] >> LOW_BITS);
        atomicAdd(hist + f, 1u);
    }
}

// smem variant: 2^15 bins of u32 (128 KB) on the top 15 bits
__global__ void __launch_bounds__(1024)
hist_smem15(const uint64_t* __restrict__ keys, uint64_t n,
            uint32_t* __restrict__ hist) {
    extern __shared__ uint32_t sh[];
    for (int i = threadIdx.x; i < 32768; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
    for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < n;
            i += stride) {
        uint32_t f = uint32_t(keys[i] >> (LOW_BITS + 1));
        atomicAdd(sh + f, 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 32768; i += blockDim.x) {
        uint32_t v = sh[i];
        if (v) atomicAdd(hist + i, v);
    }
}

__global__ void __launch_bounds__(1024)
scan_fine(const uint32_t* __restrict__ hist, uint32_t* __restrict__ base,
          uint32_t* __restrict__ cursor) {
    // single CTA exclusive scan of 65536 counts
    __shared__ uint32_t part[1024];
    const int per = N_FINE / 1024;
    uint32_t s = 0;
    for (int j = 0; j < per; j++) s += hist[threadIdx.x * per + j];
    part[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t run = 0;
        for (int j = 0; j < 1024; j++) {
            uint32_t v = part[j];
            part[j] = run;
            run += v;
        }
    }
    __syncthreads();
    uint32_t run = part[threadIdx.x];
    for (int j = 0; j < per; j++) {
        int b = threadIdx.x * per + j;
        base[b] = run;
        cursor[b] = run;
        run += hist[b];
    }
    if (threadIdx.x == 1023) base[N_FINE] = run;
}

template <int ITEMS>
__global__ void __launch_bounds__(256)
part_atomic(const uint64_t* __restrict__ keys, uint64_t n,
            uint32_t* __restrict__ cursor, uint64_t* __restrict__ out) {
    uint64_t base = (blockIdx.x * uint64_t(blockDim.x)) * ITEMS + threadIdx.x;
    uint64_t k[ITEMS];
    uint32_t pos[ITEMS];
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
        uint64_t i = base + uint64_t(j) * blockDim.x;
        k[j] = i < n ? keys[i] : ~0ull;
    }
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
        uint64_t i = base + uint64_t(j) * blockDim.x;
        if (i < n) pos[j] = atomicAdd(cursor + uint32_t(k[j] >> LOW_BITS), 1u);
    }
#pragma unroll
    for (int j = 0; j < ITEMS; j++) {
        uint64_t i = base + uint64_t(j) * blockDim.x;
        if (i < n) {
            out[pos[j]] = ((k[j] & ((1ull << LOW_BITS) - 1)) << 32) | uint32_t(i);
        }
    }
}

// ---- in-smem bucket sort ---------------------------------------------------------
constexpr int BS_THREADS = 512;
constexpr int BS_WARPS = BS_THREADS / 32;
constexpr int BS_ITEMS = 8;
constexpr int CAP = BS_THREADS * BS_ITEMS;  // 4096

struct BSmem {
    uint64_t buf[2][CAP];
    uint32_t whist[BS_WARPS][256];
    uint32_t dbase[256];
    uint32_t wsum[BS_WARPS];
    uint32_t masks[BS_WARPS][256];
};

// MODE 0: match.any   1: 8 ballots   2: atomicOr masks in shared memory
template <int PASSES, int MODE>
__global__ void __launch_bounds__(BS_THREADS, 2)
bucket_sort(const uint64_t* __restrict__ in, uint64_t* __restrict__ out,
            const uint2* __restrict__ units, int n_units) {
    extern __shared__ __align__(16) unsigned char raw[];
    BSmem& sm = *reinterpret_cast<BSmem*>(raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lt = (1u << lane) - 1;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
        const uint2 un = units[u];
        const uint32_t start = un.x, n = un.y - un.x;
        if (MODE == 2) {
#pragma unroll
            for (int i = 0; i < 8; i++) sm.masks[warp][lane + 32 * i] = 0;
        }
        for (uint32_t j = tid; j < n; j += BS_THREADS) sm.buf[0][j] = in[start + j];
        __syncthreads();
        int cur = 0;
#pragma unroll 1
        for (int pass = 0; pass < PASSES; pass++) {
            const int shift = 32 + 8 * pass;
#pragma unroll
            for (int i = 0; i < 8; i++) sm.whist[warp][lane + 32 * i] = 0;
            __syncwarp();
            uint64_t rec[BS_ITEMS];
            uint32_t rank[BS_ITEMS], dig[BS_ITEMS];
            const uint32_t wbase = warp * (32 * BS_ITEMS);
#pragma unroll
            for (int i = 0; i < BS_ITEMS; i++) {
                uint32_t idx = wbase + i * 32 + lane;
                bool valid = idx < n;
                rec[i] = valid ? sm.buf[cur][idx] : 0;
                uint32_t d = valid ? uint32_t(rec[i] >> shift) & 255u : 256u;
                dig[i] = d;
                if (wbase + i * 32 < n) {  // warp-uniform
                    uint32_t peers;
                    if (MODE == 0) {
                        peers = __match_any_sync(0xffffffffu, d);
                    } else if (MODE == 1) {
                        peers = 0xffffffffu;
#pragma unroll
                        for (int b = 0; b < 9; b++) {
                            uint32_t v = __ballot_sync(0xffffffffu, (d >> b) & 1);
                            peers &= ((d >> b) & 1) ? v : ~v;
                        }
                    } else {
                        uint32_t* mm = sm.masks[warp];
                        if (d < 256) atomicOr(mm + d, 1u << lane);
                        __syncwarp();
                        peers = d < 256 ? mm[d] : (1u << lane);
                        __syncwarp();
                        if (d < 256) mm[d] = 0;
                    }
                    int leader = __ffs(peers) - 1;
                    uint32_t old = 0;
                    if (lane == leader && d < 256) {
                        old = sm.whist[warp][d];
                        sm.whist[warp][d] = old + __popc(peers);
                    }
                    old = __shfl_sync(0xffffffffu, old, leader);
                    rank[i] = old + __popc(peers & lt);
                    __syncwarp();
                }
            }
            __syncthreads();
            if (tid < 256) {
                uint32_t c = 0;
#pragma unroll
                for (int w = 0; w < BS_WARPS; w++) {
                    uint32_t t = sm.whist[w][tid];
                    sm.whist[w][tid] = c;
                    c += t;
                }
                uint32_t incl = c;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o) incl += v;
                }
                if (lane == 31) sm.wsum[warp] = incl;
                sm.dbase[tid] = incl - c;  // warp-local exclusive
            }
            __syncthreads();
            if (tid < 256) {
                uint32_t b = 0;
                for (int w = 0; w < warp; w++) b += sm.wsum[w];
                sm.dbase[tid] += b;
            }
            __syncthreads();
#pragma unroll
            for (int i = 0; i < BS_ITEMS; i++) {
                uint32_t d = dig[i];
                if (d < 256) {
                    sm.buf[cur ^ 1][sm.dbase[d] + sm.whist[warp][d] + rank[i]] = rec[i];
                }
            }
            __syncthreads();
            cur ^= 1;
        }
        for (uint32_t j = tid; j < n; j += BS_THREADS) out[start + j] = sm.buf[cur][j];
        __syncthreads();
    }
}

static float time_ms(cudaEvent_t a, cudaEvent_t b) {
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    return ms;
}

int main(int argc, char** argv) {
    uint64_t n = argc > 1 ? strtoull(argv[1], 0, 10) : 56000000ull;
    int reps = argc > 2 ? atoi(argv[2]) : 5;
    int dup = argc > 3 ? atoi(argv[3]) : 17;
    int verify = argc > 4 ? atoi(argv[4]) : 1;
    uint64_t *keys, *part, *sorted;
    uint32_t *hist, *base, *cursor;
    uint2* units_d;
    CK(cudaMalloc(&keys, n * 8));
    CK(cudaMalloc(&part, n * 8));
    CK(cudaMalloc(&sorted, n * 8));
    CK(cudaMalloc(&hist, (N_FINE + 1) * 4));
    CK(cudaMalloc(&base, (N_FINE + 1) * 4));
    CK(cudaMalloc(&cursor, (N_FINE + 1) * 4));
    CK(cudaMalloc(&units_d, N_FINE * sizeof(uint2)));
    uint8_t* flush;
    CK(cudaMalloc(&flush, 512u << 20));
    gen_keys<<<unsigned((n + 255) / 256), 256>>>(keys, n, 12345, dup);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    CK(cudaFuncSetAttribute(hist_smem15, cudaFuncAttributeMaxDynamicSharedMemorySize,
                            32768 * 4));
#define SETATTR(P, M) CK(cudaFuncSetAttribute(bucket_sort<P, M>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(BSmem))))
    SETATTR(0, 0); SETATTR(1, 0); SETATTR(3, 0); SETATTR(3, 1); SETATTR(3, 2);
    std::vector<uint32_t> h_base(N_FINE + 1);
    std::vector<uint2> units;
    for (int r = 0; r < reps; r++) {
        CK(cudaMemset(flush, r, 512u << 20));
        CK(cudaMemset(hist, 0, (N_FINE + 1) * 4));
        cudaEventRecord(e0);
        hist_red<<<148 * 16, 256>>>(keys, n, hist);
        cudaEventRecord(e1);
        CK(cudaDeviceSynchronize());
        printf("hist_red(2^16 global RED)  %.3f ms\n", time_ms(e0, e1));

        CK(cudaMemset(flush, r, 512u << 20));
        CK(cudaMemset(cursor, 0, (N_FINE + 1) * 4));
        cudaEventRecord(e0);
        hist_smem15<<<148, 1024, 32768 * 4>>>(keys, n, cursor);
        cudaEventRecord(e1);
        CK(cudaDeviceSynchronize());
        printf("hist_smem15(2^15 smem)     %.3f ms\n", time_ms(e0, e1));

        cudaEventRecord(e0);
        scan_fine<<<1, 1024>>>(hist, base, cursor);
        cudaEventRecord(e1);
        CK(cudaDeviceSynchronize());
        printf("scan_fine                  %.3f ms\n", time_ms(e0, e1));

        CK(cudaMemset(flush, r, 512u << 20));
        cudaEventRecord(e0);
        part_atomic<8><<<unsigned((n + 2047) / 2048), 256>>>(keys, n, cursor, part);
        cudaEventRecord(e1);
        CK(cudaDeviceSynchronize());
        float ms = time_ms(e0, e1);
        printf("part_atomic<8>             %.3f ms  (%.1f GB/s on 16 B/rec)\n", ms,
               n * 16.0 / ms / 1e6);

        if (r == 0) {
            CK(cudaMemcpy(h_base.data(), base, (N_FINE + 1) * 4, cudaMemcpyDeviceToHost));
            uint32_t s = 0, big = 0;
            for (uint32_t f = 0; f < N_FINE;) {
                uint32_t e = f + 1;
                while (e < N_FINE && h_base[e + 1] - h_base[f] <= uint32_t(CAP)) e++;
                if (h_base[e] - h_base[f] > uint32_t(CAP)) {
                    big++;  // oversized fine bin: split arbitrarily (bench only)
                    for (uint32_t p = h_base[f]; p < h_base[e]; p += CAP) {
                        units.push_back(make_uint2(p, std::min(p + CAP, h_base[e])));
                    }
                } else {
                    units.push_back(make_uint2(h_base[f], h_base[e]));
                }
                f = e;
                s++;
            }
            printf("units: %zu (oversized fine bins: %u), mean fill %.1f of %d\n",
                   units.size(), big, double(n) / units.size(), CAP);
            CK(cudaMemcpy(units_d, units.data(), units.size() * sizeof(uint2),
                          cudaMemcpyHostToDevice));
        }
#define RUN(P, M)                                                                  \
        CK(cudaMemset(flush, r, 512u << 20));                                      \
        cudaEventRecord(e0);                                                       \
        bucket_sort<P, M><<<148 * 2, BS_THREADS, sizeof(BSmem)>>>(part, sorted, units_d, int(units.size())); \
        cudaEventRecord(e1);                                                       \
        CK(cudaDeviceSynchronize());                                               \
        ms = time_ms(e0, e1);                                                      \
        printf("bucket_sort<%d passes, mode %d>  %.3f ms\n", P, M, ms);
        RUN(0, 0) RUN(1, 0) RUN(3, 1) RUN(3, 2) RUN(3, 0)
        CK(cudaGetLastError());
    }
    if (verify) {
        std::vector<uint64_t> hp(n), hs(n);
        CK(cudaMemcpy(hp.data(), part, n * 8, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(hs.data(), sorted, n * 8, cudaMemcpyDeviceToHost));
        uint64_t bad = 0;
        for (const uint2& u : units) {
            uint64_t a = 0, b = 0;
            for (uint32_t i = u.x; i < u.y; i++) {
                a += mix(hp[i]);
                b += mix(hs[i]);
                if (i > u.x && (hs[i] >> 32) < (hs[i - 1] >> 32)) bad++;
            }
            if (a != b) bad++;
        }
        printf("verify: %llu bad of %llu\n", (unsigned long long)bad,
               (unsigned long long)n);
    }
    return 0;
}
