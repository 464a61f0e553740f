// Throughput of match.any vs a ballot-based equivalent for 8-bit digits (sm_100a).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t match_ballot8(uint32_t d) {
    uint32_t peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < 8; b++) {
        const bool bit = (d >> b) & 1;
        const uint32_t m = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? m : ~m;
    }
    return peers;
}

template <int MODE, int ILP>
__global__ void __launch_bounds__(512, 2) bench(uint32_t* out, int iters, uint32_t seed) {
    uint32_t x[ILP], acc = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) x[i] = (threadIdx.x * 2654435761u + i * 40503u + seed) >> 7;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) {
            uint32_t d = x[i] & 255u;
            uint32_t p = MODE == 0 ? __match_any_sync(0xffffffffu, d) : match_ballot8(d);
            acc += __popc(p);
            x[i] = x[i] * 1664525u + 1013904223u + (p & 1);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int MODE, int ILP>
void run(const char* name) {
    uint32_t* out;
    const int grid = 148 * 2, iters = 2000;
    cudaMalloc(&out, grid * 512 * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    bench<MODE, ILP><<<grid, 512>>>(out, 10, 1);
    cudaEventRecord(e0);
    bench<MODE, ILP><<<grid, 512>>>(out, iters, 1);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    double warp_ops = double(grid) * 16 * iters * ILP;
    printf("%-28s ILP=%d  %.3f ms  %.2f G warp-match/s  (%.1f SM-cycles per warp-match @1.965GHz)\n",
           name, ILP, ms, warp_ops / ms / 1e6, ms * 1e-3 * 1.965e9 * 148 / warp_ops);
    cudaFree(out);
}

int main() {
    run<0, 1>("match.any");
    run<0, 4>("match.any");
    run<0, 8>("match.any");
    run<1, 1>("8 ballots");
    run<1, 4>("8 ballots");
    run<1, 8>("8 ballots");
    return 0;
}
