// scan_kernels.cu -- pack (ASCII -> 2 bit + N mask), digit histograms of all
// canonical k-mers, histogram scan.  sm_100a.
#include "kernels.h"

namespace npge_b200 {

// ---- pack_2bit --------------------------------------------------------------
// One thread packs 32 bases: two 16-byte loads -> two u32 of 2-bit codes and one
// u32 of N bits.  Replaces the per-base virtual Sequence::char_at
// (src/model/Sequence.cpp:230-237, CompactLowNSequence::char_at_impl :533-540)
// the reference pays twice per window; the layout is add_hunk's (:561-596).
__device__ __forceinline__ void pack16(uint4 v, uint32_t& codes, uint32_t& ns,
                                       uint32_t& bad) {
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
    codes = 0;
    ns = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
#pragma unroll
        for (int b = 0; b < 4; b++) {
            uint32_t c = (w[i] >> (8 * b)) & 0xff;
            // A 0x41, C 0x43, G 0x47, T 0x54, N 0x4E: (c>>1)&3 = A0 C1 G3 T2,
            // remapped to the reference's A0 T1 G2 C3 by a 4-entry table
            uint32_t x = (c >> 1) & 3;
            uint32_t code = (0x9Cu >> (2 * x)) & 3;
            uint32_t is_n = (c == 'N');
            bool ok = (c == 'A') | (c == 'T') | (c == 'G') | (c == 'C') | is_n;
            bad += ok ? 0 : 1;
            code = (is_n | !ok) ? 0 : code;
            int j = i * 4 + b;
            codes |= code << (2 * j);
            ns |= (is_n | (ok ? 0u : 1u)) << j;
        }
    }
}

__global__ void __launch_bounds__(256)
pack_kernel(const uint4* __restrict__ ascii, uint32_t* __restrict__ packed,
            uint32_t* __restrict__ nmask, uint64_t n_groups32,
            uint32_t* __restrict__ counters) {
    uint64_t t = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
    uint32_t bad = 0;
    if (t < n_groups32) {
        uint4 a = __ldg(ascii + 2 * t);
        uint4 b = __ldg(ascii + 2 * t + 1);
        uint32_t c0, c1, n0, n1;
        pack16(a, c0, n0, bad);
        pack16(b, c1, n1, bad);
        reinterpret_cast<uint2*>(packed)[t] = make_uint2(c0, c1);
        nmask[t] = n0 | (n1 << 16);
    }
    bad = __reduce_add_sync(0xffffffffu, bad);
    if (bad && (threadIdx.x & 31) == 0) {
        atomicAdd(counters + kCntBadInput, bad);
    }
}

void launch_pack(const uint8_t* ascii, uint32_t* packed, uint32_t* nmask,
                 uint64_t n_padded, uint32_t* counters, cudaStream_t s) {
    uint64_t groups = n_padded / 32;
    if (groups == 0) return;
    uint32_t blocks = uint32_t((groups + 255) / 256);
    pack_kernel<<<blocks, 256, 0, s>>>(reinterpret_cast<const uint4*>(ascii),
                                       packed, nmask, groups, counters);
}

// ---- digit histograms -------------------------------------------------------
// One read of the packed genome produces the histogram of every radix digit of
// every valid canonical k-mer (the records are not materialised).  One thread
// walks 32 consecutive windows, sliding a 128-bit register window with funnel
// shifts.  Persistent grid: CTAs loop over chunks and flush their
// shared-memory histograms once.
template <int PASSES>
__global__ void __launch_bounds__(256)
hist_kernel(const uint32_t* __restrict__ packed,
            const uint32_t* __restrict__ nmask, uint64_t n_groups32, int k,
            int width, int capbits, uint32_t* __restrict__ hist) {
    extern __shared__ uint32_t sh[];
    const int cap = 1 << capbits;
    for (int i = threadIdx.x; i < PASSES * cap; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
    const uint64_t kmask = k == 32 ? ~0ull : ((1ull << (2 * k)) - 1);
    const uint32_t nm = k == 32 ? ~0u : ((1u << k) - 1);
    const uint32_t dmask = (1u << width) - 1;
    for (uint64_t t = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
            t < n_groups32; t += stride) {
        // 64 bases of codes and 64 bits of N mask starting at position 32 t
        // (the buffers carry one spare group past n_groups32, see upload)
        uint2 wl = __ldg(reinterpret_cast<const uint2*>(packed) + t);
        uint2 wh = __ldg(reinterpret_cast<const uint2*>(packed) + t + 1);
        uint64_t lo = (uint64_t(wl.y) << 32) | wl.x;
        uint64_t hi = (uint64_t(wh.y) << 32) | wh.x;
        uint64_t mm = (uint64_t(__ldg(nmask + t + 1)) << 32) | __ldg(nmask + t);
#pragma unroll 4
        for (int j = 0; j < 32; j++) {
            uint64_t dir = lo & kmask;
            bool valid = (uint32_t(mm) & nm) == 0;
            if (valid) {
                uint64_t rev = revcomp_key(dir, k);
                uint64_t h = dir < rev ? dir : rev;
#pragma unroll
                for (int p = 0; p < PASSES; p++) {
                    // digits past the key's top bit are 0 (h < 2^(2k))
                    atomicAdd(&sh[p * cap + (uint32_t(h >> (p * width)) & dmask)], 1u);
                }
            }
            lo = (lo >> 2) | (hi << 62);
            hi >>= 2;
            mm >>= 1;
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < PASSES * cap; i += blockDim.x) {
        uint32_t v = sh[i];
        if (v) atomicAdd(hist + i, v);
    }
}

void launch_hist(const SortPlan& plan, const uint32_t* packed,
                 const uint32_t* nmask, uint64_t n_padded, int k, uint32_t* hist,
                 int n_sms, cudaStream_t s) {
    uint64_t groups = n_padded / 32;
    if (groups == 0) return;
    uint64_t want = (groups + 255) / 256;
    uint32_t blocks = uint32_t(want < uint64_t(n_sms) * 4 ? want : uint64_t(n_sms) * 4);
    const size_t smem = size_t(plan.passes) * plan.bins() * 4;
    switch (plan.passes) {
#define NPGE_HIST_CASE(P)                                                      \
    case P:                                                                    \
        cudaFuncSetAttribute(hist_kernel<P>,                                   \
                             cudaFuncAttributeMaxDynamicSharedMemorySize,      \
                             int(smem));                                       \
        hist_kernel<P><<<blocks, 256, smem, s>>>(packed, nmask, groups, k,     \
                                                  plan.width, plan.capbits, hist); \
        break;
        NPGE_HIST_CASE(1) NPGE_HIST_CASE(2) NPGE_HIST_CASE(3) NPGE_HIST_CASE(4)
        NPGE_HIST_CASE(5) NPGE_HIST_CASE(6) NPGE_HIST_CASE(7) NPGE_HIST_CASE(8)
        NPGE_HIST_CASE(9) NPGE_HIST_CASE(10) NPGE_HIST_CASE(11) NPGE_HIST_CASE(12)
        NPGE_HIST_CASE(13) NPGE_HIST_CASE(14) NPGE_HIST_CASE(15) NPGE_HIST_CASE(16)
#undef NPGE_HIST_CASE
    }
}

// ---- histogram scan -----------------------------------------------------------
// counts -> exclusive digit bases, one block of `bins` threads per pass.
__global__ void __launch_bounds__(1024)
scan_hist_kernel(uint32_t* __restrict__ hist, uint32_t* __restrict__ counters) {
    __shared__ uint32_t warp_sums[32];
    const int bins = blockDim.x;
    uint32_t* h = hist + blockIdx.x * bins;
    uint32_t v = h[threadIdx.x];
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
        if ((threadIdx.x & 31) >= o) incl += u;
    }
    if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = incl;
    __syncthreads();
    uint32_t base = 0;
    for (int w = 0; w < int(threadIdx.x >> 5); w++) base += warp_sums[w];
    h[threadIdx.x] = base + incl - v;
    if (blockIdx.x == 0 && int(threadIdx.x) == bins - 1) {
        counters[kCntValid] = base + incl;
    }
}

void launch_scan_hist(const SortPlan& plan, uint32_t* hist, uint32_t* counters,
                      cudaStream_t s) {
    scan_hist_kernel<<<plan.passes, plan.bins(), 0, s>>>(hist, counters);
}

}  // namespace npge_b200
