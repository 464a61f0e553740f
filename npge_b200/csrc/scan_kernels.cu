// scan_kernels.cu -- pack (ASCII -> 2 bit + N mask).  sm_100a.
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

}  // namespace npge_b200
