// block_hash_kernels.cu -- the reference's block_hash (src/model/block_hash.cpp:
// 29-54) for every block of a flat block list, on the device: the value
// MoveUnchanged (src/algo/MoveUnchanged.cpp:37-49) and blockset_hash
// (block_hash.cpp:125-130) are built from.  SURVEY 8(f) rank 1 / 8(a) a19.
//
// One thread per block (anchor blocks hold a few fragments each, and there are
// millions of them): keys of the ids (block_hash_core.h) -> which of the two id
// lists is the smaller one (by their minima) -> sort that list -> fold the bytes
// of the joint string.  Blocks of up to kFastFrags fragments keep their keys in
// shared memory, interleaved by thread (a warp working on key i of its 32
// blocks touches 512 contiguous bytes).  Larger blocks, and blocks whose two
// lists start with the same id, build both lists in a slice of a global scratch
// buffer (heap sort above 24 keys).
// The first version kept the keys in per-thread local memory (768 B a thread):
// with every SM full that is more than the L2 holds, and the kernel moved 2 GB
// through DRAM for 150 MB of input (profiles/r01j_block_hash.md).
#include "block_hash_core.h"
#include "kernels.h"

namespace npge_b200 {

namespace {

constexpr int kFastFrags = 20;   // keys per thread in shared memory (40 KB per CTA)
constexpr int kThreads = 128;

__global__ void __launch_bounds__(kThreads) block_hash_kernel(BlockHashArgs a) {
    const uint64_t b = a.b0 + uint64_t(blockIdx.x) * kThreads + threadIdx.x;
    uint64_t h = 0;
    bool counted = false;
    if (b < a.b1) {
        const uint64_t f0 = a.block_off[b], f1 = a.block_off[b + 1];
        if (f1 < f0 || f0 < a.f_base || f1 > a.f_end) {
            atomicOr(a.error, 1u);  // offsets not ascending
        } else {
            bool ok = true;
            for (uint64_t f = f0; f < f1; f++) ok = ok && a.frag_seq[f] < a.n_seqs;
            if (!ok) {
                atomicOr(a.error, 2u);  // sequence index out of range
            } else {
                extern __shared__ __align__(16) unsigned char raw[];
                const KeySlots slots = {reinterpret_cast<IdKey*>(raw) + threadIdx.x, kThreads};
                const FlatBlocks fb = {a.frag_seq, a.min_pos, a.max_pos, a.frag_len, a.frag_ori,
                                       a.seq_rank, a.name_words, a.name_pos, a.name_len, a.bcd4,
                                       a.name_slow};
                h = block_hash_flat(fb, f0, f1, slots, kFastFrags,
                                    reinterpret_cast<IdKey*>(a.scratch) + 2 * (f0 - a.f_base));
                // blockset_hash skips blocks of <= 1 fragment (HashGroup::
                // create_task_impl, block_hash.cpp:101-113)
                counted = f1 - f0 > 1;
            }
        }
        a.hashes[b] = h;
    }
    uint64_t x = counted ? h : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x ^= __shfl_xor_sync(0xffffffffu, x, o);
    if ((threadIdx.x & 31) == 0 && x) {
        atomicXor(reinterpret_cast<unsigned long long*>(a.xor_out),
                  static_cast<unsigned long long>(x));
    }
}

// cut_block[i] = first block whose first fragment is >= i * cap (i <= n_cuts;
// the last one = n_blocks); cut_frag[i] = its first fragment.
__global__ void block_hash_cuts_kernel(const uint64_t* block_off, uint64_t n_blocks,
                                       uint64_t cap, uint32_t n_cuts, uint64_t* cut_block,
                                       uint64_t* cut_frag) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n_cuts) return;
    uint64_t lo = 0, hi = n_blocks;  // answer in [lo, hi]
    if (i < n_cuts) {
        const uint64_t want = uint64_t(i) * cap;
        while (lo < hi) {
            const uint64_t mid = (lo + hi) >> 1;
            if (block_off[mid] >= want) hi = mid; else lo = mid + 1;
        }
    } else {
        lo = n_blocks;
    }
    cut_block[i] = lo;
    cut_frag[i] = block_off[lo];
}

}  // namespace

void launch_block_hash_cuts(const uint64_t* block_off, uint64_t n_blocks, uint64_t cap,
                            uint32_t n_cuts, uint64_t* cut_block, uint64_t* cut_frag,
                            cudaStream_t s) {
    block_hash_cuts_kernel<<<(n_cuts + 1 + 127) / 128, 128, 0, s>>>(
        block_off, n_blocks, cap, n_cuts, cut_block, cut_frag);
}

void launch_block_hash(const BlockHashArgs& a, cudaStream_t s) {
    if (a.b1 <= a.b0) return;
    const uint64_t n = a.b1 - a.b0;
    block_hash_kernel<<<uint32_t((n + kThreads - 1) / kThreads), kThreads,
                        kFastFrags * kThreads * sizeof(IdKey), s>>>(a);
}

}  // namespace npge_b200
