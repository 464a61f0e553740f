// fragment_kernels.cu -- FragmentFinder: every exact
// occurrence of a pattern, on either strand, in the uploaded sequences.
//
// Reference: src/algo/FragmentFinder.cpp:42-171.  It rolls the same two hashes
// as AnchorFinder over every sequence and compares them with the pattern's
// hash, verifying by string compare when the hash cannot be trusted (pattern
// longer than 32, N in the pattern or in the window).  Net effect: window == P
// gives a fragment with ori = +1, revcomp(window) == P one with ori = -1, N
// matching only N.  Here the compare is done directly on the packed genome:
// 32 bases (one u64) + 32 N-mask bits per step, against the packed pattern and
// the packed reverse complement of the pattern.
//
// Two launches over the genome (count per tile, then ordered write) and one
// single-CTA scan between them; the output is in scan order (sequence by size
// descending, position ascending, forward before reverse), cut to max-matches,
// which is what a one-worker run of the reference produces.
#include <mutex>

#include "kernels.h"

namespace npge_b200 {

namespace {

constexpr int kFT = 512;  // threads per CTA; 8 consecutive positions per thread

__device__ __forceinline__ uint32_t mask_bits(const uint32_t* __restrict__ M, uint32_t rel,
                                              int len) {
    const uint32_t w = rel >> 5, s = rel & 31;
    uint32_t m = __funnelshift_r(M[w], M[w + 1], s);
    if (len < 32) m &= (1u << len) - 1;
    return m;
}

// does the text at position g equal the pattern of strand `strand`?
__device__ __forceinline__ bool equals_at(const FragmentArgs& a, uint32_t g, int strand) {
    const uint64_t* pk = a.pat_key + size_t(strand) * a.n_chunks;
    const uint32_t* pm = a.pat_mask + size_t(strand) * a.n_chunks;
    for (uint32_t c = 0; c < a.n_chunks; c++) {
        const int len = int(min(32u, a.pattern_len - 32u * c));
        const uint32_t rel = g + 32u * c;
        if (forward_key(a.packed, rel, len) != pk[c]) return false;
        if (mask_bits(a.nmask, rel, len) != pm[c]) return false;
    }
    return true;
}

// sequence of position g (scan-order index), via the chunk look-up table
__device__ __forceinline__ uint32_t seq_of(const FragmentArgs& a, uint32_t g) {
    const uint32_t q = g >> a.lut_shift;
    uint32_t lo = a.seq_lut[q], hi = a.seq_lut[q + 1] + 1;  // answer in [lo, hi)
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (a.seq_start[mid] <= g) lo = mid; else hi = mid;
    }
    return lo;
}

// bit 0: forward match at g, bit 1: reverse match
__device__ __forceinline__ uint32_t matches_at(const FragmentArgs& a, uint64_t g) {
    if (g + a.pattern_len > a.n_positions) return 0;
    uint32_t m = 0;
    if (equals_at(a, uint32_t(g), 0)) m |= 1;
    if (equals_at(a, uint32_t(g), 1)) m |= 2;
    if (m) {  // the window must lie inside one sequence (a pattern with N could
              // otherwise match across the pad between two sequences)
        const uint32_t s = seq_of(a, uint32_t(g));
        if (g + a.pattern_len > a.seq_start[s + 1] - 1) m = 0;
    }
    return m;
}

__device__ __forceinline__ uint32_t block_sum_excl(uint32_t v, uint32_t* tmp, uint32_t* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += u;
    }
    if (lane == 31) tmp[warp] = incl;
    __syncthreads();
    uint32_t base = 0, all = 0;
    for (int w = 0; w < kFT / 32; w++) {
        const uint32_t t = tmp[w];
        if (w < warp) base += t;
        all += t;
    }
    if (total) *total = all;
    return base + incl - v;
}

// MODE 0: tile_count[tile] = matches in the tile.  MODE 1: write them in order.
template <int MODE>
__global__ void __launch_bounds__(kFT)
fragment_kernel(FragmentArgs a) {
    __shared__ uint32_t tmp[kFT / 32];
    const uint64_t g0 = (uint64_t(blockIdx.x) * kFT + threadIdx.x) * 8;
    uint64_t first = 0;
    if (MODE == 1) {
        first = a.tile_offset[blockIdx.x];
        if (first >= a.max_matches) return;  // whole tile past the cut
    }
    uint32_t bits = 0, count = 0;  // 2 bits per position
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t m = matches_at(a, g0 + i);
        bits |= m << (2 * i);
        count += __popc(m);
    }
    if (MODE == 0) {
        uint32_t total;
        block_sum_excl(count, tmp, &total);
        if (threadIdx.x == 0) a.tile_count[blockIdx.x] = total;
        return;
    }
    uint64_t at = first + block_sum_excl(count, tmp, nullptr);
    for (int i = 0; i < 8 && bits; i++) {
        for (int strand = 0; strand < 2; strand++) {
            if (!((bits >> (2 * i + strand)) & 1)) continue;
            if (at < a.max_matches) {
                const uint32_t g = uint32_t(g0 + i);
                const uint32_t s = seq_of(a, g);
                a.r_seq[at] = a.seq_caller[s];
                a.r_pos[at] = g - uint32_t(a.seq_start[s]);
                a.r_ori[at] = strand == 0 ? 1 : -1;
            }
            at++;
        }
    }
}

// exclusive scan of the tile counts by one CTA; summary[0] = total
__global__ void __launch_bounds__(1024)
fragment_scan_kernel(FragmentArgs a, uint32_t n_tiles) {
    __shared__ uint64_t part[1024];
    const uint32_t per = (n_tiles + 1023) / 1024;
    const uint32_t b = min(threadIdx.x * per, n_tiles), e = min(b + per, n_tiles);
    uint64_t sum = 0;
    for (uint32_t i = b; i < e; i++) sum += a.tile_count[i];
    part[threadIdx.x] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint64_t run = 0;
        for (int i = 0; i < 1024; i++) {
            const uint64_t v = part[i];
            part[i] = run;
            run += v;
        }
        a.summary[0] = run;
    }
    __syncthreads();
    uint64_t run = part[threadIdx.x];
    for (uint32_t i = b; i < e; i++) {
        a.tile_offset[i] = run;
        run += a.tile_count[i];
    }
}

}  // namespace

uint32_t fragment_tiles(uint64_t n_padded) {
    return uint32_t((n_padded + kFT * 8 - 1) / (kFT * 8));
}

// stage 0: count + scan (summary[0] = matches before the cut); stage 1: write
void launch_fragment_find(const FragmentArgs& a, int stage, cudaStream_t s,
                          uint32_t* launches) {
    const uint32_t tiles = fragment_tiles(a.n_padded);
    if (tiles == 0) return;
    if (stage == 0) {
        fragment_kernel<0><<<tiles, kFT, 0, s>>>(a);
        fragment_scan_kernel<<<1, 1024, 0, s>>>(a, tiles);
        *launches += 2;
    } else {
        fragment_kernel<1><<<tiles, kFT, 0, s>>>(a);
        *launches += 1;
    }
}

}  // namespace npge_b200

// ==== npge_fragment_find_device_noalloc ===================================================
// The same search for callers that must not allocate (code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.  The scratch of the kernels
// (tile counts and offsets, the total, the packed pattern) lives in ONE file-scope
// device array of a fixed size, which every GPU that loads the library holds until
// exit; the caller passes the outputs as device arrays.  The entry point calls no
// cudaMalloc / cudaFree / cudaHostAlloc and no allocating library routine; on the
// success path it does not touch the host heap either (fixed-size stack buffers).
//
// Every call on a device shares the array, which keeps the previous call's contents:
// everything a call reads from it is written by the same call first (the pattern by a
// copy, the tile counts by the count kernel over its whole grid, the total and the
// offsets by the scan kernel); the total is cleared as well.  Calls that overlap on one
// GPU (two streams, two host threads, two contexts) RACE on it: serialise them.
namespace npge_b200 {
namespace {

constexpr uint32_t kNoAllocMaxTiles = NPGE_FRAGMENT_NOALLOC_MAX_POSITIONS / (kFT * 8);
constexpr uint32_t kNoAllocMaxChunks = NPGE_FRAGMENT_NOALLOC_MAX_PATTERN / 32;
// [summary u64 x 2][tile_offset u64 x tiles][pat_key u64 x 2 chunks]
// [tile_count u32 x tiles][pat_mask u32 x 2 chunks]
constexpr size_t kNoAllocBytes = 16 + size_t(kNoAllocMaxTiles) * 8 + size_t(kNoAllocMaxChunks) * 16 +
                                 size_t(kNoAllocMaxTiles) * 4 + size_t(kNoAllocMaxChunks) * 8;

// In host code this name is a host-side stand-in, NOT a device address: the address
// the kernels get comes from cudaGetSymbolAddress, once per device (below).
static __device__ __align__(256) unsigned char g_fragment_scratch[kNoAllocBytes];

constexpr int kMaxDevices = 64;
std::mutex g_scratch_mu;
unsigned char* g_scratch_ptr[kMaxDevices] = {};

// the array's address on `device` (which must be current)
int scratch_address(int device, unsigned char** out) {
    if (device < 0 || device >= kMaxDevices) {
        return set_last_error(NPGE_ANCHOR_ERR_PARAMS, "device index out of range");
    }
    std::lock_guard<std::mutex> lock(g_scratch_mu);
    if (!g_scratch_ptr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_fragment_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return set_last_error(NPGE_ANCHOR_ERR_CUDA,
                                  std::string("cudaGetSymbolAddress(fragment scratch): ") +
                                      cudaGetErrorString(e));
        }
        g_scratch_ptr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_scratch_ptr[device];
    return NPGE_ANCHOR_OK;
}

// 2-bit codes of the reference (A0 T1 G2 C3; char_to_size.hpp:24-36); N -> code 0 +
// mask bit, as pack_kernel does.  False for a byte outside ATGCN.
bool pack_pattern_fixed(const char* p, uint32_t len, uint64_t* key, uint32_t* mask) {
    for (uint32_t c = 0; c < (len + 31) / 32; c++) {
        key[c] = 0;
        mask[c] = 0;
    }
    for (uint32_t i = 0; i < len; i++) {
        uint64_t code = 0;
        switch (p[i]) {
        case 'A': code = 0; break;
        case 'T': code = 1; break;
        case 'G': code = 2; break;
        case 'C': code = 3; break;
        case 'N': mask[i / 32] |= 1u << (i % 32); break;
        default: return false;
        }
        key[i / 32] |= code << (2 * (i % 32));
    }
    return true;
}

}  // namespace
}  // namespace npge_b200

using namespace npge_b200;

#define NPGE_NA_CUDA(expr)                                                              \
    do {                                                                                \
        cudaError_t e__ = (expr);                                                       \
        if (e__ != cudaSuccess) {                                                       \
            return set_last_error(NPGE_ANCHOR_ERR_CUDA, std::string("CUDA error: ") +   \
                                  cudaGetErrorString(e__) + " at " #expr);              \
        }                                                                               \
    } while (0)

extern "C" int npge_fragment_find_device_noalloc(npge_anchor_ctx* ctx, const uint8_t* pattern,
                                                 uint32_t pattern_len, int64_t max_matches,
                                                 uint32_t* d_frag_seq, uint32_t* d_frag_min_pos,
                                                 int8_t* d_frag_ori, uint64_t capacity,
                                                 uint64_t* n_matches, uint64_t* total_matches) {
    if (n_matches) *n_matches = 0;
    if (total_matches) *total_matches = 0;
    if (!ctx || !pattern) return set_last_error(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    // the option checks of the processor: FragmentFinder.cpp:18-28, :37
    if (pattern_len == 0) {
        return set_last_error(NPGE_ANCHOR_ERR_PARAMS, "'pattern' should not be empty");
    }
    if (max_matches < 1) return set_last_error(NPGE_ANCHOR_ERR_PARAMS, "max-matches >= 1 is violated");
    if (capacity && (!d_frag_seq || !d_frag_min_pos || !d_frag_ori)) {
        return set_last_error(NPGE_ANCHOR_ERR_PARAMS, "NULL output array");
    }
    if (!ctx->uploaded || ctx->dist) {
        return set_last_error(NPGE_ANCHOR_ERR_STATE, "npge_anchor_upload has not been called");
    }
    if (uint32_t(ctx->upload_k) > pattern_len) {
        return set_last_error(NPGE_ANCHOR_ERR_STATE,
                              "sequences were uploaded with an anchor-size above the pattern length");
    }
    if (pattern_len > NPGE_FRAGMENT_NOALLOC_MAX_PATTERN) {
        return set_last_error(NPGE_ANCHOR_ERR_CAPACITY,
                              "npge_fragment_find_device_noalloc: pattern longer than "
                              "NPGE_FRAGMENT_NOALLOC_MAX_PATTERN (1024); use npge_fragment_find_device");
    }
    if (ctx->n_padded > NPGE_FRAGMENT_NOALLOC_MAX_POSITIONS) {
        return set_last_error(NPGE_ANCHOR_ERR_CAPACITY,
                              "npge_fragment_find_device_noalloc: more than "
                              "NPGE_FRAGMENT_NOALLOC_MAX_POSITIONS (2^26) positions uploaded: the "
                              "fixed scratch array is too small; use npge_fragment_find_device");
    }
    // pattern and its reverse complement (complement.hpp:19-32: N stays N), packed on the stack
    char fwd[NPGE_FRAGMENT_NOALLOC_MAX_PATTERN], rc[NPGE_FRAGMENT_NOALLOC_MAX_PATTERN];
    for (uint32_t i = 0; i < pattern_len; i++) {
        fwd[i] = char(pattern[i]);
        const char c = char(pattern[pattern_len - 1 - i]);
        rc[i] = c == 'A' ? 'T' : c == 'T' ? 'A' : c == 'G' ? 'C' : c == 'C' ? 'G' : c;
    }
    const uint32_t chunks = (pattern_len + 31) / 32;
    uint64_t key[2 * kNoAllocMaxChunks];
    uint32_t mask[2 * kNoAllocMaxChunks];
    if (!pack_pattern_fixed(fwd, pattern_len, key, mask) ||
        !pack_pattern_fixed(rc, pattern_len, key + chunks, mask + chunks)) {
        return set_last_error(NPGE_ANCHOR_ERR_INPUT, "pattern holds a byte outside ATGCN (to_atgcn first)");
    }
    NPGE_NA_CUDA(cudaSetDevice(ctx->device));
    unsigned char* scratch = nullptr;
    {
        const int rc_ = scratch_address(ctx->device, &scratch);
        if (rc_ != NPGE_ANCHOR_OK) return rc_;
    }
    const uint32_t tiles = fragment_tiles(ctx->n_padded);  // <= kNoAllocMaxTiles (checked above)
    uint64_t cap = uint64_t(max_matches) < capacity ? uint64_t(max_matches) : capacity;
    FragmentArgs a = {};
    a.packed = static_cast<const uint32_t*>(ctx->packed.ptr);
    a.nmask = static_cast<const uint32_t*>(ctx->nmask.ptr);
    a.n_positions = ctx->n_positions;
    a.n_padded = ctx->n_padded;
    a.pattern_len = pattern_len;
    a.n_chunks = chunks;
    a.seq_start = static_cast<const uint64_t*>(ctx->seq_start.ptr);
    a.seq_caller = static_cast<const uint32_t*>(ctx->seq_caller.ptr);
    a.seq_lut = static_cast<const uint32_t*>(ctx->seq_lut.ptr);
    a.lut_shift = ctx->lut_shift;
    a.max_matches = cap;
    a.summary = reinterpret_cast<uint64_t*>(scratch);
    a.tile_offset = reinterpret_cast<uint64_t*>(scratch + 16);
    uint64_t* d_key = a.tile_offset + kNoAllocMaxTiles;
    a.tile_count = reinterpret_cast<uint32_t*>(d_key + 2 * kNoAllocMaxChunks);
    uint32_t* d_mask = a.tile_count + kNoAllocMaxTiles;
    a.pat_key = d_key;
    a.pat_mask = d_mask;
    a.r_seq = d_frag_seq;
    a.r_pos = d_frag_min_pos;
    a.r_ori = d_frag_ori;
    cudaStream_t st = ctx->stream;
    uint64_t total = 0;
    NPGE_NA_CUDA(cudaMemsetAsync(a.summary, 0, 16, st));
    NPGE_NA_CUDA(cudaMemcpyAsync(d_key, key, size_t(chunks) * 16, cudaMemcpyHostToDevice, st));
    NPGE_NA_CUDA(cudaMemcpyAsync(d_mask, mask, size_t(chunks) * 8, cudaMemcpyHostToDevice, st));
    if (tiles) {
        uint32_t launches = 0;
        launch_fragment_find(a, 0, st, &launches);
        launch_fragment_find(a, 1, st, &launches);
        NPGE_NA_CUDA(cudaGetLastError());
        NPGE_NA_CUDA(cudaMemcpyAsync(&total, a.summary, 8, cudaMemcpyDeviceToHost, st));
        ctx->stats.gpu_launches = launches;
    }
    NPGE_NA_CUDA(cudaStreamSynchronize(st));  // key / mask / total are locals
    if (n_matches) *n_matches = total < cap ? total : cap;
    if (total_matches) *total_matches = total;
    return NPGE_ANCHOR_OK;
}
