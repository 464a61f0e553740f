// ingest_kernels.cu -- sequence ingest on the device:
// raw sequence text (FASTA body lines with their newlines, any case, gaps,
// IUPAC ambiguity codes) -> Sequence::to_atgcn -> the padded ASCII layout that
// pack_kernel turns into 2 bits per base + N mask.
//
// Reference: Sequence::to_atgcn, src/model/Sequence.cpp:151-179 (upper-case;
// A T G C N kept; R Y M K W S B V H D become N; every other byte is dropped),
// applied by the reference on the host while reading (FastaReader / Read.cpp).
// Dropping bytes is a stream compaction: count per tile -> scan -> scatter.
#include "kernels.h"

namespace npge_b200 {

namespace {

constexpr int kIT = 512;          // threads per CTA
constexpr int kIB = kIT * 8;      // raw bytes per tile

// to_atgcn of one byte: the byte to keep, or 0 when it is dropped
__device__ __forceinline__ uint32_t clean_byte(uint32_t c) {
    if (c >= 'a' && c <= 'z') c -= 32;
    const bool base = (c == 'A') | (c == 'T') | (c == 'G') | (c == 'C') | (c == 'N');
    const bool ambiguous = (c == 'R') | (c == 'Y') | (c == 'M') | (c == 'K') | (c == 'W') |
                           (c == 'S') | (c == 'B') | (c == 'V') | (c == 'H') | (c == 'D');
    return base ? c : (ambiguous ? uint32_t('N') : 0u);
}

// the 8 raw bytes of this thread (zero past the end), cleaned; returns the count kept
__device__ __forceinline__ uint32_t load_clean8(const uint8_t* __restrict__ raw, uint64_t n,
                                                uint64_t at, uint32_t (&out)[8]) {
    uint32_t kept = 0;
    if (at + 8 <= n && (reinterpret_cast<uintptr_t>(raw + at) & 7) == 0) {
        const uint2 v = *reinterpret_cast<const uint2*>(raw + at);
#pragma unroll
        for (int i = 0; i < 8; i++) {
            out[i] = clean_byte(((i < 4 ? v.x : v.y) >> (8 * (i & 3))) & 0xff);
            kept += out[i] != 0;
        }
    } else {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            out[i] = at + i < n ? clean_byte(raw[at + i]) : 0u;
            kept += out[i] != 0;
        }
    }
    return kept;
}

__device__ __forceinline__ uint32_t block_excl(uint32_t v, uint32_t* tmp, uint32_t* total) {
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
    for (int w = 0; w < kIT / 32; w++) {
        const uint32_t t = tmp[w];
        if (w < warp) base += t;
        all += t;
    }
    if (total) *total = all;
    return base + incl - v;
}

__global__ void __launch_bounds__(kIT)
ingest_count_kernel(IngestArgs a) {
    __shared__ uint32_t tmp[kIT / 32];
    uint32_t c[8], total;
    const uint64_t at = uint64_t(blockIdx.x) * kIB + threadIdx.x * 8;
    const uint32_t kept = load_clean8(a.raw, a.n_raw, at, c);
    block_excl(kept, tmp, &total);
    if (threadIdx.x == 0) a.tile_count[blockIdx.x] = total;
}

__global__ void __launch_bounds__(1024)
ingest_scan_kernel(IngestArgs a, uint32_t n_tiles) {
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
        a.tile_offset[n_tiles] = run;
    }
    __syncthreads();
    uint64_t run = part[threadIdx.x];
    for (uint32_t i = b; i < e; i++) {
        a.tile_offset[i] = run;
        run += a.tile_count[i];
    }
}

// rank[s] = kept bytes before raw offset raw_offsets[s], s = 0 .. n_seqs: one warp each
__global__ void __launch_bounds__(256)
ingest_rank_kernel(IngestArgs a, uint32_t n_tiles) {
    const uint32_t s = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (s > a.n_seqs) return;
    const uint64_t off = a.raw_offsets[s];
    const uint64_t tile = off / kIB;
    uint32_t kept = 0;
    for (uint64_t i = tile * kIB + lane; i < off; i += 32) kept += clean_byte(a.raw[i]) != 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) kept += __shfl_xor_sync(0xffffffffu, kept, o);
    if (lane == 0) a.seq_rank[s] = a.tile_offset[tile < n_tiles ? tile : n_tiles] + kept;
}

// every kept byte goes to dst_base[s] + (its rank - seq_rank[s]) of its sequence s
__global__ void __launch_bounds__(kIT)
ingest_scatter_kernel(IngestArgs a) {
    __shared__ uint32_t tmp[kIT / 32];
    uint32_t c[8];
    const uint64_t at = uint64_t(blockIdx.x) * kIB + threadIdx.x * 8;
    const uint32_t kept = load_clean8(a.raw, a.n_raw, at, c);
    uint64_t rank = a.tile_offset[blockIdx.x] + block_excl(kept, tmp, nullptr);
    if (kept == 0 || at >= a.n_raw) return;
    // sequence of the first byte: last s with raw_offsets[s] <= at
    uint32_t lo = 0, hi = a.n_seqs;  // answer in [lo, hi)
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (a.raw_offsets[mid] <= at) lo = mid; else hi = mid;
    }
    uint32_t s = lo;
    uint64_t next = a.raw_offsets[s + 1];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        while (at + i >= next && s + 1 < a.n_seqs) {  // (empty sequences are skipped)
            s++;
            next = a.raw_offsets[s + 1];
        }
        if (c[i]) {
            const uint64_t base = a.dst_base[s];
            if (base != ~0ull) a.ascii[base + (rank - a.seq_rank[s])] = uint8_t(c[i]);
            rank++;
        }
    }
}

}  // namespace

uint32_t ingest_tiles(uint64_t n_raw) { return uint32_t((n_raw + kIB - 1) / kIB); }

// stage 0: count, scan, per-sequence ranks (seq_rank[0 .. n_seqs]);
// stage 1: scatter the cleaned bytes to their places in the padded layout
void launch_ingest(const IngestArgs& a, int stage, cudaStream_t s, uint32_t* launches) {
    const uint32_t tiles = ingest_tiles(a.n_raw);
    if (stage == 0) {
        if (tiles) ingest_count_kernel<<<tiles, kIT, 0, s>>>(a);
        ingest_scan_kernel<<<1, 1024, 0, s>>>(a, tiles);
        ingest_rank_kernel<<<(a.n_seqs + 1 + 7) / 8, 256, 0, s>>>(a, tiles);
        *launches += 3;
    } else if (tiles) {
        ingest_scatter_kernel<<<tiles, kIT, 0, s>>>(a);
        *launches += 1;
    }
}

}  // namespace npge_b200
