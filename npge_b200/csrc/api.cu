// api.cu -- the C ABI (include/npge_anchor.h): host orchestration of the CUDA
// path.  No CPU fallback: every compute entry point needs a CUDA device.
#include <algorithm>
#include <functional>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <new>
#include <utility>

#include "block_hash_core.h"
#include "block_hash_names.h"
#include "fetch_expand.h"
#include "kernels.h"

using namespace npge_b200;

namespace {

thread_local std::string g_last_error;

int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

}  // namespace

namespace npge_b200 {
int set_last_error(int code, const std::string& msg) { return fail(code, msg); }
}  // namespace npge_b200

namespace {

#define NPGE_CUDA(expr)                                                        \
    do {                                                                       \
        cudaError_t e__ = (expr);                                              \
        if (e__ != cudaSuccess) {                                              \
            return fail(e__ == cudaErrorMemoryAllocation                       \
                            ? NPGE_ANCHOR_ERR_MEMORY : NPGE_ANCHOR_ERR_CUDA,    \
                        std::string("CUDA error: ") + cudaGetErrorString(e__) + \
                            " at " #expr);                                     \
        }                                                                      \
    } while (0)

// Every device buffer ends in kCanary bytes of 0xA5 past its usable size:
// npge_anchor_debug_check reads them back (compute-sanitizer is closed on the pool this
// was developed on; an out-of-bounds store past the end of a buffer lands here first).
constexpr size_t kCanary = 256;

int ensure(DeviceBuffer& b, size_t bytes) {
    if (bytes == 0) bytes = 256;
    if (b.bytes >= bytes) return NPGE_ANCHOR_OK;
    if (b.ptr) {
        cudaFree(b.ptr);
        b.ptr = nullptr;
        b.bytes = 0;
    }
    // grow with some headroom so repeated runs on slightly larger inputs reuse
    size_t want = (bytes + bytes / 16 + 256 + 255) / 256 * 256;
    cudaError_t e = cudaMalloc(&b.ptr, want + kCanary);
    if (e != cudaSuccess) {
        cudaGetLastError();
        want = (bytes + 255) / 256 * 256;
        e = cudaMalloc(&b.ptr, want + kCanary);
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(NPGE_ANCHOR_ERR_MEMORY,
                    "device allocation of " + std::to_string(bytes) + " bytes failed");
    }
    // (cudaMalloc has synchronised the device; nothing uses the new buffer yet)
    if (cudaMemset(static_cast<uint8_t*>(b.ptr) + want, 0xA5, kCanary) != cudaSuccess ||
        cudaDeviceSynchronize() != cudaSuccess) {
        cudaGetLastError();
        cudaFree(b.ptr);
        b.ptr = nullptr;
        return fail(NPGE_ANCHOR_ERR_CUDA, "could not initialise a device buffer");
    }
    b.bytes = want;
    return NPGE_ANCHOR_OK;
}

void release(DeviceBuffer& b) {
    if (b.ptr) cudaFree(b.ptr);
    b.ptr = nullptr;
    b.bytes = 0;
}

#define NPGE_TRY(expr)                        \
    do {                                      \
        int rc__ = (expr);                    \
        if (rc__ != NPGE_ANCHOR_OK) return rc__; \
    } while (0)

template <typename T>
T* ptr(const DeviceBuffer& b) {
    return static_cast<T*>(b.ptr);
}

// Pinned host memory for the result arrays: device -> host copies run at PCIe
// speed only into page-locked memory, and cudaHostAlloc is far too slow to call
// per result, so freed blocks are kept (up to kPoolLimit bytes) and reused.
class PinnedPool {
public:
    void* take(size_t bytes) {
        {
            std::lock_guard<std::mutex> lock(mu_);
            for (size_t i = 0; i < free_.size(); i++) {
                if (free_[i].second == bytes) {
                    void* p = free_[i].first;
                    free_[i] = free_.back();
                    free_.pop_back();
                    pooled_ -= bytes;
                    return p;
                }
            }
        }
        void* p = nullptr;
        if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) {
            cudaGetLastError();
            throw std::bad_alloc();
        }
        return p;
    }
    void give(void* p, size_t bytes) {
        {
            std::lock_guard<std::mutex> lock(mu_);
            if (pooled_ + bytes <= kPoolLimit) {
                free_.push_back({p, bytes});
                pooled_ += bytes;
                return;
            }
        }
        cudaFreeHost(p);
    }
    static PinnedPool& instance() {
        static PinnedPool* pool = new PinnedPool;  // never destroyed: outlives CUDA
        return *pool;
    }
    static size_t round_up(size_t bytes) {
        size_t b = 4096;
        while (b < bytes) b <<= 1;
        return b;
    }
private:
    static constexpr size_t kPoolLimit = size_t(8) << 30;
    std::mutex mu_;
    std::vector<std::pair<void*, size_t>> free_;
    size_t pooled_ = 0;
};

// std::vector allocator on the pool; default construction leaves the element
// uninitialised (resize() must not memset half a gigabyte before the copy).
template <typename T>
struct PinnedAlloc {
    typedef T value_type;
    PinnedAlloc() {}
    template <typename U> PinnedAlloc(const PinnedAlloc<U>&) {}
    T* allocate(size_t n) {
        return static_cast<T*>(PinnedPool::instance().take(
            PinnedPool::round_up(n * sizeof(T))));
    }
    void deallocate(T* p, size_t n) {
        PinnedPool::instance().give(p, PinnedPool::round_up(n * sizeof(T)));
    }
    template <typename U> void construct(U*) {}
    template <typename U, typename... A> void construct(U* p, A&&... a) {
        ::new (static_cast<void*>(p)) U(std::forward<A>(a)...);
    }
    template <typename U> bool operator==(const PinnedAlloc<U>&) const { return true; }
    template <typename U> bool operator!=(const PinnedAlloc<U>&) const { return false; }
};

template <typename T> using PinnedVector = std::vector<T, PinnedAlloc<T>>;

struct HostResult {
    PinnedVector<uint64_t> block_offsets, block_hash;
    std::vector<uint64_t> new_used;
    PinnedVector<uint32_t> frag_seq, frag_min_pos;
    PinnedVector<int8_t> frag_ori;
};

// How large results come back (npge_anchor_fetch): plain copies of the three
// fragment arrays, or the compact transfer (fetch_compact below).  NPGE_FETCH =
// plain | compact | auto overrides the default.
enum { kFetchPlain, kFetchCompact, kFetchAuto };
constexpr int kFetchDefault = kFetchPlain;
int fetch_mode() {
    const char* e = getenv("NPGE_FETCH");
    if (e && strcmp(e, "plain") == 0) return kFetchPlain;
    if (e && strcmp(e, "compact") == 0) return kFetchCompact;
    if (e && strcmp(e, "auto") == 0) return kFetchAuto;
    return kFetchDefault;
}

int check_params(const npge_anchor_params* p) {
    if (!p) return fail(NPGE_ANCHOR_ERR_PARAMS, "params is NULL");
    // add_opt_rule("anchor-size > 0"), ("anchor-size <= 32"): AnchorFinder.cpp:49-51
    if (p->anchor_size <= 0) {
        return fail(NPGE_ANCHOR_ERR_PARAMS, "anchor-size > 0 is violated");
    }
    if (p->anchor_size > NPGE_ANCHOR_MAX_SIZE) {
        return fail(NPGE_ANCHOR_ERR_PARAMS, "anchor-size <= 32 is violated");
    }
    // max-anchor-fragments has no rule in the reference: a negative value is
    // compared as size_t (AnchorFinder.cpp:361) and never truncates.  Here
    // uint64_t(p->max_anchor_fragments) gives the same "no limit".
    return NPGE_ANCHOR_OK;
}

}  // namespace

extern "C" const char* npge_anchor_last_error(void) {
    return g_last_error.c_str();
}

extern "C" const char* npge_anchor_version(void) {
    return "npge_anchor 0.1 sm_100a";
}

extern "C" int npge_anchor_debug_plan(npge_anchor_ctx* ctx, int32_t fine_bits,
                                      int32_t first_split) {
    if (!ctx) return fail(NPGE_ANCHOR_ERR_PARAMS, "ctx is NULL");
    ctx->force_f = fine_bits;
    ctx->force_wa = first_split;
    return NPGE_ANCHOR_OK;
}

// Test hook without a device: the digit plan of a job of n_windows records.
extern "C" int npge_anchor_debug_plan_of(int32_t anchor_size, uint64_t n_windows,
                                         int32_t* bits_a, int32_t* bits_b,
                                         int32_t* sort_passes) {
    npge_anchor_params probe = {anchor_size, 0, 0};
    NPGE_TRY(check_params(&probe));
    const MsdPlan plan = make_msd_plan(anchor_size, n_windows, 0, 0);
    if (bits_a) *bits_a = plan.wa;
    if (bits_b) *bits_b = plan.wb;
    if (sort_passes) *sort_passes = plan.smem_passes;
    return NPGE_ANCHOR_OK;
}

extern "C" int npge_anchor_create(npge_anchor_ctx** out, int device) {
    if (!out) return fail(NPGE_ANCHOR_ERR_PARAMS, "ctx is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        cudaGetLastError();
        return fail(NPGE_ANCHOR_ERR_CUDA,
                    "no CUDA device: the anchor finder has no CPU fallback");
    }
    if (device < 0 || device >= count) {
        return fail(NPGE_ANCHOR_ERR_PARAMS, "device index out of range");
    }
    NPGE_CUDA(cudaSetDevice(device));
    npge_anchor_ctx* ctx = new (std::nothrow) npge_anchor_ctx;
    if (!ctx) return fail(NPGE_ANCHOR_ERR_MEMORY, "out of host memory");
    ctx->device = device;
    NPGE_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    for (auto& ev : ctx->ev) NPGE_CUDA(cudaEventCreate(&ev));
    for (auto& ev : ctx->ev_pass) NPGE_CUDA(cudaEventCreate(&ev));
    *out = ctx;
    return NPGE_ANCHOR_OK;
}

namespace {

std::vector<DeviceBuffer*> all_buffers(npge_anchor_ctx* ctx) {
    return {&ctx->ascii, &ctx->packed, &ctx->nmask, &ctx->seq_start,
                           &ctx->seq_caller, &ctx->seq_pair, &ctx->g_seq_pair, &ctx->counters, &ctx->used,
                           &ctx->found_bits, &ctx->found_tmp, &ctx->sel,
                           &ctx->r_seq, &ctx->r_pos, &ctx->r_ori,
                           &ctx->r_block_off, &ctx->r_block_hash,
                           &ctx->summary, &ctx->rows, &ctx->blk,
                           &ctx->segtab, &ctx->fine_start,
                           &ctx->tile_desc, &ctx->chunk_tab, &ctx->bstatus, &ctx->unit_wtot, &ctx->units, &ctx->rec[0], &ctx->rec[1], &ctx->rhi[0],
                           &ctx->rhi[1], &ctx->seq_lut, &ctx->g_seq_start, &ctx->g_seq_caller,
                           &ctx->g_seq_lut, &ctx->recv_rec, &ctx->recv_hi,
                           &ctx->peer_tab, &ctx->frag_pat, &ctx->frag_tiles, &ctx->frag_out,
                           &ctx->raw, &ctx->ingest, &ctx->bh_in, &ctx->bh_names,
                           &ctx->bh_scratch, &ctx->bh_out, &ctx->r_pay};
}

}  // namespace

// Reads back the guard bytes behind every device buffer of the context (after waiting for
// its stream).  *n_damaged = buffers whose guard was overwritten, i.e. some kernel stored
// past the end of that buffer; the last-error text names them.
extern "C" int npge_anchor_debug_check(npge_anchor_ctx* ctx, uint32_t* n_damaged) {
    if (!ctx || !n_damaged) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    *n_damaged = 0;
    NPGE_CUDA(cudaSetDevice(ctx->device));
    NPGE_CUDA(cudaStreamSynchronize(ctx->stream));
    std::vector<DeviceBuffer*> all = all_buffers(ctx);
    std::string which;
    uint8_t tail[kCanary];
    for (size_t i = 0; i < all.size(); i++) {
        if (!all[i]->ptr) continue;
        NPGE_CUDA(cudaMemcpy(tail, static_cast<uint8_t*>(all[i]->ptr) + all[i]->bytes, kCanary,
                             cudaMemcpyDeviceToHost));
        for (size_t j = 0; j < kCanary; j++) {
            if (tail[j] != 0xA5) {
                *n_damaged += 1;
                which += " #" + std::to_string(i) + "@+" + std::to_string(j);
                break;
            }
        }
    }
    if (*n_damaged) g_last_error = "guard bytes overwritten behind buffer(s)" + which;
    return NPGE_ANCHOR_OK;
}

extern "C" void npge_anchor_destroy(npge_anchor_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    std::vector<DeviceBuffer*> all = all_buffers(ctx);
    for (void* q : ctx->peer_opened) cudaIpcCloseMemHandle(q);
    for (DeviceBuffer* b : all) release(*b);
    for (auto& ev : ctx->ev) if (ev) cudaEventDestroy(ev);
    for (auto& ev : ctx->ev_pass) if (ev) cudaEventDestroy(ev);
    for (auto& ev : ctx->fetch_ev) if (ev) cudaEventDestroy(ev);
    for (auto& ev : ctx->copy_events) if (ev) cudaEventDestroy(ev);
    for (auto& ev : ctx->chunk_events) if (ev) cudaEventDestroy(ev);
    if (ctx->fetch_stream) cudaStreamDestroy(ctx->fetch_stream);
    if (ctx->pipe_carry) cudaFreeHost(ctx->pipe_carry);
    for (auto& cs : ctx->copy_streams) if (cs) cudaStreamDestroy(cs);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

// ---- upload: ATGCN text -> packed device genome ---------------------------------
namespace {

// Lays the sequences out in scan order (sizes[s] = clean size of caller sequence
// s), lets `fill` write the ASCII bytes of every kept sequence at its start
// (ctx->seqs: caller_index, size, start), then packs and builds the tables.
int upload_common(npge_anchor_ctx* ctx, int32_t anchor_size, const uint64_t* sizes,
                  uint32_t n_seqs, const std::function<int(cudaStream_t)>& fill) {
    ctx->uploaded = false;
    ctx->have_result = false;
    // SeqBase::make_seqs (SeqI.hpp:45-57): size >= k, by size descending; ties
    // by caller index (the reference's are unspecified)
    ctx->seqs.clear();
    for (uint32_t s = 0; s < n_seqs; s++) {
        uint64_t size = sizes[s];
        if (size >= uint64_t(anchor_size)) ctx->seqs.push_back({s, size, 0});
    }
    std::stable_sort(ctx->seqs.begin(), ctx->seqs.end(),
                     [](const SeqEntry& a, const SeqEntry& b) { return a.size > b.size; });
    uint64_t pos = 0, n_bases = 0;
    for (SeqEntry& e : ctx->seqs) {
        e.start = pos;
        pos += e.size + 1;  // one N pad after every sequence
        n_bases += e.size;
    }
    if (pos > kMaxPositions) {
        return fail(NPGE_ANCHOR_ERR_CAPACITY,
                    "more than 2^31-1 positions per device; shard the input");
    }
    ctx->n_positions = pos;
    ctx->n_bases = n_bases;
    ctx->n_padded = (pos + kPosAlign - 1) / kPosAlign * kPosAlign;
    ctx->n_caller_seqs = n_seqs;
    ctx->upload_k = anchor_size;
    const uint64_t np = ctx->n_padded;
    const size_t slack = 256;
    NPGE_TRY(ensure(ctx->ascii, np + slack));
    NPGE_TRY(ensure(ctx->packed, np / 4 + slack));
    NPGE_TRY(ensure(ctx->nmask, np / 8 + slack));
    NPGE_TRY(ensure(ctx->seq_start, (ctx->seqs.size() + 1) * 8));
    NPGE_TRY(ensure(ctx->seq_caller, (ctx->seqs.size() + 1) * 4));
    NPGE_TRY(ensure(ctx->counters, kCntWords * 4));
    cudaStream_t st = ctx->stream;
    NPGE_CUDA(cudaMemsetAsync(ctx->counters.ptr, 0, kCntWords * 4, st));
    if (np) {
        NPGE_CUDA(cudaMemsetAsync(ctx->ascii.ptr, 'N', np + slack, st));
        NPGE_CUDA(cudaMemsetAsync(ptr<uint8_t>(ctx->packed) + np / 4, 0, slack, st));
        NPGE_CUDA(cudaMemsetAsync(ptr<uint8_t>(ctx->nmask) + np / 8, 0xff, slack, st));
        NPGE_TRY(fill(st));
        std::vector<uint64_t> starts(ctx->seqs.size() + 1);
        std::vector<uint32_t> callers(ctx->seqs.size() + 1, 0);
        for (size_t i = 0; i < ctx->seqs.size(); i++) {
            starts[i] = ctx->seqs[i].start;
            callers[i] = ctx->seqs[i].caller_index;
        }
        starts[ctx->seqs.size()] = pos;
        NPGE_CUDA(cudaMemcpyAsync(ctx->seq_start.ptr, starts.data(),
                                  starts.size() * 8, cudaMemcpyHostToDevice, st));
        NPGE_CUDA(cudaMemcpyAsync(ctx->seq_caller.ptr, callers.data(),
                                  callers.size() * 4, cudaMemcpyHostToDevice, st));
        std::vector<uint2> pairs(starts.size());
        for (size_t i = 0; i < starts.size(); i++) pairs[i] = make_uint2(uint32_t(starts[i]), callers[i]);
        NPGE_TRY(ensure(ctx->seq_pair, pairs.size() * 8));
        NPGE_CUDA(cudaMemcpyAsync(ctx->seq_pair.ptr, pairs.data(), pairs.size() * 8,
                                  cudaMemcpyHostToDevice, st));
        // seq_lut[q] = scan-order sequence holding position q << lut_shift
        // (pads count for the sequence before them)
        const uint64_t n_chunks = (np >> ctx->lut_shift) + 2;
        std::vector<uint32_t> lut(n_chunks);
        {
            size_t sidx = 0;
            for (uint64_t q = 0; q < n_chunks; q++) {
                const uint64_t g = q << ctx->lut_shift;
                while (sidx + 1 < ctx->seqs.size() && ctx->seqs[sidx + 1].start <= g) sidx++;
                lut[q] = uint32_t(sidx);
            }
        }
        ctx->h_lut = lut;
        ctx->h_caller = callers;
        ctx->h_start32.resize(starts.size());
        for (size_t i = 0; i < starts.size(); i++) ctx->h_start32[i] = uint32_t(starts[i]);
        NPGE_TRY(ensure(ctx->seq_lut, n_chunks * 4));
        NPGE_CUDA(cudaMemcpyAsync(ctx->seq_lut.ptr, lut.data(), n_chunks * 4,
                                  cudaMemcpyHostToDevice, st));
        NPGE_CUDA(cudaStreamSynchronize(st));  // lut is a local
        launch_pack(ptr<uint8_t>(ctx->ascii), ptr<uint32_t>(ctx->packed),
                    ptr<uint32_t>(ctx->nmask), np, ptr<uint32_t>(ctx->counters), st);
        NPGE_CUDA(cudaGetLastError());
        uint32_t bad = 0;
        NPGE_CUDA(cudaMemcpyAsync(&bad, ptr<uint32_t>(ctx->counters) + kCntBadInput,
                                  4, cudaMemcpyDeviceToHost, st));
        NPGE_CUDA(cudaStreamSynchronize(st));
        if (bad) {
            return fail(NPGE_ANCHOR_ERR_INPUT,
                        std::to_string(bad) + " bytes outside ATGCN: run "
                        "Sequence::to_atgcn on the input first");
        }
    }
    ctx->uploaded = true;
    ctx->dist = false;
    ctx->gseqs = ctx->seqs;
    ctx->g_positions = ctx->n_positions;
    ctx->g_padded = ctx->n_padded;
    return NPGE_ANCHOR_OK;
}

}  // namespace

extern "C" int npge_anchor_upload(npge_anchor_ctx* ctx, int32_t anchor_size,
                                  const uint8_t* bases,
                                  const uint64_t* seq_offsets, uint32_t n_seqs) {
    if (!ctx) return fail(NPGE_ANCHOR_ERR_PARAMS, "ctx is NULL");
    npge_anchor_params probe = {anchor_size, 0, 0};
    NPGE_TRY(check_params(&probe));
    if (n_seqs && (!seq_offsets || (!bases && seq_offsets[n_seqs] > 0))) {
        return fail(NPGE_ANCHOR_ERR_PARAMS, "bases / seq_offsets is NULL");
    }
    std::vector<uint64_t> sizes(n_seqs);
    for (uint32_t s = 0; s < n_seqs; s++) {
        if (seq_offsets[s + 1] < seq_offsets[s]) {
            return fail(NPGE_ANCHOR_ERR_PARAMS, "seq_offsets must be non-decreasing");
        }
        sizes[s] = seq_offsets[s + 1] - seq_offsets[s];
    }
    NPGE_CUDA(cudaSetDevice(ctx->device));
    return upload_common(ctx, anchor_size, sizes.data(), n_seqs, [&](cudaStream_t st) -> int {
        for (const SeqEntry& e : ctx->seqs) {
            // (cudaMemcpyDefault: `bases` may also be device memory, e.g. the
            // sequences a multi-GPU job has just exchanged)
            NPGE_CUDA(cudaMemcpyAsync(ptr<uint8_t>(ctx->ascii) + e.start,
                                      bases + seq_offsets[e.caller_index], e.size,
                                      cudaMemcpyDefault, st));
        }
        return NPGE_ANCHOR_OK;
    });
}

// ---- upload_raw: to_atgcn on the device ---------------------------
extern "C" int npge_anchor_upload_raw(npge_anchor_ctx* ctx, int32_t anchor_size,
                                      const uint8_t* raw, const uint64_t* raw_offsets,
                                      uint32_t n_seqs, uint64_t* clean_sizes) {
    if (!ctx) return fail(NPGE_ANCHOR_ERR_PARAMS, "ctx is NULL");
    npge_anchor_params probe = {anchor_size, 0, 0};
    NPGE_TRY(check_params(&probe));
    if (n_seqs && (!raw_offsets || (!raw && raw_offsets[n_seqs] > 0))) {
        return fail(NPGE_ANCHOR_ERR_PARAMS, "raw / raw_offsets is NULL");
    }
    for (uint32_t s = 0; s < n_seqs; s++) {
        if (raw_offsets[s + 1] < raw_offsets[s]) {
            return fail(NPGE_ANCHOR_ERR_PARAMS, "raw_offsets must be non-decreasing");
        }
    }
    NPGE_CUDA(cudaSetDevice(ctx->device));
    ctx->uploaded = false;
    const uint64_t n_raw = n_seqs ? raw_offsets[n_seqs] - raw_offsets[0] : 0;
    const uint32_t tiles = ingest_tiles(n_raw);
    cudaStream_t st = ctx->stream;
    NPGE_TRY(ensure(ctx->raw, n_raw + 16));
    NPGE_TRY(ensure(ctx->ingest, size_t(tiles + 1) * 12 + (size_t(n_seqs) + 1) * 24 + 64));
    uint8_t* w = ptr<uint8_t>(ctx->ingest);
    IngestArgs a = {};
    a.raw = ptr<uint8_t>(ctx->raw);
    a.n_raw = n_raw;
    a.n_seqs = n_seqs;
    a.tile_offset = reinterpret_cast<uint64_t*>(w);
    a.seq_rank = a.tile_offset + tiles + 1;
    uint64_t* d_offsets = a.seq_rank + n_seqs + 1;
    uint64_t* d_base = d_offsets + n_seqs + 1;
    a.raw_offsets = d_offsets;
    a.dst_base = d_base;
    a.tile_count = reinterpret_cast<uint32_t*>(d_base + n_seqs + 1);
    std::vector<uint64_t> rel(n_seqs + 1, 0);  // offsets relative to the first byte
    for (uint32_t s = 0; s <= n_seqs && n_seqs; s++) rel[s] = raw_offsets[s] - raw_offsets[0];
    if (n_raw) {
        NPGE_CUDA(cudaMemcpyAsync(ctx->raw.ptr, raw + raw_offsets[0], n_raw,
                                  cudaMemcpyHostToDevice, st));
    }
    NPGE_CUDA(cudaMemcpyAsync(d_offsets, rel.data(), rel.size() * 8, cudaMemcpyHostToDevice, st));
    uint32_t launches = 0;
    launch_ingest(a, 0, st, &launches);
    NPGE_CUDA(cudaGetLastError());
    std::vector<uint64_t> rank(n_seqs + 1, 0), sizes(n_seqs, 0);
    NPGE_CUDA(cudaMemcpyAsync(rank.data(), a.seq_rank, rank.size() * 8, cudaMemcpyDeviceToHost, st));
    NPGE_CUDA(cudaStreamSynchronize(st));
    for (uint32_t s = 0; s < n_seqs; s++) sizes[s] = rank[s + 1] - rank[s];
    if (clean_sizes) std::copy(sizes.begin(), sizes.end(), clean_sizes);
    return upload_common(ctx, anchor_size, sizes.data(), n_seqs, [&](cudaStream_t s2) -> int {
        std::vector<uint64_t> base(n_seqs + 1, ~0ull);  // dropped sequences: no place
        for (const SeqEntry& e : ctx->seqs) base[e.caller_index] = e.start;
        IngestArgs b = a;
        b.ascii = ptr<uint8_t>(ctx->ascii);
        NPGE_CUDA(cudaMemcpyAsync(d_base, base.data(), base.size() * 8,
                                  cudaMemcpyHostToDevice, s2));
        launch_ingest(b, 1, s2, &launches);
        NPGE_CUDA(cudaGetLastError());
        NPGE_CUDA(cudaStreamSynchronize(s2));  // `base` is a local
        return NPGE_ANCHOR_OK;
    });
}

// ---- find_device, MSD partition + shared-memory bucket sort ---------------------------
namespace {

// The arguments of the bucket-sort kernels from the context's buffers.
BucketArgs bucket_args(npge_anchor_ctx* ctx, const MsdPlan& plan, const MsdBuffers& mb,
                       uint64_t* scratch_rec, uint32_t* scratch_hi, uint64_t n_used,
                       uint64_t limit) {
    BucketArgs a = {};
    a.plan = plan;
    a.rec_w = mb.rec[1];
    a.hi_w = mb.hi[1];
    a.scratch_rec = scratch_rec;
    a.scratch_hi = scratch_hi;
    a.fine_start = mb.fine_start;
    a.units = static_cast<uint4*>(ctx->units.ptr);
    a.counters = ptr<uint32_t>(ctx->counters);
    a.used = ptr<uint64_t>(ctx->used);
    a.n_used = n_used;
    if (ctx->dist) {  // positions are job-wide: map them with the job-wide table
        a.seq_start = ptr<uint64_t>(ctx->g_seq_start);
        a.seq_caller = ptr<uint32_t>(ctx->g_seq_caller);
        a.seq_lut = ptr<uint32_t>(ctx->g_seq_lut);
        a.seq_pair = ptr<uint2>(ctx->g_seq_pair);
    } else {
        a.seq_start = ptr<uint64_t>(ctx->seq_start);
        a.seq_caller = ptr<uint32_t>(ctx->seq_caller);
        a.seq_lut = ptr<uint32_t>(ctx->seq_lut);
        a.seq_pair = ptr<uint2>(ctx->seq_pair);
    }
    a.lut_shift = ctx->lut_shift;
    a.max_fragments = limit;
    a.found_bits = ptr<uint32_t>(ctx->found_bits);
    a.sel = ptr<uint8_t>(ctx->sel);
    a.r_seq = ptr<uint32_t>(ctx->r_seq);
    a.r_pos = ptr<uint32_t>(ctx->r_pos);
    a.r_ori = ptr<int8_t>(ctx->r_ori);
    a.r_block_off = ptr<uint64_t>(ctx->r_block_off);
    a.r_block_hash = ptr<uint64_t>(ctx->r_block_hash);
    a.unit_tot = ptr<uint64_t>(ctx->bstatus);
    a.unit_wtot = ptr<uint32_t>(ctx->unit_wtot);
    a.summary = ptr<uint64_t>(ctx->summary);
    return a;
}

int fix_cut_resident(npge_anchor_ctx* ctx, uint64_t limit);

constexpr int kPipeRanges = 8;                           // unit ranges of the pipelined call
constexpr uint64_t kPipeMaxRecords = uint64_t(1) << 28;  // above: 3 GiB+ of pinned buffers

int find_device_msd(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                    const uint64_t* used_sorted, uint64_t n_used,
                    npge_anchor_stats* stats) {
    const int k = p->anchor_size;
    const bool similar = p->anchor_similar != 0;
    const uint64_t np = ctx->n_padded;
    const uint64_t max_records = ctx->n_positions;  // upper bound on windows
    if (max_records > kMaxRecords) {
        return fail(NPGE_ANCHOR_ERR_CAPACITY,
                    "more than 2^30-1 windows per device; shard the input");
    }
    const MsdPlan plan = make_msd_plan(k, max_records, ctx->force_f, ctx->force_wa);
    const uint64_t limit = uint64_t(p->max_anchor_fragments);
    const uint64_t max_frags = std::min<uint64_t>(max_records, limit);
    const size_t bins_a = size_t(1) << plan.capa;
    const size_t summary_words = bucket_summary_words();
    const uint32_t max_win = bucket_max_windows(plan, max_records);
    const uint32_t max_seg = uint32_t(bins_a);
    NPGE_TRY(ensure(ctx->rows, msd_rows_needed(plan, np, max_records, max_seg) * 4));
    NPGE_TRY(ensure(ctx->blk, msd_blk_needed(plan, np, max_records, max_seg) * 4));
    NPGE_TRY(ensure(ctx->segtab, seg_table_words(max_seg, plan) * 4));
    NPGE_TRY(ensure(ctx->fine_start, ((bins_a << plan.wb) + 1) * 4));
    NPGE_TRY(ensure(ctx->tile_desc, (msd_tiles_b(max_records, max_seg) + 1) * 16));
    for (int b = 0; b < 2; b++) {
        NPGE_TRY(ensure(ctx->rec[b], (max_records + 16) * 8));
        if (plan.wide) NPGE_TRY(ensure(ctx->rhi[b], (max_records + 16) * 4));
    }
    NPGE_TRY(ensure(ctx->bstatus, size_t(max_win + 1) * 8 + bins_a * 4));
    NPGE_TRY(ensure(ctx->unit_wtot, size_t(max_win + 1) * (kSortThreads / 32) * 4));
    NPGE_TRY(ensure(ctx->units, size_t(max_win + 1) * 16));
    NPGE_TRY(ensure(ctx->counters, kCntWords * 4));
    NPGE_TRY(ensure(ctx->summary, summary_words * 8));
    NPGE_TRY(ensure(ctx->used, n_used * 8));
    const bool compact = ctx->layout == 1;  // payloads only (npge_anchor_set_result_layout)
    const bool with_pay = compact || fetch_mode() != kFetchPlain;
    if (!compact) {
        NPGE_TRY(ensure(ctx->r_seq, max_frags * 4));
        NPGE_TRY(ensure(ctx->r_pos, max_frags * 4));
        NPGE_TRY(ensure(ctx->r_ori, max_frags));
    }
    if (with_pay) NPGE_TRY(ensure(ctx->r_pay, max_frags * 4));
    NPGE_TRY(ensure(ctx->r_block_off, (max_frags / 2 + 2) * 8));
    NPGE_TRY(ensure(ctx->r_block_hash, (max_frags / 2 + 2) * 8));
    if (similar) {
        NPGE_TRY(ensure(ctx->found_bits, np / 8 + 256));
        NPGE_TRY(ensure(ctx->sel, max_records + 16));
    }
    cudaStream_t st = ctx->stream;
    uint32_t launches = 0;
    if (n_used) {
        NPGE_CUDA(cudaMemcpyAsync(ctx->used.ptr, used_sorted, n_used * 8,
                                  cudaMemcpyHostToDevice, st));
    }
    NPGE_CUDA(cudaEventRecord(ctx->ev[0], st));
    NPGE_CUDA(cudaMemsetAsync(ctx->counters.ptr, 0, kCntWords * 4, st));
    NPGE_CUDA(cudaMemsetAsync(ctx->summary.ptr, 0, summary_words * 8, st));
    if (similar) {
        NPGE_CUDA(cudaMemsetAsync(ctx->found_bits.ptr, 0, np / 8 + 256, st));
        NPGE_CUDA(cudaMemsetAsync(ctx->sel.ptr, 0, max_records + 16, st));
    }
    MsdBuffers mb = {};
    mb.rows = ptr<uint32_t>(ctx->rows);
    mb.blk = ptr<uint32_t>(ctx->blk);
    mb.seg = carve_seg_tables(ptr<uint32_t>(ctx->segtab), max_seg, plan);
    mb.fine_start = ptr<uint32_t>(ctx->fine_start);
    mb.tile_desc = ptr<uint4>(ctx->tile_desc);
    for (int b = 0; b < 2; b++) {
        mb.rec[b] = ptr<uint64_t>(ctx->rec[b]);
        mb.hi[b] = ptr<uint32_t>(ctx->rhi[b]);
    }
    if (np) {
        PosMap pm;
        launch_msd_pass_a(plan, ptr<uint32_t>(ctx->packed), ptr<uint32_t>(ctx->nmask),
                          np, mb, pm, ptr<uint32_t>(ctx->counters), st, ctx->ev_pass,
                          &launches);
        launch_msd_pass_b(plan, mb.rec[0], mb.hi[0], max_records, max_seg, mb, st,
                          ctx->ev_pass, &launches);
    } else {
        for (int i = 0; i < 4; i++) NPGE_CUDA(cudaEventRecord(ctx->ev_pass[i], st));
    }
    NPGE_CUDA(cudaEventRecord(ctx->ev[2], st));
    BucketArgs a = bucket_args(ctx, plan, mb, mb.rec[0], mb.hi[0], n_used, limit);
    a.r_pay = with_pay ? ptr<uint32_t>(ctx->r_pay) : nullptr;
    if (compact) {
        a.r_seq = nullptr;
        a.r_pos = nullptr;
        a.r_ori = nullptr;
    }
    // (no cut possible and no second sweep: the result of a unit range is final as soon
    // as that range has been emitted)
    const bool piped = ctx->pipe && compact && !similar && np && limit >= max_records &&
                       max_records <= kPipeMaxRecords;
    if (ctx->pipe) ctx->pipe->used = piped;
    if (piped) {
        launch_bucket_stage(a, kBucketUnits, max_records, st, &launches);
        uint32_t n_units = 0;
        NPGE_CUDA(cudaMemcpyAsync(&n_units, ptr<uint32_t>(ctx->counters) + kCntWindows, 4,
                                  cudaMemcpyDeviceToHost, st));
        NPGE_CUDA(cudaStreamSynchronize(st));
        if (!ctx->fetch_stream) {
            NPGE_CUDA(cudaStreamCreateWithFlags(&ctx->fetch_stream, cudaStreamNonBlocking));
        }
        if (!ctx->pipe_carry) {
            NPGE_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&ctx->pipe_carry), kPipeRanges * 8,
                                    cudaHostAllocDefault));
        }
        while (ctx->fetch_ev.size() < size_t(kPipeRanges)) {
            cudaEvent_t ev = nullptr;
            NPGE_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
            ctx->fetch_ev.push_back(ev);
        }
        const uint32_t per = (n_units + kPipeRanges - 1) / kPipeRanges;
        int n_ranges = 0;
        for (int c = 0; c < kPipeRanges; c++) {
            const uint32_t u0 = uint32_t(c) * per, u1 = std::min(n_units, u0 + per);
            if (u0 >= u1) break;
            a.unit_first = u0;
            a.unit_end = u1;
            launch_bucket_stage(a, kBucketSort, max_records, st, &launches);
            launch_bucket_stage(a, kBucketScan, max_records, st, &launches);
            launch_bucket_stage(a, kBucketEmit, max_records, st, &launches);
            NPGE_CUDA(cudaMemcpyAsync(&ctx->pipe_carry[c], ptr<uint64_t>(ctx->summary) + kSumCarry,
                                      8, cudaMemcpyDeviceToHost, st));
            NPGE_CUDA(cudaEventRecord(ctx->fetch_ev[c], st));
            n_ranges++;
        }
        a.unit_first = 0;
        a.unit_end = 0;
        if (n_ranges == 0) launch_bucket_stage(a, kBucketScan, max_records, st, &launches);
        NPGE_CUDA(cudaEventRecord(ctx->ev[4], st));
        launch_bucket_stage(a, kBucketFinal, max_records, st, &launches);
        NPGE_CUDA(cudaGetLastError());
        // every finished range goes to the host while the next ones are sorted
        uint64_t f0 = 0, b0 = 0;
        for (int c = 0; c < n_ranges; c++) {
            NPGE_CUDA(cudaEventSynchronize(ctx->fetch_ev[c]));
            const uint64_t f1 = ctx->pipe_carry[c] & 0x7fffffffull, b1 = ctx->pipe_carry[c] >> 31;
            cudaStream_t fs = ctx->fetch_stream;
            if (f1 > f0) {
                NPGE_CUDA(cudaMemcpyAsync(ctx->pipe->pos_ori + f0, ptr<uint32_t>(ctx->r_pay) + f0,
                                          (f1 - f0) * 4, cudaMemcpyDeviceToHost, fs));
            }
            if (b1 > b0) {
                NPGE_CUDA(cudaMemcpyAsync(ctx->pipe->block_off + b0,
                                          ptr<uint64_t>(ctx->r_block_off) + b0, (b1 - b0) * 8,
                                          cudaMemcpyDeviceToHost, fs));
                NPGE_CUDA(cudaMemcpyAsync(ctx->pipe->block_hash + b0,
                                          ptr<uint64_t>(ctx->r_block_hash) + b0, (b1 - b0) * 8,
                                          cudaMemcpyDeviceToHost, fs));
            }
            f0 = f1;
            b0 = b1;
        }
        ctx->pipe->block_off[b0] = f0;  // the closing offset
    } else {
        launch_bucket(a, similar, np ? max_records : 0, st, &ctx->ev[4], &launches);
        NPGE_CUDA(cudaGetLastError());
    }
    NPGE_CUDA(cudaEventRecord(ctx->ev[3], st));
    uint64_t summary[kSumWords];
    uint32_t counters_h[kCntWords];
    NPGE_CUDA(cudaMemcpyAsync(summary, ctx->summary.ptr, sizeof(summary),
                              cudaMemcpyDeviceToHost, st));
    NPGE_CUDA(cudaMemcpyAsync(counters_h, ctx->counters.ptr, sizeof(counters_h),
                              cudaMemcpyDeviceToHost, st));
    NPGE_CUDA(cudaStreamSynchronize(st));
    npge_anchor_stats& s = ctx->stats;
    memset(&s, 0, sizeof(s));
    s.n_positions = ctx->n_bases;
    s.n_windows = counters_h[kCntValid];
    s.n_fragments = summary[kSumEmitted];
    s.n_blocks = summary[kSumBlocksEmitted];
    s.digest[0] = summary[kSumDigest0];
    s.digest[1] = summary[kSumDigest1];
    s.radix_passes = uint32_t(2 + plan.smem_passes);
    s.gpu_launches = launches;
    // stages: [count A + scan] [pass A] [count B + scan] [pass B] [bucket sort] [emit]
    cudaEventElapsedTime(&s.ms_hist, ctx->ev[0], ctx->ev_pass[0]);
    cudaEventElapsedTime(&s.ms_pass[0], ctx->ev_pass[0], ctx->ev_pass[1]);
    cudaEventElapsedTime(&s.ms_pass[1], ctx->ev_pass[1], ctx->ev_pass[2]);
    cudaEventElapsedTime(&s.ms_pass[2], ctx->ev_pass[2], ctx->ev_pass[3]);
    cudaEventElapsedTime(&s.ms_pass[3], ctx->ev[2], ctx->ev[4]);  // bucket sort
    cudaEventElapsedTime(&s.ms_pass[4], ctx->ev[4], ctx->ev[3]);  // scan + emit
    cudaEventElapsedTime(&s.ms_sort, ctx->ev[0], ctx->ev[2]);
    cudaEventElapsedTime(&s.ms_group, ctx->ev[2], ctx->ev[3]);
    cudaEventElapsedTime(&s.ms_total, ctx->ev[0], ctx->ev[3]);
    const int model_passes = (2 * k + 7) / 8;
    s.bytes_algorithmic = uint64_t(double(ctx->n_bases) * (32.375 + 24.0 * model_passes)) +
                          9 * s.n_fragments + 16 * s.n_blocks;
    // compulsory traffic of this build: genome read twice; records written by
    // pass A, read by count B and pass B, written by pass B, read by the bucket
    // sort (twice more in anchor-similar mode); results written
    const double rec = plan.wide ? 12.0 : 8.0;
    s.bytes_moved = uint64_t(double(ctx->n_bases) * 0.75 +
                             double(s.n_windows) * rec * (similar ? 7.0 : 5.0)) +
                    (compact ? 4 : 9) * s.n_fragments + 16 * s.n_blocks;
    ctx->total_selected = summary[kSumFragments];
    ctx->plan = plan;
    ctx->have_pay = with_pay;
    ctx->result_compact = compact;
    ctx->last_params = *p;
    NPGE_TRY(fix_cut_resident(ctx, limit));
    ctx->have_result = true;
    if (stats) *stats = s;
    return NPGE_ANCHOR_OK;
}

}  // namespace

// ---- find_device: hist -> radix sort -> grouping, all on the device ---------------
extern "C" int npge_anchor_find_device(npge_anchor_ctx* ctx,
                                       const npge_anchor_params* p,
                                       const uint64_t* used_sorted,
                                       uint64_t n_used,
                                       npge_anchor_stats* stats) {
    if (!ctx) return fail(NPGE_ANCHOR_ERR_PARAMS, "ctx is NULL");
    NPGE_TRY(check_params(p));
    if (!ctx->uploaded) {
        return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_upload has not been called");
    }
    if (p->anchor_size != ctx->upload_k) {
        return fail(NPGE_ANCHOR_ERR_STATE,
                    "anchor-size differs from the one given to npge_anchor_upload");
    }
    if (n_used && !used_sorted) {
        return fail(NPGE_ANCHOR_ERR_PARAMS, "used_sorted is NULL");
    }
    NPGE_CUDA(cudaSetDevice(ctx->device));
    ctx->have_result = false;
    return find_device_msd(ctx, p, used_sorted, n_used, stats);
}

extern "C" int npge_anchor_get_stats(npge_anchor_ctx* ctx,
                                     npge_anchor_stats* stats) {
    if (!ctx || !stats) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    if (!ctx->have_result) return fail(NPGE_ANCHOR_ERR_STATE, "no result yet");
    *stats = ctx->stats;
    return NPGE_ANCHOR_OK;
}

// ---- fetch: device result -> host arrays ------------------------------------------
namespace {

// The reference cuts the list sorted by (hash, Sequence*, pos') at
// max-anchor-fragments (AnchorFinder.cpp:361-363), pos' = pos + size for the
// reverse strand (:295-301).  The device order inside a group is the scan order,
// so only the group the cut falls into can differ.  Right after the emit kernels
// (find_device / dist_emit have synchronised anyway) the kept members of that one
// group are rebuilt in the reference's order and patched into the resident result,
// a group cut down to one member is dropped (a singleton of the reference's list:
// no block, but its hash still counts as used, :385), and the digest is corrected:
// stats, resident arrays and npge_anchor_fetch all describe the same answer.
// The unit holding the cut record was written back sorted; its bounds and the cut
// record's index are in the counters.
int fix_cut_resident(npge_anchor_ctx* ctx, uint64_t limit) {
    ctx->lone_tail = false;
    npge_anchor_stats& st = ctx->stats;
    if (!(ctx->total_selected > limit && limit > 0) || st.n_blocks == 0) return NPGE_ANCHOR_OK;
    cudaStream_t s = ctx->stream;
    uint32_t c[kCntWords];
    uint64_t tail[2] = {0, 0};  // offset and hash of the last block
    NPGE_CUDA(cudaMemcpyAsync(c, ctx->counters.ptr, sizeof(c), cudaMemcpyDeviceToHost, s));
    NPGE_CUDA(cudaMemcpyAsync(&tail[0], ptr<uint64_t>(ctx->r_block_off) + st.n_blocks - 1, 8,
                              cudaMemcpyDeviceToHost, s));
    NPGE_CUDA(cudaMemcpyAsync(&tail[1], ptr<uint64_t>(ctx->r_block_hash) + st.n_blocks - 1, 8,
                              cudaMemcpyDeviceToHost, s));
    NPGE_CUDA(cudaStreamSynchronize(s));
    const uint32_t cut = c[kCntCutIndex], cs = c[kCntCutStart], cn = c[kCntCutSize];
    if (cn == 0 || cut < cs || cut >= cs + cn) {
        return fail(NPGE_ANCHOR_ERR_CUDA, "internal: cut record outside its unit");
    }
    std::vector<uint64_t> rec(cn);
    std::vector<uint32_t> hi(ctx->plan.wide ? cn : 0);
    NPGE_CUDA(cudaMemcpyAsync(rec.data(), ptr<uint64_t>(ctx->rec[1]) + cs, size_t(cn) * 8,
                              cudaMemcpyDeviceToHost, s));
    if (ctx->plan.wide) {
        NPGE_CUDA(cudaMemcpyAsync(hi.data(), ptr<uint32_t>(ctx->rhi[1]) + cs, size_t(cn) * 4,
                                  cudaMemcpyDeviceToHost, s));
    }
    NPGE_CUDA(cudaStreamSynchronize(s));
    auto K = [&](uint32_t i) {
        return (uint64_t(ctx->plan.wide ? hi[i] : 0u) << 32) | (rec[i] >> 32);
    };
    const uint32_t ci = cut - cs;
    uint32_t gs = ci, ge = ci + 1;
    while (gs > 0 && K(gs - 1) == K(ci)) gs--;
    while (ge < cn && K(ge) == K(ci)) ge++;
    const uint64_t first = tail[0], key = tail[1];
    if (first >= limit || ci - gs + 1 != limit - first) {
        return fail(NPGE_ANCHOR_ERR_CUDA, "internal: cut group is not the last block");
    }
    const uint64_t keep = limit - first;
    struct Member { uint32_t seq; uint64_t pos2; uint32_t pos; int8_t ori; uint32_t pay; };
    auto member = [&](uint32_t i) {
        const uint32_t pw = uint32_t(rec[i]);
        const uint64_t g = pw >> 1;
        size_t a = 0, b = ctx->gseqs.size();
        while (b - a > 1) {
            size_t mid = (a + b) / 2;
            if (ctx->gseqs[mid].start <= g) a = mid; else b = mid;
        }
        const SeqEntry& e = ctx->gseqs[a];
        const uint32_t pos = uint32_t(g - e.start);
        const bool direct = pw & 1;
        return Member{e.caller_index, direct ? pos : pos + e.size, pos, int8_t(direct ? 1 : -1), pw};
    };
    auto term = [&](const Member& m, uint64_t& d0, uint64_t& d1, bool add) {
        const uint64_t t = (uint64_t(m.seq) << 33) | (uint64_t(m.pos) << 1) | (m.ori == 1 ? 1 : 0);
        uint64_t x, y;
        digest_terms(t, key, x, y);
        d0 += add ? x : 0 - x;
        d1 += add ? y : 0 - y;
    };
    std::vector<Member> members;
    members.reserve(ge - gs);
    for (uint32_t i = gs; i < ge; i++) members.push_back(member(i));
    uint64_t d0 = st.digest[0], d1 = st.digest[1];
    for (uint64_t j = 0; j < keep; j++) term(members[j], d0, d1, false);  // the device's choice
    std::sort(members.begin(), members.end(), [](const Member& x, const Member& y) {
        return x.seq != y.seq ? x.seq < y.seq : x.pos2 < y.pos2;
    });
    if (keep == 1) {
        ctx->lone_tail = true;
        ctx->lone_hash = key;
        st.n_blocks -= 1;      // r_block_off[n_blocks] == first == the new fragment count
        st.n_fragments -= 1;
    } else {
        std::vector<uint32_t> q(keep), o(keep), y(keep);
        std::vector<int8_t> r(keep);
        for (uint64_t j = 0; j < keep; j++) {
            term(members[j], d0, d1, true);
            q[j] = members[j].seq;
            o[j] = members[j].pos;
            r[j] = members[j].ori;
            y[j] = members[j].pay;
        }
        if (!ctx->result_compact) {
            NPGE_CUDA(cudaMemcpyAsync(ptr<uint32_t>(ctx->r_seq) + first, q.data(), keep * 4,
                                      cudaMemcpyHostToDevice, s));
            NPGE_CUDA(cudaMemcpyAsync(ptr<uint32_t>(ctx->r_pos) + first, o.data(), keep * 4,
                                      cudaMemcpyHostToDevice, s));
            NPGE_CUDA(cudaMemcpyAsync(ptr<int8_t>(ctx->r_ori) + first, r.data(), keep,
                                      cudaMemcpyHostToDevice, s));
        }
        if (ctx->have_pay) {
            NPGE_CUDA(cudaMemcpyAsync(ptr<uint32_t>(ctx->r_pay) + first, y.data(), keep * 4,
                                      cudaMemcpyHostToDevice, s));
        }
        NPGE_CUDA(cudaStreamSynchronize(s));
    }
    st.digest[0] = d0;
    st.digest[1] = d1;
    return NPGE_ANCHOR_OK;
}

}  // namespace

// ---- compact fetch ------------------------------------------------------------------
// The fragment arrays of a large result are 9 bytes per fragment and PCIe is the
// slowest link of the host-buffer call.  The emit kernel also leaves the 4-byte
// payload (position in the packed genome << 1 | strand) of every fragment; only
// that crosses the bus, in chunks, and host threads turn each chunk into
// (sequence, position, strand) while the next one is copied.
namespace {

// Persistent worker threads: creating a dozen std::threads per call costs about
// as much as the work they would do.
class WorkerPool {
public:
    static WorkerPool& instance() {
        static WorkerPool* pool = new WorkerPool;  // never destroyed: threads are detached
        return *pool;
    }
    // one job at a time: hold this from start() to wait()
    std::mutex& serial() { return serial_; }
    void start(const std::function<void(int)>& job) {
        std::lock_guard<std::mutex> lock(mu_);
        job_ = job;
        pending_ = n_;
        gen_++;
        cv_.notify_all();
    }
    void wait() {
        std::unique_lock<std::mutex> lock(mu_);
        done_.wait(lock, [this] { return pending_ == 0; });
    }
private:
    WorkerPool() {
        unsigned hw = std::thread::hardware_concurrency();
        int n = int(hw ? hw : 4) - 1;  // the caller's thread drives the copies
        if (const char* e = getenv("NPGE_FETCH_THREADS")) n = atoi(e);
        n_ = std::max(1, std::min(n, 32));
        for (int i = 0; i < n_; i++) std::thread([this, i] { loop(i); }).detach();
    }
    void loop(int index) {
        uint64_t seen = 0;
        while (true) {
            std::function<void(int)> job;
            {
                std::unique_lock<std::mutex> lock(mu_);
                cv_.wait(lock, [&] { return gen_ != seen; });
                seen = gen_;
                job = job_;
            }
            job(index);
            std::lock_guard<std::mutex> lock(mu_);
            if (--pending_ == 0) done_.notify_all();
        }
    }
    std::mutex mu_, serial_;
    std::condition_variable cv_, done_;
    std::function<void(int)> job_;
    uint64_t gen_ = 0;
    int pending_ = 0, n_ = 1;
};

constexpr uint64_t kFetchCompactMin = uint64_t(1) << 20;  // fragments; below: plain copies
constexpr uint64_t kFetchChunk = uint64_t(1) << 22;       // fragments per PCIe chunk
constexpr uint64_t kFetchPiece = uint64_t(1) << 16;       // fragments per unit of host work

bool fetch_compact_wanted(const npge_anchor_ctx* ctx, uint64_t nf, uint64_t limit) {
    (void)limit;
    if (ctx->result_compact) return true;  // there is nothing else to copy
    if (!ctx->have_pay || ctx->dist || nf == 0) return false;
    const int mode = fetch_mode();
    if (mode == kFetchPlain) return false;
    if (mode == kFetchCompact) return true;
    return nf >= kFetchCompactMin;  // auto: only where PCIe time dominates
}

int fetch_compact(npge_anchor_ctx* ctx, HostResult* r, uint64_t nf, uint64_t nb) {
    cudaStream_t st = ctx->stream;
    PinnedVector<uint32_t> pay;
    try {
        pay.resize(nf);
    } catch (std::bad_alloc&) {
        return fail(NPGE_ANCHOR_ERR_MEMORY, "out of host memory for the result");
    }
    const uint64_t n_chunks = (nf + kFetchChunk - 1) / kFetchChunk;
    while (ctx->fetch_ev.size() < n_chunks) {
        cudaEvent_t ev = nullptr;
        NPGE_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        ctx->fetch_ev.push_back(ev);
    }
    for (uint64_t c = 0; c < n_chunks; c++) {
        const uint64_t lo = c * kFetchChunk, hi = std::min(nf, lo + kFetchChunk);
        NPGE_CUDA(cudaMemcpyAsync(pay.data() + lo, ptr<uint32_t>(ctx->r_pay) + lo, (hi - lo) * 4,
                                  cudaMemcpyDeviceToHost, st));
        NPGE_CUDA(cudaEventRecord(ctx->fetch_ev[c], st));
    }
    NPGE_CUDA(cudaMemcpyAsync(r->block_offsets.data(), ctx->r_block_off.ptr, (nb + 1) * 8,
                              cudaMemcpyDeviceToHost, st));
    if (nb) {
        NPGE_CUDA(cudaMemcpyAsync(r->block_hash.data(), ctx->r_block_hash.ptr, nb * 8,
                                  cudaMemcpyDeviceToHost, st));
    }
    // the workers take pieces in order and wait until the chunk of their piece
    // has arrived; this thread announces the arrivals
    std::atomic<uint64_t> ready(0), next(0);
    std::atomic<bool> abort(false);
    const uint64_t n_pieces = (nf + kFetchPiece - 1) / kFetchPiece;
    uint32_t* seq = r->frag_seq.data();
    uint32_t* pos = r->frag_min_pos.data();
    int8_t* ori = r->frag_ori.data();
    const uint32_t* pay_h = pay.data();
    WorkerPool& pool = WorkerPool::instance();
    std::lock_guard<std::mutex> one_job(pool.serial());
    pool.start([&, ctx, nf, n_pieces, seq, pos, ori, pay_h](int) {
        while (true) {
            const uint64_t piece = next.fetch_add(1);
            if (piece >= n_pieces) return;
            const uint64_t lo = piece * kFetchPiece, hi = std::min(nf, lo + kFetchPiece);
            while (ready.load(std::memory_order_acquire) < hi) {
                if (abort.load()) return;
                std::this_thread::yield();
            }
            expand_payloads(pay_h, lo, hi, ctx->h_lut.data(), ctx->lut_shift,
                            ctx->h_start32.data(), ctx->h_caller.data(), seq, pos, ori);
        }
    });
    cudaError_t err = cudaSuccess;
    for (uint64_t c = 0; c < n_chunks && err == cudaSuccess; c++) {
        err = cudaEventSynchronize(ctx->fetch_ev[c]);
        if (err == cudaSuccess) {
            ready.store(std::min(nf, (c + 1) * kFetchChunk), std::memory_order_release);
        }
    }
    if (err != cudaSuccess) abort.store(true);
    pool.wait();  // the job refers to this frame: never return before it is done
    NPGE_CUDA(err);
    NPGE_CUDA(cudaStreamSynchronize(st));
    return NPGE_ANCHOR_OK;
}

}  // namespace

extern "C" int npge_anchor_fetch(npge_anchor_ctx* ctx, npge_anchor_result* out) {
    if (!ctx || !out) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    memset(out, 0, sizeof(*out));
    if (!ctx->have_result) {
        return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_find_device has not run");
    }
    NPGE_CUDA(cudaSetDevice(ctx->device));
    HostResult* r = new (std::nothrow) HostResult;
    if (!r) return fail(NPGE_ANCHOR_ERR_MEMORY, "out of host memory");
    const uint64_t nf = ctx->stats.n_fragments, nb = ctx->stats.n_blocks;
    const uint64_t limit = uint64_t(ctx->last_params.max_anchor_fragments);
    int rc = NPGE_ANCHOR_OK;
    try {
        r->frag_seq.resize(nf);
        r->frag_min_pos.resize(nf);
        r->frag_ori.resize(nf);
        r->block_offsets.resize(nb + 1);
        r->block_hash.resize(nb);
    } catch (std::bad_alloc&) {
        delete r;
        return fail(NPGE_ANCHOR_ERR_MEMORY, "out of host memory for the result");
    }
    const uint64_t* used_ptr = nullptr;
    uint64_t used_n = 0;
    auto body = [&]() -> int {
        cudaStream_t st = ctx->stream;
        if (fetch_compact_wanted(ctx, nf, limit)) {
            NPGE_TRY(fetch_compact(ctx, r, nf, nb));
        } else {
            if (nf) {
                NPGE_CUDA(cudaMemcpyAsync(r->frag_seq.data(), ctx->r_seq.ptr, nf * 4,
                                          cudaMemcpyDeviceToHost, st));
                NPGE_CUDA(cudaMemcpyAsync(r->frag_min_pos.data(), ctx->r_pos.ptr, nf * 4,
                                          cudaMemcpyDeviceToHost, st));
                NPGE_CUDA(cudaMemcpyAsync(r->frag_ori.data(), ctx->r_ori.ptr, nf,
                                          cudaMemcpyDeviceToHost, st));
            }
            NPGE_CUDA(cudaMemcpyAsync(r->block_offsets.data(), ctx->r_block_off.ptr,
                                      (nb + 1) * 8, cudaMemcpyDeviceToHost, st));
            if (nb) {
                NPGE_CUDA(cudaMemcpyAsync(r->block_hash.data(), ctx->r_block_hash.ptr,
                                          nb * 8, cudaMemcpyDeviceToHost, st));
            }
            NPGE_CUDA(cudaStreamSynchronize(st));
        }
        out->digest[0] = ctx->stats.digest[0];
        out->digest[1] = ctx->stats.digest[1];
        const bool lone_tail = ctx->lone_tail;
        const uint64_t lone_hash = ctx->lone_hash;
        // used_hashes.push_back for every distinct hash of the list but the
        // first (AnchorFinder.cpp:366-387)
        // (several GPUs: every hash of this rank's range; the caller drops the
        // job's first one)
        const size_t b0 = ctx->dist ? 0 : 1;
        if (lone_tail && (ctx->dist || !r->block_hash.empty())) {
            r->new_used.assign(r->block_hash.begin() + std::min(b0, r->block_hash.size()),
                               r->block_hash.end());
            r->new_used.push_back(lone_hash);
            used_ptr = r->new_used.data();
            used_n = r->new_used.size();
        } else if (r->block_hash.size() > b0) {
            // the list IS the tail of block_hash: no copy (millions of blocks in
            // parity mode)
            used_ptr = r->block_hash.data() + b0;
            used_n = r->block_hash.size() - b0;
        }
        return NPGE_ANCHOR_OK;
    };
    rc = body();
    if (rc != NPGE_ANCHOR_OK) {
        delete r;
        return rc;
    }
    out->n_blocks = r->block_hash.size();
    out->n_fragments = r->frag_seq.size();
    out->block_offsets = r->block_offsets.data();
    out->block_hash = r->block_hash.data();
    out->frag_seq = r->frag_seq.data();
    out->frag_min_pos = r->frag_min_pos.data();
    out->frag_ori = r->frag_ori.data();
    out->n_new_used = used_n;
    out->new_used = used_ptr;
    out->owner_ = r;
    return NPGE_ANCHOR_OK;
}

// ---- the compact result: 4 bytes per fragment over PCIe ----------------------------------
namespace {

struct CompactHost {
    PinnedVector<uint64_t> block_offsets, block_hash;
    PinnedVector<uint32_t> pos_ori;
    std::vector<uint64_t> new_used, layout_start;
    std::vector<uint32_t> layout_seq;
};

}  // namespace

extern "C" int npge_anchor_set_result_layout(npge_anchor_ctx* ctx, int32_t layout) {
    if (!ctx) return fail(NPGE_ANCHOR_ERR_PARAMS, "ctx is NULL");
    if (layout != NPGE_ANCHOR_LAYOUT_FULL && layout != NPGE_ANCHOR_LAYOUT_COMPACT) {
        return fail(NPGE_ANCHOR_ERR_PARAMS, "unknown result layout");
    }
    ctx->layout = layout;
    return NPGE_ANCHOR_OK;
}

namespace {

// `r` holds the first nf fragments / nb blocks (+ closing offset): the layout table,
// new_used and the fields of `out`
void finish_compact(npge_anchor_ctx* ctx, CompactHost* r, uint64_t nf, uint64_t nb,
                    npge_anchor_compact_result* out) {
    // (a rank of a multi-GPU job: positions are job-wide, the layout is the job's)
    const auto& seqs = ctx->dist ? ctx->gseqs : ctx->seqs;
    r->layout_start.resize(seqs.size() + 1);
    r->layout_seq.resize(seqs.size() + 1);
    uint64_t end = 0;
    for (size_t i = 0; i < seqs.size(); i++) {
        r->layout_start[i] = seqs[i].start;
        r->layout_seq[i] = seqs[i].caller_index;
        end = seqs[i].start + seqs[i].size + 1;
    }
    r->layout_start[seqs.size()] = end;
    // used_hashes.push_back for every distinct hash of the list but the first
    // (AnchorFinder.cpp:366-387); a lone cut tail still counts (:385)
    // (several GPUs: every hash of this rank's range; the caller drops the job's first)
    const uint64_t b0 = ctx->dist ? 0 : 1;
    if (ctx->lone_tail && (ctx->dist || nb)) {
        r->new_used.assign(r->block_hash.data() + std::min(b0, nb), r->block_hash.data() + nb);
        r->new_used.push_back(ctx->lone_hash);
        out->new_used = r->new_used.data();
        out->n_new_used = r->new_used.size();
    } else if (nb > b0) {
        out->new_used = r->block_hash.data() + b0;  // the list IS the tail of block_hash
        out->n_new_used = nb - b0;
    }
    out->n_blocks = nb;
    out->n_fragments = nf;
    out->block_offsets = r->block_offsets.data();
    out->block_hash = r->block_hash.data();
    out->frag_pos_ori = r->pos_ori.data();
    out->n_layout = uint32_t(seqs.size());
    out->layout_start = r->layout_start.data();
    out->layout_seq = r->layout_seq.data();
    out->anchor_size = ctx->last_params.anchor_size;
    out->digest[0] = ctx->stats.digest[0];
    out->digest[1] = ctx->stats.digest[1];
    out->owner_ = r;
}

}  // namespace

extern "C" int npge_anchor_fetch_compact(npge_anchor_ctx* ctx, npge_anchor_compact_result* out) {
    if (!ctx || !out) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    memset(out, 0, sizeof(*out));
    if (!ctx->have_result) {
        return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_find_device has not run");
    }
    if (!ctx->have_pay) {
        return fail(NPGE_ANCHOR_ERR_STATE,
                    "the resident result holds no payloads: npge_anchor_set_result_layout("
                    "NPGE_ANCHOR_LAYOUT_COMPACT) before npge_anchor_find_device");
    }
    NPGE_CUDA(cudaSetDevice(ctx->device));
    CompactHost* r = new (std::nothrow) CompactHost;
    if (!r) return fail(NPGE_ANCHOR_ERR_MEMORY, "out of host memory");
    const uint64_t nf = ctx->stats.n_fragments, nb = ctx->stats.n_blocks;
    try {
        r->pos_ori.resize(nf);
        r->block_offsets.resize(nb + 1);
        r->block_hash.resize(nb);
    } catch (std::bad_alloc&) {
        delete r;
        return fail(NPGE_ANCHOR_ERR_MEMORY, "out of host memory for the result");
    }
    auto body = [&]() -> int {
        cudaStream_t st = ctx->stream;
        if (nf) {
            NPGE_CUDA(cudaMemcpyAsync(r->pos_ori.data(), ctx->r_pay.ptr, nf * 4,
                                      cudaMemcpyDeviceToHost, st));
        }
        NPGE_CUDA(cudaMemcpyAsync(r->block_offsets.data(), ctx->r_block_off.ptr, (nb + 1) * 8,
                                  cudaMemcpyDeviceToHost, st));
        if (nb) {
            NPGE_CUDA(cudaMemcpyAsync(r->block_hash.data(), ctx->r_block_hash.ptr, nb * 8,
                                      cudaMemcpyDeviceToHost, st));
        }
        NPGE_CUDA(cudaStreamSynchronize(st));
        return NPGE_ANCHOR_OK;
    };
    const int rc = body();
    if (rc != NPGE_ANCHOR_OK) {
        delete r;
        return rc;
    }
    finish_compact(ctx, r, nf, nb, out);
    return NPGE_ANCHOR_OK;
}

namespace npge_b200 {

void* pinned_take(size_t bytes) {
    return PinnedPool::instance().take(PinnedPool::round_up(bytes));
}

void pinned_give(void* p, size_t bytes) {
    if (p) PinnedPool::instance().give(p, PinnedPool::round_up(bytes));
}

int fetch_compact_into(npge_anchor_ctx* ctx, uint32_t* pos_ori, uint64_t* block_off,
                       uint64_t* block_hash) {
    if (!ctx || !ctx->have_result || !ctx->have_pay) {
        return fail(NPGE_ANCHOR_ERR_STATE, "no compact resident result");
    }
    NPGE_CUDA(cudaSetDevice(ctx->device));
    const uint64_t nf = ctx->stats.n_fragments, nb = ctx->stats.n_blocks;
    cudaStream_t st = ctx->stream;
    if (nf) {
        NPGE_CUDA(cudaMemcpyAsync(pos_ori, ctx->r_pay.ptr, nf * 4, cudaMemcpyDeviceToHost, st));
    }
    if (nb) {
        NPGE_CUDA(cudaMemcpyAsync(block_off, ctx->r_block_off.ptr, nb * 8,
                                  cudaMemcpyDeviceToHost, st));
        NPGE_CUDA(cudaMemcpyAsync(block_hash, ctx->r_block_hash.ptr, nb * 8,
                                  cudaMemcpyDeviceToHost, st));
    }
    NPGE_CUDA(cudaStreamSynchronize(st));
    return NPGE_ANCHOR_OK;
}

}  // namespace npge_b200

extern "C" void npge_anchor_compact_result_free(npge_anchor_compact_result* out) {
    if (!out) return;
    delete static_cast<CompactHost*>(out->owner_);
    memset(out, 0, sizeof(*out));
}

extern "C" int npge_anchor_find_compact(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                                        const uint8_t* bases, const uint64_t* seq_offsets,
                                        uint32_t n_seqs, const uint64_t* used_sorted,
                                        uint64_t n_used, npge_anchor_compact_result* out) {
    if (!out) return fail(NPGE_ANCHOR_ERR_PARAMS, "out is NULL");
    memset(out, 0, sizeof(*out));
    if (!ctx) return fail(NPGE_ANCHOR_ERR_PARAMS, "ctx is NULL");
    NPGE_TRY(check_params(p));
    NPGE_TRY(npge_anchor_upload(ctx, p->anchor_size, bases, seq_offsets, n_seqs));
    // Exact mode without a cut (the case with a result of the size of the input): host
    // buffers for the largest possible result, filled range by range while the device is
    // still sorting (find_device_msd).  Anything else: find, then fetch.
    CompactHost* r = nullptr;
    npge_anchor_ctx::FetchPipe pipe = {nullptr, nullptr, nullptr, false};
    const uint64_t max_records = ctx->n_positions;
    if (!p->anchor_similar && uint64_t(p->max_anchor_fragments) >= max_records &&
        max_records > 0 && max_records <= kPipeMaxRecords) {
        r = new (std::nothrow) CompactHost;
        try {
            if (r) {
                r->pos_ori.resize(max_records);
                r->block_offsets.resize(max_records / 2 + 2);
                r->block_hash.resize(max_records / 2 + 2);
                pipe.pos_ori = r->pos_ori.data();
                pipe.block_off = r->block_offsets.data();
                pipe.block_hash = r->block_hash.data();
            }
        } catch (std::bad_alloc&) {  // not enough pinned memory: the plain way
            delete r;
            r = nullptr;
        }
    }
    const int before = ctx->layout;
    ctx->layout = NPGE_ANCHOR_LAYOUT_COMPACT;
    ctx->pipe = r ? &pipe : nullptr;
    int rc = npge_anchor_find_device(ctx, p, used_sorted, n_used, nullptr);
    ctx->pipe = nullptr;
    ctx->layout = before;
    if (rc == NPGE_ANCHOR_OK && r && pipe.used) {
        cudaError_t e = cudaStreamSynchronize(ctx->fetch_stream);  // the last ranges
        if (e != cudaSuccess) {
            delete r;
            return fail(NPGE_ANCHOR_ERR_CUDA, std::string("CUDA error: ") + cudaGetErrorString(e));
        }
        const uint64_t nf = ctx->stats.n_fragments, nb = ctx->stats.n_blocks;
        r->pos_ori.resize(nf);
        r->block_offsets.resize(nb + 1);
        r->block_hash.resize(nb);
        finish_compact(ctx, r, nf, nb, out);
        return NPGE_ANCHOR_OK;
    }
    if (r && ctx->fetch_stream) cudaStreamSynchronize(ctx->fetch_stream);
    delete r;
    if (rc != NPGE_ANCHOR_OK) return rc;
    return npge_anchor_fetch_compact(ctx, out);
}

extern "C" void npge_anchor_result_free(npge_anchor_result* out) {
    if (!out) return;
    delete static_cast<HostResult*>(out->owner_);
    memset(out, 0, sizeof(*out));
}

extern "C" int npge_anchor_find(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                                const uint8_t* bases, const uint64_t* seq_offsets,
                                uint32_t n_seqs, const uint64_t* used_sorted,
                                uint64_t n_used, npge_anchor_result* out) {
    if (!out) return fail(NPGE_ANCHOR_ERR_PARAMS, "out is NULL");
    memset(out, 0, sizeof(*out));
    NPGE_TRY(check_params(p));
    NPGE_TRY(npge_anchor_upload(ctx, p->anchor_size, bases, seq_offsets, n_seqs));
    NPGE_TRY(npge_anchor_find_device(ctx, p, used_sorted, n_used, nullptr));
    return npge_anchor_fetch(ctx, out);
}

// ==== several GPUs: the per-rank stages between the caller's collectives ==========
namespace {

MsdBuffers dist_msd_buffers(npge_anchor_ctx* ctx, const MsdPlan& plan, uint32_t max_seg) {
    MsdBuffers mb = {};
    mb.rows = ptr<uint32_t>(ctx->rows);
    mb.blk = ptr<uint32_t>(ctx->blk);
    mb.seg = carve_seg_tables(ptr<uint32_t>(ctx->segtab), max_seg, plan);
    mb.fine_start = ptr<uint32_t>(ctx->fine_start);
    mb.tile_desc = ptr<uint4>(ctx->tile_desc);
    for (int b = 0; b < 2; b++) {
        mb.rec[b] = ptr<uint64_t>(ctx->rec[b]);
        mb.hi[b] = ptr<uint32_t>(ctx->rhi[b]);
    }
    return mb;
}

}  // namespace

// Position-range upload of one rank: the job's sequences, in scan order
// with one pad after each, form ONE position space [0, P); this rank generates the
// windows that start in [g0, g0 + n_own) and holds the bases of [g0, g0 + n_have)
// (n_have - n_own <= k - 1: the halo completing its last windows).  A sequence may
// be cut anywhere: GPUs are balanced by positions, not by sequences.
extern "C" int npge_anchor_dist_upload_range(npge_anchor_ctx* ctx, int32_t anchor_size,
                                             const uint8_t* bases, uint32_t n_pieces,
                                             const uint64_t* src_off, const uint64_t* dst_pos,
                                             const uint64_t* len, uint64_t g0, uint64_t n_own,
                                             uint64_t n_have, const uint64_t* global_start,
                                             const uint32_t* global_caller,
                                             uint32_t n_global) {
    if (!ctx) return fail(NPGE_ANCHOR_ERR_PARAMS, "ctx is NULL");
    npge_anchor_params probe = {anchor_size, 0, 0};
    NPGE_TRY(check_params(&probe));
    if (!global_start || (n_global && !global_caller) ||
        (n_pieces && (!bases || !src_off || !dst_pos || !len))) {
        return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL table");
    }
    const uint64_t g_positions = global_start[n_global];
    if (g_positions > kMaxPositions) {
        return fail(NPGE_ANCHOR_ERR_CAPACITY, "more than 2^31-1 positions in the job");
    }
    if (n_own > n_have || g0 + n_have > g_positions ||
        n_have - n_own >= uint64_t(anchor_size)) {
        return fail(NPGE_ANCHOR_ERR_PARAMS, "bad position range");
    }
    for (uint32_t i = 0; i < n_pieces; i++) {
        if (dst_pos[i] + len[i] > n_have) {
            return fail(NPGE_ANCHOR_ERR_PARAMS, "piece outside the rank's range");
        }
    }
    NPGE_CUDA(cudaSetDevice(ctx->device));
    ctx->uploaded = false;
    ctx->have_result = false;
    ctx->seqs.clear();
    ctx->gseqs.clear();
    uint64_t own_bases = 0;
    for (uint32_t i = 0; i < n_global; i++) {
        if (global_start[i + 1] <= global_start[i]) {
            return fail(NPGE_ANCHOR_ERR_PARAMS, "global_start must be increasing");
        }
        const uint64_t b = global_start[i], e = global_start[i + 1] - 1;  // bases of sequence i
        ctx->gseqs.push_back({global_caller[i], e - b, b});
        const uint64_t lo = std::max(b, g0), hi = std::min(e, g0 + n_own);
        if (hi > lo) own_bases += hi - lo;
    }
    ctx->n_positions = n_have;
    ctx->n_bases = own_bases;
    ctx->n_padded = (n_have + kPosAlign - 1) / kPosAlign * kPosAlign;
    ctx->n_caller_seqs = n_global;
    ctx->upload_k = anchor_size;
    ctx->g_positions = g_positions;
    ctx->g_padded = (g_positions + kPosAlign - 1) / kPosAlign * kPosAlign;
    ctx->dist_g0 = g0;
    ctx->dist_n_own = n_own;
    const uint64_t np = ctx->n_padded;
    const size_t slack = 256;
    NPGE_TRY(ensure(ctx->ascii, np + slack));
    NPGE_TRY(ensure(ctx->packed, np / 4 + slack));
    NPGE_TRY(ensure(ctx->nmask, np / 8 + slack));
    NPGE_TRY(ensure(ctx->counters, kCntWords * 4));
    cudaStream_t st = ctx->stream;
    NPGE_CUDA(cudaMemsetAsync(ctx->counters.ptr, 0, kCntWords * 4, st));
    NPGE_CUDA(cudaMemsetAsync(ctx->ascii.ptr, 'N', np + slack, st));
    NPGE_CUDA(cudaMemsetAsync(ptr<uint8_t>(ctx->packed) + np / 4, 0, slack, st));
    NPGE_CUDA(cudaMemsetAsync(ptr<uint8_t>(ctx->nmask) + np / 8, 0xff, slack, st));
    for (uint32_t i = 0; i < n_pieces; i++) {
        if (len[i] == 0) continue;
        // (cudaMemcpyDefault: `bases` may also be device memory, e.g. the pieces a
        // multi-process job has just exchanged)
        NPGE_CUDA(cudaMemcpyAsync(ptr<uint8_t>(ctx->ascii) + dst_pos[i], bases + src_off[i],
                                  len[i], cudaMemcpyDefault, st));
    }
    // the job-wide tables (the same on every rank) for the emit side
    const uint64_t n_chunks = (ctx->g_padded >> ctx->lut_shift) + 2;
    std::vector<uint32_t> lut(n_chunks);
    {
        size_t sidx = 0;
        for (uint64_t q = 0; q < n_chunks; q++) {
            const uint64_t g = q << ctx->lut_shift;
            while (sidx + 1 < n_global && global_start[sidx + 1] <= g) sidx++;
            lut[q] = uint32_t(sidx);
        }
    }
    std::vector<uint32_t> callers(size_t(n_global) + 1, 0);
    for (uint32_t i = 0; i < n_global; i++) callers[i] = global_caller[i];
    std::vector<uint2> pairs(size_t(n_global) + 1);
    for (uint32_t i = 0; i <= n_global; i++) {
        pairs[i] = make_uint2(uint32_t(global_start[i]), callers[i]);
    }
    // host copies: npge_anchor_fetch expands a compact resident result with them
    ctx->h_lut = lut;
    ctx->h_caller = callers;
    ctx->h_start32.resize(size_t(n_global) + 1);
    for (uint32_t i = 0; i <= n_global; i++) ctx->h_start32[i] = uint32_t(global_start[i]);
    NPGE_TRY(ensure(ctx->g_seq_start, (size_t(n_global) + 1) * 8));
    NPGE_TRY(ensure(ctx->g_seq_caller, (size_t(n_global) + 1) * 4));
    NPGE_TRY(ensure(ctx->g_seq_lut, n_chunks * 4));
    NPGE_TRY(ensure(ctx->g_seq_pair, pairs.size() * 8));
    NPGE_CUDA(cudaMemcpyAsync(ctx->g_seq_start.ptr, global_start, (size_t(n_global) + 1) * 8,
                              cudaMemcpyHostToDevice, st));
    NPGE_CUDA(cudaMemcpyAsync(ctx->g_seq_caller.ptr, callers.data(), callers.size() * 4,
                              cudaMemcpyHostToDevice, st));
    NPGE_CUDA(cudaMemcpyAsync(ctx->g_seq_lut.ptr, lut.data(), n_chunks * 4,
                              cudaMemcpyHostToDevice, st));
    NPGE_CUDA(cudaMemcpyAsync(ctx->g_seq_pair.ptr, pairs.data(), pairs.size() * 8,
                              cudaMemcpyHostToDevice, st));
    NPGE_CUDA(cudaStreamSynchronize(st));  // the tables are locals
    if (np) {
        launch_pack(ptr<uint8_t>(ctx->ascii), ptr<uint32_t>(ctx->packed),
                    ptr<uint32_t>(ctx->nmask), np, ptr<uint32_t>(ctx->counters), st);
        NPGE_CUDA(cudaGetLastError());
        uint32_t bad = 0;
        NPGE_CUDA(cudaMemcpyAsync(&bad, ptr<uint32_t>(ctx->counters) + kCntBadInput, 4,
                                  cudaMemcpyDeviceToHost, st));
        NPGE_CUDA(cudaStreamSynchronize(st));
        if (bad) {
            return fail(NPGE_ANCHOR_ERR_INPUT, std::to_string(bad) + " bytes outside ATGCN: "
                        "run Sequence::to_atgcn on the input first");
        }
    }
    ctx->dist = true;
    ctx->uploaded = true;
    return NPGE_ANCHOR_OK;
}

namespace {

PosMap dist_pos_map(npge_anchor_ctx* ctx) {
    PosMap pm;
    pm.g0 = uint32_t(ctx->dist_g0);
    pm.win_limit = uint32_t(ctx->dist_n_own);
    return pm;
}

// count A (+ the local scatter when `scatter`); fills hist_out with the records
// per A bucket
int dist_count_a(npge_anchor_ctx* ctx, const npge_anchor_params* p, bool scatter,
                 uint32_t* hist_out, uint32_t hist_cap, uint32_t* n_bins) {
    if (!ctx || !hist_out || !n_bins) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    NPGE_TRY(check_params(p));
    if (!ctx->uploaded || !ctx->dist) {
        return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_dist_upload_range has not been called");
    }
    if (p->anchor_size != ctx->upload_k) {
        return fail(NPGE_ANCHOR_ERR_STATE, "anchor-size differs from the uploaded one");
    }
    NPGE_CUDA(cudaSetDevice(ctx->device));
    ctx->have_result = false;
    ctx->a_counted = false;
    const MsdPlan plan = make_msd_plan(p->anchor_size, ctx->g_positions, ctx->force_f,
                                       ctx->force_wa);
    const uint32_t bins_a = 1u << plan.capa;
    *n_bins = bins_a;
    const uint64_t np = ctx->n_padded, max_records = ctx->n_positions;
    const int chunks = scatter ? 1 : ctx->dist_chunks;  // (dist_pass_a: one chunk)
    if (hist_cap < bins_a * uint32_t(chunks)) return fail(NPGE_ANCHOR_ERR_PARAMS, "hist_out too small");
    NPGE_TRY(ensure(ctx->rows, msd_rows_needed(plan, np, max_records, bins_a) * 4));
    NPGE_TRY(ensure(ctx->blk, (msd_blk_needed(plan, np, max_records, bins_a) +
                               size_t(kMaxChunks + 1) * bins_a) * 4));
    NPGE_TRY(ensure(ctx->segtab, seg_table_words(bins_a, plan) * 4));
    NPGE_TRY(ensure(ctx->chunk_tab, size_t(chunks) * (2 * bins_a + 1) * 4));
    if (scatter) {
        NPGE_TRY(ensure(ctx->rec[0], (max_records + 16) * 8));
        if (plan.wide) NPGE_TRY(ensure(ctx->rhi[0], (max_records + 16) * 4));
    }
    NPGE_TRY(ensure(ctx->counters, kCntWords * 4));
    cudaStream_t st = ctx->stream;
    memset(ctx->dist_ms, 0, sizeof(ctx->dist_ms));
    ctx->dist_launches = 0;
    NPGE_CUDA(cudaMemsetAsync(ctx->counters.ptr, 0, kCntWords * 4, st));
    MsdBuffers mb = dist_msd_buffers(ctx, plan, bins_a);
    // a_start[c][bins + 1]: first record of every (chunk, bucket) of the pass-A output
    std::vector<uint32_t> a_start(size_t(chunks) * (bins_a + 1), 0);
    PassAChunks ch = {chunks, ptr<uint32_t>(ctx->chunk_tab),
                      ptr<uint32_t>(ctx->chunk_tab) + size_t(chunks) * bins_a, nullptr};
    if (np) {
        NPGE_CUDA(cudaEventRecord(ctx->ev[0], st));
        launch_msd_pass_a(plan, ptr<uint32_t>(ctx->packed), ptr<uint32_t>(ctx->nmask), np,
                          mb, dist_pos_map(ctx), ptr<uint32_t>(ctx->counters), st,
                          ctx->ev_pass, &ctx->dist_launches, scatter ? 0 : 1,
                          scatter ? nullptr : &ch);
        NPGE_CUDA(cudaGetLastError());
        NPGE_CUDA(cudaMemcpyAsync(a_start.data(), scatter ? mb.seg.a_start : ch.chunk_start,
                                  a_start.size() * 4, cudaMemcpyDeviceToHost, st));
        NPGE_CUDA(cudaStreamSynchronize(st));
        cudaEventElapsedTime(&ctx->dist_ms[0], ctx->ev[0], ctx->ev_pass[0]);
        if (scatter) cudaEventElapsedTime(&ctx->dist_ms[1], ctx->ev_pass[0], ctx->ev_pass[1]);
    }
    for (int c = 0; c < chunks; c++) {
        for (uint32_t a = 0; a < bins_a; a++) {
            const size_t i = size_t(c) * (bins_a + 1) + a;
            hist_out[size_t(c) * bins_a + a] = a_start[i + 1] - a_start[i];
        }
    }
    ctx->a_chunks = chunks;
    ctx->plan = plan;
    ctx->a_counted = true;
    ctx->a_scattered = scatter;
    ctx->a_start_h = a_start;
    return NPGE_ANCHOR_OK;
}

}  // namespace

extern "C" int npge_anchor_dist_pass_a(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                                       uint32_t* hist_out, uint32_t hist_cap,
                                       uint32_t* n_bins) {
    return dist_count_a(ctx, p, true, hist_out, hist_cap, n_bins);
}

extern "C" int npge_anchor_dist_count_a(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                                        uint32_t* hist_out, uint32_t hist_cap,
                                        uint32_t* n_bins) {
    return dist_count_a(ctx, p, false, hist_out, hist_cap, n_bins);
}

// ---- peer exchange: receive buffers shared through CUDA IPC -------------------------
extern "C" int npge_anchor_dist_peer_close(npge_anchor_ctx* ctx) {
    if (!ctx) return fail(NPGE_ANCHOR_ERR_PARAMS, "ctx is NULL");
    NPGE_CUDA(cudaSetDevice(ctx->device));
    NPGE_CUDA(cudaStreamSynchronize(ctx->stream));
    for (void* q : ctx->peer_opened) cudaIpcCloseMemHandle(q);
    ctx->peer_opened.clear();
    ctx->peer_rec.clear();
    ctx->peer_hi.clear();
    return NPGE_ANCHOR_OK;
}

extern "C" int npge_anchor_dist_peer_alloc(npge_anchor_ctx* ctx, uint64_t capacity,
                                           uint8_t* handles_out) {
    if (!ctx || !handles_out) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    static_assert(2 * sizeof(cudaIpcMemHandle_t) <= NPGE_ANCHOR_IPC_BYTES, "handle size");
    if (capacity > kMaxRecords) {
        return fail(NPGE_ANCHOR_ERR_CAPACITY, "more than 2^30-1 records on one rank");
    }
    if (!ctx->peer_opened.empty()) {
        return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_dist_peer_close first");
    }
    NPGE_CUDA(cudaSetDevice(ctx->device));
    NPGE_TRY(ensure(ctx->recv_rec, (capacity + 16) * 8));
    NPGE_TRY(ensure(ctx->recv_hi, (capacity + 16) * 4));
    ctx->peer_capacity = capacity;
    cudaIpcMemHandle_t h[2];
    NPGE_CUDA(cudaIpcGetMemHandle(&h[0], ctx->recv_rec.ptr));
    NPGE_CUDA(cudaIpcGetMemHandle(&h[1], ctx->recv_hi.ptr));
    memset(handles_out, 0, NPGE_ANCHOR_IPC_BYTES);
    memcpy(handles_out, h, sizeof(h));
    return NPGE_ANCHOR_OK;
}

extern "C" int npge_anchor_dist_peer_open(npge_anchor_ctx* ctx, uint32_t world,
                                          uint32_t rank, const uint8_t* handles_all) {
    if (!ctx || !handles_all) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    if (rank >= world) return fail(NPGE_ANCHOR_ERR_PARAMS, "rank >= world");
    if (!ctx->recv_rec.ptr || !ctx->peer_capacity) {
        return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_dist_peer_alloc first");
    }
    NPGE_TRY(npge_anchor_dist_peer_close(ctx));
    ctx->peer_rec.assign(world, nullptr);
    ctx->peer_hi.assign(world, nullptr);
    ctx->dist_rank = rank;
    for (uint32_t r = 0; r < world; r++) {
        if (r == rank) {
            ctx->peer_rec[r] = ctx->recv_rec.ptr;
            ctx->peer_hi[r] = ctx->recv_hi.ptr;
            continue;
        }
        cudaIpcMemHandle_t h[2];
        memcpy(h, handles_all + size_t(r) * NPGE_ANCHOR_IPC_BYTES, sizeof(h));
        void* q[2] = {nullptr, nullptr};
        for (int i = 0; i < 2; i++) {
            cudaError_t e = cudaIpcOpenMemHandle(&q[i], h[i], cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) {
                cudaGetLastError();
                npge_anchor_dist_peer_close(ctx);
                return fail(NPGE_ANCHOR_ERR_CUDA,
                            std::string("cudaIpcOpenMemHandle (no peer access between "
                                        "the ranks' GPUs?): ") + cudaGetErrorString(e));
            }
            ctx->peer_opened.push_back(q[i]);
        }
        ctx->peer_rec[r] = q[0];
        ctx->peer_hi[r] = q[1];
    }
    return NPGE_ANCHOR_OK;
}

// Contexts of ONE process (npge_anchor_multi_*): the receive buffers of the other
// ranks are plain device pointers; peer access between the devices is enabled here.
extern "C" int npge_anchor_dist_peer_attach(npge_anchor_ctx* ctx, uint32_t world, uint32_t rank,
                                            void* const* rec, void* const* hi,
                                            const int* devices) {
    if (!ctx || !rec || !hi || !devices) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    if (rank >= world) return fail(NPGE_ANCHOR_ERR_PARAMS, "rank >= world");
    if (!ctx->recv_rec.ptr || !ctx->peer_capacity) {
        return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_dist_peer_alloc first");
    }
    NPGE_TRY(npge_anchor_dist_peer_close(ctx));
    NPGE_CUDA(cudaSetDevice(ctx->device));
    ctx->peer_rec.assign(world, nullptr);
    ctx->peer_hi.assign(world, nullptr);
    ctx->dist_rank = rank;
    for (uint32_t r = 0; r < world; r++) {
        if (devices[r] != ctx->device) {
            int can = 0;
            NPGE_CUDA(cudaDeviceCanAccessPeer(&can, ctx->device, devices[r]));
            if (!can) return fail(NPGE_ANCHOR_ERR_CUDA, "no peer access between the devices");
            cudaError_t e = cudaDeviceEnablePeerAccess(devices[r], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                return fail(NPGE_ANCHOR_ERR_CUDA, std::string("cudaDeviceEnablePeerAccess: ") +
                            cudaGetErrorString(e));
            }
            cudaGetLastError();
        }
        ctx->peer_rec[r] = r == rank ? ctx->recv_rec.ptr : rec[r];
        ctx->peer_hi[r] = r == rank ? ctx->recv_hi.ptr : hi[r];
    }
    return NPGE_ANCHOR_OK;
}

// The found bitmap of an anchor-similar job inside one process: this rank's bitmap
// becomes the OR of all ranks' (found[r]: their current bitmaps, peer pointers).  The
// result is built in a second buffer; call npge_anchor_dist_found_swap on every rank
// after all of them have finished reading.
extern "C" int npge_anchor_dist_found_or(npge_anchor_ctx* ctx, uint32_t world,
                                         const void* const* found) {
    if (!ctx || !found || world == 0 || world > 8) return fail(NPGE_ANCHOR_ERR_PARAMS, "bad arguments");
    NPGE_CUDA(cudaSetDevice(ctx->device));
    const size_t bytes = (ctx->g_padded / 8 + 256) / 16 * 16;
    NPGE_TRY(ensure(ctx->found_tmp, ctx->g_padded / 8 + 256));
    const uint32_t* srcs[8];
    for (uint32_t r = 0; r < world; r++) srcs[r] = static_cast<const uint32_t*>(found[r]);
    launch_found_or(srcs, int(world), ptr<uint32_t>(ctx->found_tmp), bytes, ctx->stream,
                    &ctx->dist_launches);
    NPGE_CUDA(cudaGetLastError());
    NPGE_CUDA(cudaStreamSynchronize(ctx->stream));
    return NPGE_ANCHOR_OK;
}

extern "C" int npge_anchor_dist_found_swap(npge_anchor_ctx* ctx) {
    if (!ctx) return fail(NPGE_ANCHOR_ERR_PARAMS, "ctx is NULL");
    std::swap(ctx->found_bits, ctx->found_tmp);
    return NPGE_ANCHOR_OK;
}

extern "C" int npge_anchor_dist_scatter_peer(npge_anchor_ctx* ctx,
                                             const npge_anchor_params* p,
                                             const uint32_t* dest_rank,
                                             const int64_t* dest_off, uint32_t n_bins) {
    if (!ctx || !dest_rank || !dest_off) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    NPGE_TRY(check_params(p));
    if (!ctx->dist || !ctx->a_counted) {
        return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_dist_count_a has not run");
    }
    const MsdPlan plan = ctx->plan;
    const uint32_t bins_a = 1u << plan.capa;
    if (n_bins != bins_a) return fail(NPGE_ANCHOR_ERR_PARAMS, "n_bins differs from the plan");
    if (ctx->peer_rec.empty()) {
        return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_dist_peer_open has not been called");
    }
    NPGE_CUDA(cudaSetDevice(ctx->device));
    std::vector<unsigned long long> tab(2 * size_t(bins_a));
    for (uint32_t a = 0; a < bins_a; a++) {
        if (dest_rank[a] >= ctx->peer_rec.size()) {
            return fail(NPGE_ANCHOR_ERR_PARAMS, "dest_rank out of range");
        }
        // addresses may lie before the buffer; only in-range elements are written
        tab[a] = (unsigned long long)(uintptr_t(ctx->peer_rec[dest_rank[a]]) +
                                      dest_off[a] * 8);
        tab[bins_a + a] = (unsigned long long)(uintptr_t(ctx->peer_hi[dest_rank[a]]) +
                                               dest_off[a] * 4);
    }
    cudaStream_t st = ctx->stream;
    NPGE_TRY(ensure(ctx->peer_tab, tab.size() * 8));
    NPGE_CUDA(cudaMemcpyAsync(ctx->peer_tab.ptr, tab.data(), tab.size() * 8,
                              cudaMemcpyHostToDevice, st));
    const uint64_t np = ctx->n_padded;
    if (np) {
        MsdBuffers mb = dist_msd_buffers(ctx, plan, bins_a);
        PosMap pm = dist_pos_map(ctx);
        pm.peer_rec = ptr<unsigned long long>(ctx->peer_tab);
        pm.peer_hi = pm.peer_rec + bins_a;
        NPGE_CUDA(cudaEventRecord(ctx->ev[0], st));
        launch_msd_pass_a(plan, ptr<uint32_t>(ctx->packed), ptr<uint32_t>(ctx->nmask), np,
                          mb, pm, ptr<uint32_t>(ctx->counters), st, nullptr,
                          &ctx->dist_launches, 2);
        NPGE_CUDA(cudaGetLastError());
        NPGE_CUDA(cudaEventRecord(ctx->ev[1], st));
    }
    NPGE_CUDA(cudaStreamSynchronize(st));  // the peers may read after their barrier
    if (np) cudaEventElapsedTime(&ctx->dist_ms[1], ctx->ev[0], ctx->ev[1]);
    ctx->a_counted = false;
    return NPGE_ANCHOR_OK;
}

// pass A into this rank's own memory, asynchronously: the caller overlaps the
// histogram all-gather with it; npge_anchor_dist_exchange follows
extern "C" int npge_anchor_dist_scatter_local(npge_anchor_ctx* ctx,
                                              const npge_anchor_params* p) {
    if (!ctx) return fail(NPGE_ANCHOR_ERR_PARAMS, "ctx is NULL");
    NPGE_TRY(check_params(p));
    if (!ctx->dist || !ctx->a_counted) {
        return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_dist_count_a has not run");
    }
    NPGE_CUDA(cudaSetDevice(ctx->device));
    const MsdPlan plan = ctx->plan;
    const uint32_t bins_a = 1u << plan.capa;
    const uint64_t np = ctx->n_padded, max_records = ctx->n_positions;
    NPGE_TRY(ensure(ctx->rec[0], (max_records + 16) * 8));
    if (plan.wide) NPGE_TRY(ensure(ctx->rhi[0], (max_records + 16) * 4));
    cudaStream_t st = ctx->stream;
    while (ctx->chunk_events.size() < size_t(kMaxChunks)) {
        cudaEvent_t ev = nullptr;
        NPGE_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        ctx->chunk_events.push_back(ev);
    }
    NPGE_CUDA(cudaEventRecord(ctx->ev_pass[0], st));
    if (np) {
        MsdBuffers mb = dist_msd_buffers(ctx, plan, bins_a);
        const int chunks = ctx->a_chunks;
        PassAChunks ch = {chunks, ptr<uint32_t>(ctx->chunk_tab),
                          ptr<uint32_t>(ctx->chunk_tab) + size_t(chunks) * bins_a,
                          ctx->chunk_events.data()};
        launch_msd_pass_a(plan, ptr<uint32_t>(ctx->packed), ptr<uint32_t>(ctx->nmask), np,
                          mb, dist_pos_map(ctx), ptr<uint32_t>(ctx->counters), st, nullptr,
                          &ctx->dist_launches, 2, &ch);
        NPGE_CUDA(cudaGetLastError());
    } else {
        for (int c = 0; c < ctx->a_chunks; c++) NPGE_CUDA(cudaEventRecord(ctx->chunk_events[c], st));
    }
    NPGE_CUDA(cudaEventRecord(ctx->ev_pass[1], st));
    ctx->a_scattered = true;
    ctx->a_scatter_async = true;
    return NPGE_ANCHOR_OK;
}

// Pass A of the following npge_anchor_dist_count_a / scatter_local / exchange runs in
// `chunks` chunks of tiles (1..8), the copies of a finished chunk overlapping the
// ranking of the next.  The same on every rank of a job.
extern "C" int npge_anchor_dist_set_chunks(npge_anchor_ctx* ctx, int32_t chunks) {
    if (!ctx) return fail(NPGE_ANCHOR_ERR_PARAMS, "ctx is NULL");
    if (chunks < 1 || chunks > kMaxChunks) return fail(NPGE_ANCHOR_ERR_PARAMS, "chunks must be 1..8");
    ctx->dist_chunks = chunks;
    return NPGE_ANCHOR_OK;
}

// ---- the exchange, planned and driven from here ---------------------------------------
// hist_all = the all-gathered per-bucket record counts of every rank ([world][bins]).
// Builds the job's tables (dist_plan.h, the same on every rank), then sends this rank's
// slice of every destination into the owner's receive buffer: one asynchronous
// device-to-device copy per destination, each on its own stream, so that all NVSwitch
// flows run at once and on the copy engines (the SMs stay free for pass A of a later
// call or of another context).  The copies wait for pass A (npge_anchor_dist_scatter_local,
// same context) through an event.  n_recv_out[world]: records every rank receives.
// NPGE_ANCHOR_ERR_CAPACITY: a receive buffer is too small (n_recv_out is valid: regrow
// with npge_anchor_dist_peer_alloc / open and call again).
extern "C" int npge_anchor_dist_exchange(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                                         const uint32_t* hist_all, uint32_t world,
                                         uint32_t rank, uint64_t* n_recv_out) {
    if (!ctx || !hist_all) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    NPGE_TRY(check_params(p));
    if (!ctx->dist || !ctx->a_counted || !ctx->a_scattered) {
        return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_dist_scatter_local has not run");
    }
    if (rank >= world || ctx->peer_rec.size() != world || ctx->dist_rank != rank) {
        return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_dist_peer_open does not match world/rank");
    }
    const MsdPlan plan = ctx->plan;
    const uint32_t bins_a = 1u << plan.capa;
    NPGE_CUDA(cudaSetDevice(ctx->device));
    DistPlan& dp = ctx->dplan;
    const uint32_t chunks = uint32_t(ctx->a_chunks);
    // hist_all: [world][chunks][bins] = [world * chunks sources][bins]
    dist_make_plan(hist_all, world, chunks, bins_a, rank, ctx->a_start_h.data(), dp);
    ctx->planned = false;
    uint64_t max_recv = 0;
    for (uint32_t d = 0; d < world; d++) {
        if (n_recv_out) n_recv_out[d] = dp.n_recv[d];
        max_recv = std::max(max_recv, dp.n_recv[d]);
    }
    for (uint32_t c = 0; c < chunks; c++) {
        for (uint32_t a = 0; a < bins_a; a++) {
            const size_t i = size_t(c) * (bins_a + 1) + a;
            if (hist_all[(size_t(rank) * chunks + c) * bins_a + a] !=
                ctx->a_start_h[i + 1] - ctx->a_start_h[i]) {
                return fail(NPGE_ANCHOR_ERR_PARAMS,
                            "hist_all[rank] differs from this rank's count A");
            }
        }
    }
    if (max_recv > ctx->peer_capacity) {
        return fail(NPGE_ANCHOR_ERR_CAPACITY, "a receive buffer is too small for " +
                    std::to_string(max_recv) + " records");
    }
    while (ctx->copy_streams.size() < world) {
        cudaStream_t cs = nullptr;
        cudaEvent_t ev = nullptr;
        NPGE_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
        NPGE_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        ctx->copy_streams.push_back(cs);
        ctx->copy_events.push_back(ev);
    }
    cudaStream_t st = ctx->stream;
    NPGE_CUDA(cudaEventRecord(ctx->ev[0], st));  // (pass A has been enqueued before)
    // chunk by chunk, as pass A finishes them; inside a chunk in ring order (rank r
    // starts with the owner after itself)
    for (uint32_t c = 0; c < chunks; c++) {
        for (uint32_t step = 0; step < world; step++) {
            const uint32_t d = (rank + 1 + step) % world;
            const size_t i = size_t(c) * world + d;
            const uint64_t first = dp.send_first[i], n = dp.send_n[i], to = dp.send_to[i];
            if (n == 0) continue;
            cudaStream_t cs = ctx->copy_streams[d];
            NPGE_CUDA(cudaStreamWaitEvent(cs, ctx->chunk_events[c], 0));
            NPGE_CUDA(cudaMemcpyAsync(static_cast<uint64_t*>(ctx->peer_rec[d]) + to,
                                      ptr<uint64_t>(ctx->rec[0]) + first, n * 8,
                                      cudaMemcpyDeviceToDevice, cs));
            if (plan.wide) {
                NPGE_CUDA(cudaMemcpyAsync(static_cast<uint32_t*>(ctx->peer_hi[d]) + to,
                                          ptr<uint32_t>(ctx->rhi[0]) + first, n * 4,
                                          cudaMemcpyDeviceToDevice, cs));
            }
        }
    }
    for (uint32_t d = 0; d < world; d++) {
        NPGE_CUDA(cudaEventRecord(ctx->copy_events[d], ctx->copy_streams[d]));
        NPGE_CUDA(cudaStreamWaitEvent(st, ctx->copy_events[d], 0));
    }
    NPGE_CUDA(cudaEventRecord(ctx->ev[1], st));
    NPGE_CUDA(cudaStreamSynchronize(st));  // every slice has landed (the peers wait at a barrier)
    // dist_ms[1]: pass A; dist_ms[7]: what the copies add after its last chunk (the
    // rest of them overlapped it)
    cudaEventElapsedTime(&ctx->dist_ms[7], ctx->ev_pass[1], ctx->ev[1]);
    if (ctx->a_scatter_async) {
        cudaEventElapsedTime(&ctx->dist_ms[1], ctx->ev_pass[0], ctx->ev_pass[1]);
        ctx->a_scatter_async = false;
    }
    cudaEventElapsedTime(&ctx->dist_exchange_ms, ctx->ev_pass[0], ctx->ev[1]);
    ctx->a_counted = false;
    ctx->planned = true;
    return NPGE_ANCHOR_OK;
}

int npge_anchor_dist_sort_impl(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                               int32_t stage, uint64_t n_recv, uint32_t n_seg,
                               const uint32_t* seg_begin, const uint32_t* seg_end,
                               const uint32_t* seg_a, const uint32_t* seg_tile,
                               const uint32_t* a_tile0, const uint32_t* a_start,
                               const uint64_t* used_sorted, uint64_t n_used,
                               uint64_t* totals);

// npge_anchor_dist_sort over the tables of the last npge_anchor_dist_exchange.
extern "C" int npge_anchor_dist_sort_planned(npge_anchor_ctx* ctx,
                                             const npge_anchor_params* p, int32_t stage,
                                             const uint64_t* used_sorted, uint64_t n_used,
                                             uint64_t* totals) {
    if (!ctx) return fail(NPGE_ANCHOR_ERR_PARAMS, "ctx is NULL");
    if (!ctx->planned) return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_dist_exchange has not run");
    const DistPlan& dp = ctx->dplan;
    if (stage == 0) {
        npge_anchor_dist_ptrs unused;
        NPGE_TRY(npge_anchor_dist_buffers(ctx, dp.n_recv[dp.rank], &unused));
    }
    return npge_anchor_dist_sort_impl(ctx, p, stage, dp.n_recv[dp.rank], dp.n_seg,
                                      dp.seg_begin.data(), dp.seg_end.data(), dp.seg_a.data(),
                                      dp.seg_tile.data(), dp.a_tile0.data(), dp.a_start.data(),
                                      used_sorted, n_used, totals);
}

// The tables alone, on the host (tests; no device needed): out arrays sized by the
// caller: cuts[world + 1], n_recv[world], dest_rank[bins], dest_off[bins],
// seg_begin / seg_end / seg_a [world * bins], seg_tile [world * bins + 1],
// a_tile0 / a_start [bins + 1].  Returns the number of segments in *n_seg.
extern "C" int npge_anchor_dist_plan_tables(const uint32_t* hist_all, uint32_t world,
                                            uint32_t chunks, uint32_t bins, uint32_t rank,
                                            uint32_t* cuts,
                                            uint64_t* n_recv, uint32_t* dest_rank,
                                            int64_t* dest_off, uint32_t* n_seg,
                                            uint32_t* seg_begin, uint32_t* seg_end,
                                            uint32_t* seg_a, uint32_t* seg_tile,
                                            uint32_t* a_tile0, uint32_t* a_start) {
    if (!hist_all || !world || !bins || !chunks || rank >= world) {
        return fail(NPGE_ANCHOR_ERR_PARAMS, "bad arguments");
    }
    DistPlan dp;
    dist_make_plan(hist_all, world, chunks, bins, rank, nullptr, dp);
    auto put = [](auto* dst, const auto& v) {
        if (dst) std::copy(v.begin(), v.end(), dst);
    };
    put(cuts, dp.cuts);
    put(n_recv, dp.n_recv);
    put(dest_rank, dp.dest_rank);
    put(dest_off, dp.dest_off);
    put(seg_begin, dp.seg_begin);
    put(seg_end, dp.seg_end);
    put(seg_a, dp.seg_a);
    put(seg_tile, dp.seg_tile);
    put(a_tile0, dp.a_tile0);
    put(a_start, dp.a_start);
    if (n_seg) *n_seg = dp.n_seg;
    return NPGE_ANCHOR_OK;
}

extern "C" int npge_anchor_dist_buffers(npge_anchor_ctx* ctx, uint64_t n_recv,
                                        npge_anchor_dist_ptrs* out) {
    if (!ctx || !out) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    if (!ctx->dist) return fail(NPGE_ANCHOR_ERR_STATE, "not a multi-GPU context");
    if (n_recv > kMaxRecords) {
        return fail(NPGE_ANCHOR_ERR_CAPACITY, "more than 2^30-1 records on one rank");
    }
    NPGE_CUDA(cudaSetDevice(ctx->device));
    const MsdPlan& plan = ctx->plan;
    if (!ctx->peer_rec.empty()) {  // mapped by the peers: must not move
        if (n_recv > ctx->peer_capacity) {
            return fail(NPGE_ANCHOR_ERR_STATE, "n_recv exceeds the peer buffers' capacity");
        }
    } else {
        NPGE_TRY(ensure(ctx->recv_rec, (n_recv + 16) * 8));
        if (plan.wide) NPGE_TRY(ensure(ctx->recv_hi, (n_recv + 16) * 4));
    }
    const size_t found_bytes = ctx->g_padded / 8 + 256;
    NPGE_TRY(ensure(ctx->found_bits, found_bytes));
    ctx->n_recv = n_recv;
    memset(out, 0, sizeof(*out));
    out->send_rec = ctx->rec[0].ptr;
    out->send_hi = plan.wide ? ctx->rhi[0].ptr : nullptr;
    out->recv_rec = ctx->recv_rec.ptr;
    out->recv_hi = plan.wide ? ctx->recv_hi.ptr : nullptr;
    out->found_bits = ctx->found_bits.ptr;
    out->found_words = found_bytes / 4;
    out->wide = plan.wide;
    out->n_bins = 1 << plan.capa;
    return NPGE_ANCHOR_OK;
}

// Stage 0: pass B over the received (bucket, source) segments, unit packing, and
// the first bucket stage (exact: sort + count + scan -> totals; anchor-similar:
// sort + found bits).  Stage 1 (anchor-similar, after the caller has OR-reduced
// the found bitmap): candidate rule + count + scan -> totals.
extern "C" int npge_anchor_dist_sort(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                                     int32_t stage, uint64_t n_recv, uint32_t n_seg,
                                     const uint32_t* seg_begin, const uint32_t* seg_end,
                                     const uint32_t* seg_a, const uint32_t* seg_tile,
                                     const uint32_t* a_tile0, const uint32_t* a_start,
                                     const uint64_t* used_sorted, uint64_t n_used,
                                     uint64_t* totals) {
    return npge_anchor_dist_sort_impl(ctx, p, stage, n_recv, n_seg, seg_begin, seg_end, seg_a,
                                      seg_tile, a_tile0, a_start, used_sorted, n_used, totals);
}

int npge_anchor_dist_sort_impl(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                               int32_t stage, uint64_t n_recv, uint32_t n_seg,
                               const uint32_t* seg_begin, const uint32_t* seg_end,
                               const uint32_t* seg_a, const uint32_t* seg_tile,
                               const uint32_t* a_tile0, const uint32_t* a_start,
                               const uint64_t* used_sorted, uint64_t n_used,
                               uint64_t* totals) {
    if (!ctx || !totals) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    NPGE_TRY(check_params(p));
    if (!ctx->dist || !ctx->uploaded) {
        return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_dist_upload_range has not been called");
    }
    if (n_recv != ctx->n_recv) {
        return fail(NPGE_ANCHOR_ERR_STATE, "n_recv differs from npge_anchor_dist_buffers");
    }
    if (n_used && !used_sorted) return fail(NPGE_ANCHOR_ERR_PARAMS, "used_sorted is NULL");
    NPGE_CUDA(cudaSetDevice(ctx->device));
    const MsdPlan plan = ctx->plan;
    const bool similar = p->anchor_similar != 0;
    const uint32_t bins_a = 1u << plan.capa;
    const uint64_t max_records = n_recv;
    const uint32_t max_win = bucket_max_windows(plan, max_records);
    const size_t summary_words = bucket_summary_words();
    cudaStream_t st = ctx->stream;
    totals[0] = totals[1] = 0;
    if (stage == 0) {
        if (!seg_begin || !seg_end || !seg_a || !seg_tile || !a_tile0 || !a_start) {
            return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL segment table");
        }
        if (n_seg == 0 || seg_tile[n_seg] > (max_records + kTile - 1) / kTile + n_seg) {
            return fail(NPGE_ANCHOR_ERR_PARAMS, "bad segment tables");
        }
        const uint32_t max_seg = std::max(n_seg, bins_a);
        NPGE_TRY(ensure(ctx->rows, msd_rows_needed(plan, ctx->n_padded, max_records, max_seg) * 4));
        NPGE_TRY(ensure(ctx->blk, msd_blk_needed(plan, ctx->n_padded, max_records, max_seg) * 4));
        NPGE_TRY(ensure(ctx->segtab, seg_table_words(max_seg, plan) * 4));
        NPGE_TRY(ensure(ctx->fine_start, ((size_t(bins_a) << plan.wb) + 1) * 4));
        NPGE_TRY(ensure(ctx->tile_desc, (msd_tiles_b(max_records, max_seg) + 1) * 16));
        NPGE_TRY(ensure(ctx->rec[1], (max_records + 16) * 8));
        if (plan.wide) NPGE_TRY(ensure(ctx->rhi[1], (max_records + 16) * 4));
        NPGE_TRY(ensure(ctx->bstatus, size_t(max_win + 1) * 8 + bins_a * 4));
        NPGE_TRY(ensure(ctx->unit_wtot, size_t(max_win + 1) * (kSortThreads / 32) * 4));
        NPGE_TRY(ensure(ctx->units, size_t(max_win + 1) * 16));
        NPGE_TRY(ensure(ctx->summary, summary_words * 8));
        NPGE_TRY(ensure(ctx->used, n_used * 8));
        if (similar) NPGE_TRY(ensure(ctx->sel, max_records + 16));
        ctx->dist_max_seg = max_seg;
        MsdBuffers mb = dist_msd_buffers(ctx, plan, max_seg);
        mb.seg.n_seg = n_seg;
        if (n_used) {
            NPGE_CUDA(cudaMemcpyAsync(ctx->used.ptr, used_sorted, n_used * 8,
                                      cudaMemcpyHostToDevice, st));
        }
        NPGE_CUDA(cudaMemcpyAsync(mb.seg.seg_begin, seg_begin, n_seg * 4, cudaMemcpyHostToDevice, st));
        NPGE_CUDA(cudaMemcpyAsync(mb.seg.seg_end, seg_end, n_seg * 4, cudaMemcpyHostToDevice, st));
        NPGE_CUDA(cudaMemcpyAsync(mb.seg.seg_a, seg_a, n_seg * 4, cudaMemcpyHostToDevice, st));
        NPGE_CUDA(cudaMemcpyAsync(mb.seg.seg_tile, seg_tile, (n_seg + 1) * 4, cudaMemcpyHostToDevice, st));
        NPGE_CUDA(cudaMemcpyAsync(mb.seg.a_tile0, a_tile0, (bins_a + 1) * 4, cudaMemcpyHostToDevice, st));
        NPGE_CUDA(cudaMemcpyAsync(mb.seg.a_start, a_start, (bins_a + 1) * 4, cudaMemcpyHostToDevice, st));
        NPGE_CUDA(cudaMemsetAsync(ctx->summary.ptr, 0, summary_words * 8, st));
        if (similar) {
            NPGE_CUDA(cudaMemsetAsync(ctx->found_bits.ptr, 0, ctx->g_padded / 8 + 256, st));
            NPGE_CUDA(cudaMemsetAsync(ctx->sel.ptr, 0, max_records + 16, st));
        }
        NPGE_CUDA(cudaEventRecord(ctx->ev[0], st));
        if (max_records) {
            launch_msd_pass_b(plan, ptr<uint64_t>(ctx->recv_rec), ptr<uint32_t>(ctx->recv_hi),
                              max_records, max_seg, mb, st, ctx->ev_pass, &ctx->dist_launches);
        } else {
            NPGE_CUDA(cudaEventRecord(ctx->ev_pass[2], st));
            NPGE_CUDA(cudaEventRecord(ctx->ev_pass[3], st));
        }
        BucketArgs a = bucket_args(ctx, plan, mb, ptr<uint64_t>(ctx->recv_rec),
                                   ptr<uint32_t>(ctx->recv_hi), n_used, 0);
        launch_bucket_stage(a, kBucketUnits, max_records, st, &ctx->dist_launches);
        launch_bucket_stage(a, similar ? kBucketMark : kBucketSort, max_records, st,
                            &ctx->dist_launches);
        NPGE_CUDA(cudaEventRecord(ctx->ev[1], st));
        if (!similar) launch_bucket_stage(a, kBucketScan, max_records, st, &ctx->dist_launches);
        NPGE_CUDA(cudaGetLastError());
        ctx->dist_n_used = n_used;
    } else if (stage == 1) {
        if (!similar) return fail(NPGE_ANCHOR_ERR_PARAMS, "stage 1 is for anchor-similar");
        MsdBuffers mb = dist_msd_buffers(ctx, plan, ctx->dist_max_seg);
        BucketArgs a = bucket_args(ctx, plan, mb, ptr<uint64_t>(ctx->recv_rec),
                                   ptr<uint32_t>(ctx->recv_hi), ctx->dist_n_used, 0);
        NPGE_CUDA(cudaEventRecord(ctx->ev[0], st));
        launch_bucket_stage(a, kBucketCount, max_records, st, &ctx->dist_launches);
        launch_bucket_stage(a, kBucketScan, max_records, st, &ctx->dist_launches);
        NPGE_CUDA(cudaGetLastError());
        NPGE_CUDA(cudaEventRecord(ctx->ev[1], st));
    } else {
        return fail(NPGE_ANCHOR_ERR_PARAMS, "stage must be 0 or 1");
    }
    uint64_t summary[kSumWords] = {};
    if (max_records && (stage == 1 || !similar)) {
        NPGE_CUDA(cudaMemcpyAsync(summary, ctx->summary.ptr, sizeof(summary),
                                  cudaMemcpyDeviceToHost, st));
    }
    NPGE_CUDA(cudaStreamSynchronize(st));
    float ms = 0;
    if (stage == 0) {
        cudaEventElapsedTime(&ctx->dist_ms[2], ctx->ev[0], ctx->ev_pass[2]);
        cudaEventElapsedTime(&ctx->dist_ms[3], ctx->ev_pass[2], ctx->ev_pass[3]);
        cudaEventElapsedTime(&ctx->dist_ms[4], ctx->ev_pass[3], ctx->ev[1]);
    } else {
        cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]);
        ctx->dist_ms[5] += ms;
    }
    totals[0] = summary[kSumFragments];
    totals[1] = summary[kSumBlocks];
    return NPGE_ANCHOR_OK;
}

// Write the anchors of this rank's hash range, at most local_limit fragments
// (the job's max-anchor-fragments minus the totals of the ranks before this one).
extern "C" int npge_anchor_dist_emit(npge_anchor_ctx* ctx, const npge_anchor_params* p,
                                     uint64_t local_limit, npge_anchor_stats* stats) {
    if (!ctx) return fail(NPGE_ANCHOR_ERR_PARAMS, "ctx is NULL");
    NPGE_TRY(check_params(p));
    if (!ctx->dist || !ctx->uploaded) {
        return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_dist_upload_range has not been called");
    }
    NPGE_CUDA(cudaSetDevice(ctx->device));
    const MsdPlan plan = ctx->plan;
    const bool similar = p->anchor_similar != 0;
    const uint64_t max_records = ctx->n_recv;
    const uint64_t max_frags = std::min<uint64_t>(max_records, local_limit);
    // compact layout: (job-wide position << 1 | strand) only, as on one GPU
    const bool compact = ctx->layout == 1;
    if (compact) {
        NPGE_TRY(ensure(ctx->r_pay, max_frags * 4));
    } else {
        NPGE_TRY(ensure(ctx->r_seq, max_frags * 4));
        NPGE_TRY(ensure(ctx->r_pos, max_frags * 4));
        NPGE_TRY(ensure(ctx->r_ori, max_frags));
    }
    NPGE_TRY(ensure(ctx->r_block_off, (max_frags / 2 + 2) * 8));
    NPGE_TRY(ensure(ctx->r_block_hash, (max_frags / 2 + 2) * 8));
    cudaStream_t st = ctx->stream;
    MsdBuffers mb = dist_msd_buffers(ctx, plan, ctx->dist_max_seg);
    BucketArgs a = bucket_args(ctx, plan, mb, ptr<uint64_t>(ctx->recv_rec),
                               ptr<uint32_t>(ctx->recv_hi), ctx->dist_n_used, local_limit);
    if (compact) {
        a.r_pay = ptr<uint32_t>(ctx->r_pay);
        a.r_seq = nullptr;
        a.r_pos = nullptr;
        a.r_ori = nullptr;
    }
    NPGE_CUDA(cudaEventRecord(ctx->ev[0], st));
    launch_bucket_stage(a, similar ? kBucketEmitSimilar : kBucketEmit, max_records, st,
                        &ctx->dist_launches);
    launch_bucket_stage(a, kBucketFinal, max_records, st, &ctx->dist_launches);
    NPGE_CUDA(cudaGetLastError());
    NPGE_CUDA(cudaEventRecord(ctx->ev[1], st));
    uint64_t summary[kSumWords];
    NPGE_CUDA(cudaMemcpyAsync(summary, ctx->summary.ptr, sizeof(summary),
                              cudaMemcpyDeviceToHost, st));
    NPGE_CUDA(cudaStreamSynchronize(st));
    cudaEventElapsedTime(&ctx->dist_ms[6], ctx->ev[0], ctx->ev[1]);
    npge_anchor_stats& s = ctx->stats;
    memset(&s, 0, sizeof(s));
    s.n_positions = ctx->n_bases;
    s.n_windows = max_records;  // records this rank sorted
    s.n_fragments = summary[kSumEmitted];
    s.n_blocks = summary[kSumBlocksEmitted];
    s.digest[0] = summary[kSumDigest0];
    s.digest[1] = summary[kSumDigest1];
    s.radix_passes = uint32_t(2 + plan.smem_passes);
    s.gpu_launches = ctx->dist_launches;
    // stages as on one GPU: [count A] | [pass A] [count B] [pass B] [bucket sort
    // (+ candidate count)] [scan + emit]
    s.ms_hist = ctx->dist_ms[0];
    for (int i = 0; i < 3; i++) s.ms_pass[i] = ctx->dist_ms[i + 1];
    s.ms_pass[0] += ctx->dist_ms[7];  // bulk exchange: what the copies add to pass A
    s.ms_pass[5] = ctx->dist_ms[7];
    s.ms_pass[6] = ctx->dist_exchange_ms;  // start of pass A .. last copy landed
    s.ms_pass[3] = ctx->dist_ms[4] + ctx->dist_ms[5];
    s.ms_pass[4] = ctx->dist_ms[6];
    s.ms_sort = ctx->dist_ms[0] + ctx->dist_ms[1] + ctx->dist_ms[2] + ctx->dist_ms[3] +
                ctx->dist_ms[7];
    s.ms_group = ctx->dist_ms[4] + ctx->dist_ms[5] + ctx->dist_ms[6];
    s.ms_total = s.ms_sort + s.ms_group;  // device time of this rank's kernels
    const int model_passes = (2 * p->anchor_size + 7) / 8;
    s.bytes_algorithmic = uint64_t(double(ctx->n_bases) * (32.375 + 24.0 * model_passes)) +
                          9 * s.n_fragments + 16 * s.n_blocks;
    const double rec = plan.wide ? 12.0 : 8.0;
    s.bytes_moved = uint64_t(double(ctx->n_bases) * 0.75 + double(ctx->n_positions) * rec +
                             double(max_records) * rec * (similar ? 6.0 : 4.0)) +
                    (compact ? 4 : 9) * s.n_fragments + 16 * s.n_blocks;
    ctx->total_selected = summary[kSumFragments];
    ctx->have_pay = compact;
    ctx->result_compact = compact;
    ctx->last_params = *p;
    ctx->last_params.max_anchor_fragments = int64_t(std::min<uint64_t>(local_limit, uint64_t(INT64_MAX)));
    NPGE_TRY(fix_cut_resident(ctx, local_limit));
    ctx->have_result = true;
    if (stats) *stats = s;
    return NPGE_ANCHOR_OK;
}

// ==== FragmentFinder (src/algo/FragmentFinder.cpp) =================================
namespace {

// 2-bit codes of the reference (A0 T1 G2 C3; char_to_size.hpp:24-36); N -> code 0
// + mask bit, like pack_kernel.  Returns false for a byte outside ATGCN.
bool pack_pattern(const std::string& p, std::vector<uint64_t>& key, std::vector<uint32_t>& mask) {
    const size_t chunks = (p.size() + 31) / 32;
    key.assign(chunks, 0);
    mask.assign(chunks, 0);
    for (size_t i = 0; i < p.size(); i++) {
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

struct FragmentHost {
    PinnedVector<uint32_t> seq, pos;
    PinnedVector<int8_t> ori;
};

}  // namespace

extern "C" int npge_fragment_find_device(npge_anchor_ctx* ctx, const uint8_t* pattern,
                                         uint32_t pattern_len, int64_t max_matches,
                                         npge_fragment_result* out) {
    if (!out) return fail(NPGE_ANCHOR_ERR_PARAMS, "out is NULL");
    memset(out, 0, sizeof(*out));
    if (!ctx || !pattern) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    // the option checks of the processor: FragmentFinder.cpp:18-28, :37
    if (pattern_len == 0) return fail(NPGE_ANCHOR_ERR_PARAMS, "'pattern' should not be empty");
    if (max_matches < 1) return fail(NPGE_ANCHOR_ERR_PARAMS, "max-matches >= 1 is violated");
    if (!ctx->uploaded || ctx->dist) {
        return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_upload has not been called");
    }
    if (uint32_t(ctx->upload_k) > pattern_len) {
        return fail(NPGE_ANCHOR_ERR_STATE,
                    "sequences were uploaded with an anchor-size above the pattern length");
    }
    std::string fwd(reinterpret_cast<const char*>(pattern), pattern_len), rc(pattern_len, 'N');
    for (uint32_t i = 0; i < pattern_len; i++) {
        char c = fwd[pattern_len - 1 - i];  // complement.hpp:19-32: N stays N
        rc[i] = c == 'A' ? 'T' : c == 'T' ? 'A' : c == 'G' ? 'C' : c == 'C' ? 'G' : c;
    }
    std::vector<uint64_t> key[2];
    std::vector<uint32_t> mask[2];
    if (!pack_pattern(fwd, key[0], mask[0]) || !pack_pattern(rc, key[1], mask[1])) {
        return fail(NPGE_ANCHOR_ERR_INPUT, "pattern holds a byte outside ATGCN (to_atgcn first)");
    }
    NPGE_CUDA(cudaSetDevice(ctx->device));
    const uint32_t chunks = uint32_t(key[0].size());
    const uint32_t tiles = fragment_tiles(ctx->n_padded);
    const uint64_t cap = std::min<uint64_t>(uint64_t(max_matches), 2 * ctx->n_positions);
    NPGE_TRY(ensure(ctx->frag_pat, size_t(chunks) * 24));
    NPGE_TRY(ensure(ctx->frag_tiles, size_t(tiles + 1) * 12 + 16));
    NPGE_TRY(ensure(ctx->frag_out, cap * 9 + 64));
    cudaStream_t st = ctx->stream;
    uint8_t* pat = ptr<uint8_t>(ctx->frag_pat);
    for (int s = 0; s < 2; s++) {
        NPGE_CUDA(cudaMemcpyAsync(pat + size_t(s) * chunks * 8, key[s].data(), chunks * 8,
                                  cudaMemcpyHostToDevice, st));
        NPGE_CUDA(cudaMemcpyAsync(pat + size_t(chunks) * 16 + size_t(s) * chunks * 4,
                                  mask[s].data(), chunks * 4, cudaMemcpyHostToDevice, st));
    }
    FragmentArgs a = {};
    a.packed = ptr<uint32_t>(ctx->packed);
    a.nmask = ptr<uint32_t>(ctx->nmask);
    a.n_positions = ctx->n_positions;
    a.n_padded = ctx->n_padded;
    a.pat_key = reinterpret_cast<const uint64_t*>(pat);
    a.pat_mask = reinterpret_cast<const uint32_t*>(pat + size_t(chunks) * 16);
    a.pattern_len = pattern_len;
    a.n_chunks = chunks;
    a.seq_start = ptr<uint64_t>(ctx->seq_start);
    a.seq_caller = ptr<uint32_t>(ctx->seq_caller);
    a.seq_lut = ptr<uint32_t>(ctx->seq_lut);
    a.lut_shift = ctx->lut_shift;
    a.max_matches = cap;
    uint8_t* tl = ptr<uint8_t>(ctx->frag_tiles);
    a.summary = reinterpret_cast<uint64_t*>(tl);
    a.tile_offset = reinterpret_cast<uint64_t*>(tl + 16);
    a.tile_count = reinterpret_cast<uint32_t*>(tl + 16 + size_t(tiles) * 8);
    uint8_t* ro = ptr<uint8_t>(ctx->frag_out);
    a.r_seq = reinterpret_cast<uint32_t*>(ro);
    a.r_pos = reinterpret_cast<uint32_t*>(ro + cap * 4 + 16 - (cap * 4) % 16);
    a.r_ori = reinterpret_cast<int8_t*>(reinterpret_cast<uint8_t*>(a.r_pos) + cap * 4);
    uint32_t launches = 0;
    uint64_t total = 0;
    if (tiles) {
        launch_fragment_find(a, 0, st, &launches);
        launch_fragment_find(a, 1, st, &launches);
        NPGE_CUDA(cudaGetLastError());
        NPGE_CUDA(cudaMemcpyAsync(&total, a.summary, 8, cudaMemcpyDeviceToHost, st));
    }
    NPGE_CUDA(cudaStreamSynchronize(st));
    const uint64_t n = std::min<uint64_t>(total, cap);
    FragmentHost* r = new (std::nothrow) FragmentHost;
    if (!r) return fail(NPGE_ANCHOR_ERR_MEMORY, "out of host memory");
    try {
        r->seq.resize(n);
        r->pos.resize(n);
        r->ori.resize(n);
    } catch (std::bad_alloc&) {
        delete r;
        return fail(NPGE_ANCHOR_ERR_MEMORY, "out of host memory for the result");
    }
    if (n) {
        cudaMemcpyAsync(r->seq.data(), a.r_seq, n * 4, cudaMemcpyDeviceToHost, st);
        cudaMemcpyAsync(r->pos.data(), a.r_pos, n * 4, cudaMemcpyDeviceToHost, st);
        cudaMemcpyAsync(r->ori.data(), a.r_ori, n, cudaMemcpyDeviceToHost, st);
        cudaError_t e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) {
            delete r;
            return fail(NPGE_ANCHOR_ERR_CUDA, cudaGetErrorString(e));
        }
    }
    out->n_matches = n;
    out->total_matches = total;
    out->frag_seq = r->seq.data();
    out->frag_min_pos = r->pos.data();
    out->frag_ori = r->ori.data();
    out->owner_ = r;
    ctx->stats.gpu_launches = launches;
    return NPGE_ANCHOR_OK;
}

extern "C" int npge_fragment_find(npge_anchor_ctx* ctx, const uint8_t* pattern,
                                  uint32_t pattern_len, int64_t max_matches,
                                  const uint8_t* bases, const uint64_t* seq_offsets,
                                  uint32_t n_seqs, npge_fragment_result* out) {
    if (!out) return fail(NPGE_ANCHOR_ERR_PARAMS, "out is NULL");
    memset(out, 0, sizeof(*out));
    if (pattern_len == 0) return fail(NPGE_ANCHOR_ERR_PARAMS, "'pattern' should not be empty");
    const int32_t k = int32_t(std::min<uint32_t>(pattern_len, NPGE_ANCHOR_MAX_SIZE));
    NPGE_TRY(npge_anchor_upload(ctx, k, bases, seq_offsets, n_seqs));
    return npge_fragment_find_device(ctx, pattern, pattern_len, max_matches, out);
}

extern "C" void npge_fragment_result_free(npge_fragment_result* out) {
    if (!out) return;
    delete static_cast<FragmentHost*>(out->owner_);
    memset(out, 0, sizeof(*out));
}

// ==== block_hash of flat block lists (block_hash_kernels.cu) ========================
namespace {

// Fragments per launch: bounds the key scratch (32 B per fragment) at 1 GiB.
// NPGE_BLOCK_HASH_BATCH overrides it (the tests run many small batches).
uint64_t block_hash_batch() {
    const char* e = getenv("NPGE_BLOCK_HASH_BATCH");
    const long long v = e ? atoll(e) : 0;
    return v > 0 ? uint64_t(v) : uint64_t(1) << 25;
}

// Device pointers of a flat block list; hashes of all blocks -> host.
int block_hashes_device(npge_anchor_ctx* ctx, const char* const* names, uint32_t n_seqs,
                        uint64_t n_blocks, uint64_t n_fragments, const uint64_t* d_block_off,
                        const uint32_t* d_seq, const uint32_t* d_min, const uint32_t* d_max,
                        uint32_t frag_len, const int8_t* d_ori, uint64_t* hashes,
                        uint64_t* blockset_hash) {
    if (!names && n_seqs) return fail(NPGE_ANCHOR_ERR_PARAMS, "names is NULL");
    NameTable table;
    const std::string why = build_name_table(names, n_seqs, table);
    if (!why.empty()) return fail(NPGE_ANCHOR_ERR_PARAMS, why);
    if (blockset_hash) *blockset_hash = 0;
    if (n_blocks == 0) return NPGE_ANCHOR_OK;
    cudaStream_t st = ctx->stream;
    // names: [words u64][seq_rank u32 x n_seqs][pos u32 x ranks][len u32 x ranks]
    const size_t n_words = table.words.size(), n_ranks = table.pos.size();
    static const std::vector<uint16_t> bcd4 = make_bcd4();
    // + [bcd4 u16 x 10000][slow u8 x ranks]
    NPGE_TRY(ensure(ctx->bh_names, n_words * 8 + (size_t(n_seqs) + 2 * n_ranks) * 4 +
                                       bcd4.size() * 2 + n_ranks));
    uint64_t* d_words = ptr<uint64_t>(ctx->bh_names);
    uint32_t* d_rank = reinterpret_cast<uint32_t*>(d_words + n_words);
    uint32_t* d_pos = d_rank + n_seqs;
    uint32_t* d_len = d_pos + n_ranks;
    uint16_t* d_bcd4 = reinterpret_cast<uint16_t*>(d_len + n_ranks);
    unsigned char* d_slow = nullptr;  // only for name sets that need the string order
    if (!table.slow.empty()) {
        d_slow = reinterpret_cast<unsigned char*>(d_bcd4 + bcd4.size());
        NPGE_CUDA(cudaMemcpyAsync(d_slow, table.slow.data(), n_ranks, cudaMemcpyHostToDevice, st));
    }
    NPGE_CUDA(cudaMemcpyAsync(d_bcd4, bcd4.data(), bcd4.size() * 2, cudaMemcpyHostToDevice, st));
    NPGE_CUDA(cudaMemcpyAsync(d_words, table.words.data(), n_words * 8, cudaMemcpyHostToDevice,
                              st));
    if (n_seqs) {
        NPGE_CUDA(cudaMemcpyAsync(d_rank, table.seq_rank.data(), size_t(n_seqs) * 4,
                                  cudaMemcpyHostToDevice, st));
    }
    NPGE_CUDA(cudaMemcpyAsync(d_pos, table.pos.data(), n_ranks * 4, cudaMemcpyHostToDevice, st));
    NPGE_CUDA(cudaMemcpyAsync(d_len, table.len.data(), n_ranks * 4, cudaMemcpyHostToDevice, st));
    // out: [hashes u64 x n_blocks][xor u64][error u32, pad][cut_block][cut_frag]
    const uint64_t batch = block_hash_batch();
    if ((n_fragments + batch - 1) / batch > (1u << 20)) {
        return fail(NPGE_ANCHOR_ERR_PARAMS, "NPGE_BLOCK_HASH_BATCH is too small");
    }
    const uint32_t n_cuts = std::max<uint32_t>(1, uint32_t((n_fragments + batch - 1) / batch));
    NPGE_TRY(ensure(ctx->bh_out, (n_blocks + 2 + 2 * (size_t(n_cuts) + 1)) * 8));
    uint64_t* d_hashes = ptr<uint64_t>(ctx->bh_out);
    uint64_t* d_xor = d_hashes + n_blocks;
    uint32_t* d_error = reinterpret_cast<uint32_t*>(d_xor + 1);
    uint64_t* d_cut_block = d_xor + 2;
    uint64_t* d_cut_frag = d_cut_block + n_cuts + 1;
    NPGE_CUDA(cudaMemsetAsync(d_xor, 0, 16, st));
    std::vector<uint64_t> cuts(2 * (size_t(n_cuts) + 1), 0);
    launch_block_hash_cuts(d_block_off, n_blocks, batch, n_cuts, d_cut_block,
                           d_cut_frag, st);
    NPGE_CUDA(cudaGetLastError());
    NPGE_CUDA(cudaMemcpyAsync(cuts.data(), d_cut_block, cuts.size() * 8,
                              cudaMemcpyDeviceToHost, st));
    NPGE_CUDA(cudaStreamSynchronize(st));
    const uint64_t* cut_block = cuts.data();
    const uint64_t* cut_frag = cuts.data() + n_cuts + 1;
    const uint32_t n_launch = n_cuts;
    uint64_t max_batch = 0;
    for (uint32_t i = 0; i < n_launch; i++) {
        if (cut_block[i] > cut_block[i + 1] || cut_frag[i] > cut_frag[i + 1] ||
            cut_frag[i + 1] > n_fragments) {
            return fail(NPGE_ANCHOR_ERR_PARAMS, "block_offsets are not ascending");
        }
        max_batch = std::max(max_batch, cut_frag[i + 1] - cut_frag[i]);
    }
    NPGE_TRY(ensure(ctx->bh_scratch, max_batch * 2 * sizeof(IdKey)));
    NPGE_CUDA(cudaEventRecord(ctx->ev[0], st));
    BlockHashArgs a = {};
    a.block_off = d_block_off;
    a.n_fragments = n_fragments;
    a.frag_seq = d_seq;
    a.min_pos = d_min;
    a.max_pos = d_max;
    a.frag_len = frag_len;
    a.frag_ori = d_ori;
    a.n_seqs = n_seqs;
    a.seq_rank = d_rank;
    a.name_words = d_words;
    a.name_pos = d_pos;
    a.name_len = d_len;
    a.name_slow = d_slow;
    a.bcd4 = d_bcd4;
    a.scratch = ctx->bh_scratch.ptr;
    a.hashes = d_hashes;
    a.xor_out = d_xor;
    a.error = d_error;
    // one launch per batch of blocks between two cuts (cut 0 = block 0)
    for (uint32_t i = 0; i < n_launch; i++) {
        a.b0 = cut_block[i];
        a.b1 = cut_block[i + 1];
        a.f_base = cut_frag[i];
        a.f_end = cut_frag[i + 1];
        launch_block_hash(a, st);
        NPGE_CUDA(cudaGetLastError());
    }
    NPGE_CUDA(cudaEventRecord(ctx->ev[1], st));
    uint64_t tail[2] = {0, 0};
    NPGE_CUDA(cudaMemcpyAsync(tail, d_xor, 16, cudaMemcpyDeviceToHost, st));
    if (hashes) {
        NPGE_CUDA(cudaMemcpyAsync(hashes, d_hashes, n_blocks * 8, cudaMemcpyDeviceToHost, st));
    }
    NPGE_CUDA(cudaStreamSynchronize(st));
    NPGE_CUDA(cudaEventElapsedTime(&ctx->bh_ms, ctx->ev[0], ctx->ev[1]));
    const uint32_t err = uint32_t(tail[1]);
    if (err & 1) return fail(NPGE_ANCHOR_ERR_PARAMS, "block_offsets are not ascending");
    if (err & 2) return fail(NPGE_ANCHOR_ERR_PARAMS, "frag_seq out of range");
    if (blockset_hash) *blockset_hash = tail[0];
    return NPGE_ANCHOR_OK;
}

}  // namespace

extern "C" int npge_block_hashes(npge_anchor_ctx* ctx, const char* const* names,
                                 uint32_t n_seqs, uint64_t n_blocks,
                                 const uint64_t* block_offsets, const uint32_t* frag_seq,
                                 const uint32_t* frag_min_pos, const uint32_t* frag_max_pos,
                                 uint32_t frag_len, const int8_t* frag_ori, uint64_t* hashes,
                                 uint64_t* blockset_hash) {
    if (!ctx) return fail(NPGE_ANCHOR_ERR_PARAMS, "ctx is NULL");
    if (!block_offsets) return fail(NPGE_ANCHOR_ERR_PARAMS, "block_offsets is NULL");
    if (block_offsets[0] != 0) return fail(NPGE_ANCHOR_ERR_PARAMS, "block_offsets[0] must be 0");
    const uint64_t nf = block_offsets[n_blocks];
    if (nf && (!frag_seq || !frag_min_pos || !frag_ori)) {
        return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL fragment array");
    }
    if (!frag_max_pos && frag_len == 0 && nf) {
        return fail(NPGE_ANCHOR_ERR_PARAMS, "frag_len must be > 0 without frag_max_pos");
    }
    NPGE_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    // in: [block_off u64 x n_blocks + 1][seq u32][min u32][max u32][ori i8]
    const size_t off_bytes = (n_blocks + 1) * 8;
    NPGE_TRY(ensure(ctx->bh_in, off_bytes + nf * 13 + 64));
    uint8_t* base = static_cast<uint8_t*>(ctx->bh_in.ptr);
    uint64_t* d_off = reinterpret_cast<uint64_t*>(base);
    uint32_t* d_seq = reinterpret_cast<uint32_t*>(base + off_bytes);
    uint32_t* d_min = d_seq + nf;
    uint32_t* d_max = d_min + nf;
    int8_t* d_ori = reinterpret_cast<int8_t*>(d_max + nf);
    NPGE_CUDA(cudaMemcpyAsync(d_off, block_offsets, off_bytes, cudaMemcpyHostToDevice, st));
    if (nf) {
        NPGE_CUDA(cudaMemcpyAsync(d_seq, frag_seq, nf * 4, cudaMemcpyHostToDevice, st));
        NPGE_CUDA(cudaMemcpyAsync(d_min, frag_min_pos, nf * 4, cudaMemcpyHostToDevice, st));
        if (frag_max_pos) {
            NPGE_CUDA(cudaMemcpyAsync(d_max, frag_max_pos, nf * 4, cudaMemcpyHostToDevice, st));
        }
        NPGE_CUDA(cudaMemcpyAsync(d_ori, frag_ori, nf, cudaMemcpyHostToDevice, st));
    }
    return block_hashes_device(ctx, names, n_seqs, n_blocks, nf, d_off, d_seq, d_min,
                               frag_max_pos ? d_max : nullptr, frag_len, d_ori, hashes,
                               blockset_hash);
}

extern "C" int npge_anchor_block_hashes(npge_anchor_ctx* ctx, const char* const* names,
                                        uint32_t n_seqs, uint64_t* hashes,
                                        uint64_t* blockset_hash) {
    if (!ctx) return fail(NPGE_ANCHOR_ERR_PARAMS, "ctx is NULL");
    if (!ctx->have_result) {
        return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_find_device has not run");
    }
    if (ctx->result_compact) {
        return fail(NPGE_ANCHOR_ERR_STATE, "the resident result is compact (payloads only): "
                    "find with NPGE_ANCHOR_LAYOUT_FULL, or hash fetched arrays with npge_block_hashes");
    }
    const uint32_t need = ctx->dist ? uint32_t(ctx->gseqs.size()) : ctx->n_caller_seqs;
    if (n_seqs < need) {
        return fail(NPGE_ANCHOR_ERR_PARAMS, "fewer names than uploaded sequences");
    }
    NPGE_CUDA(cudaSetDevice(ctx->device));
    return block_hashes_device(ctx, names, n_seqs, ctx->stats.n_blocks,
                               ctx->stats.n_fragments, ptr<uint64_t>(ctx->r_block_off),
                               ptr<uint32_t>(ctx->r_seq), ptr<uint32_t>(ctx->r_pos), nullptr,
                               uint32_t(ctx->last_params.anchor_size), ptr<int8_t>(ctx->r_ori),
                               hashes, blockset_hash);
}

extern "C" float npge_block_hashes_last_ms(npge_anchor_ctx* ctx) {
    return ctx ? ctx->bh_ms : 0.f;
}
