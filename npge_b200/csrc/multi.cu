// multi.cu -- AnchorFinder over several GPUs of ONE process (npge_anchor_multi_*):
// the C++ host that NPGE's Processor reaches through npge::gpu::AnchorFinder
// (set_devices).  One context and one host thread per GPU; the threads drive the
// same per-rank stages as the multi-process driver (npge_b200/dist.py), with
// in-process rendezvous instead of collectives: a barrier, a shared histogram
// table, receive buffers reached through peer access.
//
// Replaces the reference's thread pool over sequences (ReusingThreadGroup,
// src/util/thread_group.cpp:70-139; one whole sequence per task,
// src/algo/AnchorFinder.hpp:27-29): positions, not sequences, are sharded.
#include <algorithm>
#include <condition_variable>
#include <cstring>
#include <functional>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "kernels.h"

using namespace npge_b200;

namespace {

// Reusable barrier of the rank threads (C++17: no std::barrier).
class Barrier {
public:
    explicit Barrier(int n) : n_(n) {}
    void wait() {
        std::unique_lock<std::mutex> lock(mu_);
        const uint64_t gen = gen_;
        if (++count_ == n_) {
            count_ = 0;
            gen_++;
            cv_.notify_all();
        } else {
            cv_.wait(lock, [&] { return gen_ != gen; });
        }
    }
private:
    std::mutex mu_;
    std::condition_variable cv_;
    int n_, count_ = 0;
    uint64_t gen_ = 0;
};

// The rank threads of one npge_anchor_multi: created once, parked between calls (a
// call costs a condition-variable handoff per rank, not a thread, and every thread
// keeps its CUDA per-thread state).  Rank 0 runs on the caller's thread.
class RankThreads {
public:
    explicit RankThreads(int n) : n_(n) {
        for (int r = 1; r < n; r++) threads_.emplace_back([this, r] { loop(r); });
    }
    ~RankThreads() {
        {
            std::lock_guard<std::mutex> lock(mu_);
            stop_ = true;
            gen_++;
        }
        cv_.notify_all();
        for (auto& t : threads_) t.join();
    }
    // fn(rank) on every rank's thread; returns when all of them have returned
    void run(const std::function<void(int)>& fn) {
        {
            std::lock_guard<std::mutex> lock(mu_);
            fn_ = &fn;
            pending_ = n_ - 1;
            gen_++;
        }
        cv_.notify_all();
        fn(0);
        std::unique_lock<std::mutex> lock(mu_);
        done_.wait(lock, [&] { return pending_ == 0; });
        fn_ = nullptr;
    }
private:
    void loop(int r) {
        uint64_t seen = 0;
        for (;;) {
            const std::function<void(int)>* fn = nullptr;
            {
                std::unique_lock<std::mutex> lock(mu_);
                cv_.wait(lock, [&] { return gen_ != seen; });
                seen = gen_;
                if (stop_) return;
                fn = fn_;
            }
            (*fn)(r);
            {
                std::lock_guard<std::mutex> lock(mu_);
                if (--pending_ == 0) done_.notify_all();
            }
        }
    }
    int n_;
    std::vector<std::thread> threads_;
    std::mutex mu_;
    std::condition_variable cv_, done_;
    const std::function<void(int)>* fn_ = nullptr;
    uint64_t gen_ = 0;
    int pending_ = 0;
    bool stop_ = false;
};

}  // namespace

struct npge_anchor_multi {
    std::vector<int> devices;
    std::vector<npge_anchor_ctx*> ctx;
    RankThreads* threads = nullptr;
    int32_t k = 0;
    bool uploaded = false, have_result = false;
    uint64_t capacity = 0;       // records every receive buffer can take
    int chunks = 4;
    npge_anchor_params last = {0, 0, 0};
    std::vector<npge_anchor_stats> stats;  // per rank, last find_device
    npge_anchor_stats job = {};
};

namespace {

int fail(int code, const std::string& msg) { return set_last_error(code, msg); }

// Runs fn(rank) on one thread per rank; the first error (code, text) wins.  A rank
// that failed keeps arriving at the barriers (fn must not return early between them),
// so nobody is left waiting.
struct RankError {
    std::mutex mu;
    int code = NPGE_ANCHOR_OK;
    std::string text;
    void set(int c) {
        std::lock_guard<std::mutex> lock(mu);
        if (code == NPGE_ANCHOR_OK && c != NPGE_ANCHOR_OK) {
            code = c;
            text = npge_anchor_last_error();
        }
    }
    bool failed() {
        std::lock_guard<std::mutex> lock(mu);
        return code != NPGE_ANCHOR_OK;
    }
};

void run_ranks(npge_anchor_multi* m, const std::function<void(int)>& fn) {
    m->threads->run(fn);
}

int peer_setup(npge_anchor_multi* m, uint64_t capacity) {
    const int G = int(m->ctx.size());
    std::vector<void*> rec(G), hi(G);
    uint8_t handle[NPGE_ANCHOR_IPC_BYTES];
    for (int r = 0; r < G; r++) {
        int rc = npge_anchor_dist_peer_close(m->ctx[r]);
        if (rc) return rc;
    }
    for (int r = 0; r < G; r++) {
        int rc = npge_anchor_dist_peer_alloc(m->ctx[r], capacity, handle);
        if (rc) return rc;
        rec[r] = m->ctx[r]->recv_rec.ptr;
        hi[r] = m->ctx[r]->recv_hi.ptr;
    }
    for (int r = 0; r < G; r++) {
        int rc = npge_anchor_dist_peer_attach(m->ctx[r], uint32_t(G), uint32_t(r), rec.data(),
                                              hi.data(), m->devices.data());
        if (rc) return rc;
    }
    m->capacity = capacity;
    return NPGE_ANCHOR_OK;
}

}  // namespace

extern "C" int npge_anchor_multi_create(npge_anchor_multi** out, const int* device_ids,
                                        int n_devices) {
    if (!out) return fail(NPGE_ANCHOR_ERR_PARAMS, "out is NULL");
    *out = nullptr;
    if (!device_ids || n_devices < 1 || n_devices > 8) {
        return fail(NPGE_ANCHOR_ERR_PARAMS, "1..8 devices");
    }
    npge_anchor_multi* m = new (std::nothrow) npge_anchor_multi;
    if (!m) return fail(NPGE_ANCHOR_ERR_MEMORY, "out of host memory");
    for (int r = 0; r < n_devices; r++) {
        npge_anchor_ctx* c = nullptr;
        int rc = npge_anchor_create(&c, device_ids[r]);
        if (rc) {
            for (npge_anchor_ctx* x : m->ctx) npge_anchor_destroy(x);
            delete m;
            return rc;
        }
        m->ctx.push_back(c);
        m->devices.push_back(device_ids[r]);
    }
    m->stats.resize(n_devices);
    m->threads = new RankThreads(n_devices);
    *out = m;
    return NPGE_ANCHOR_OK;
}

extern "C" void npge_anchor_multi_destroy(npge_anchor_multi* m) {
    if (!m) return;
    delete m->threads;
    for (npge_anchor_ctx* c : m->ctx) npge_anchor_dist_peer_close(c);
    for (npge_anchor_ctx* c : m->ctx) npge_anchor_destroy(c);
    delete m;
}

extern "C" int npge_anchor_multi_devices(const npge_anchor_multi* m) {
    return m ? int(m->ctx.size()) : 0;
}

// ---- upload: every GPU gets its range of the job's position space ------------------
extern "C" int npge_anchor_multi_upload(npge_anchor_multi* m, int32_t anchor_size,
                                        const uint8_t* bases, const uint64_t* seq_offsets,
                                        uint32_t n_seqs) {
    if (!m) return fail(NPGE_ANCHOR_ERR_PARAMS, "multi context is NULL");
    if (n_seqs && (!seq_offsets || (!bases && seq_offsets[n_seqs] > 0))) {
        return fail(NPGE_ANCHOR_ERR_PARAMS, "bases / seq_offsets is NULL");
    }
    if (anchor_size <= 0 || anchor_size > NPGE_ANCHOR_MAX_SIZE) {
        return fail(NPGE_ANCHOR_ERR_PARAMS, anchor_size <= 0 ? "anchor-size > 0 is violated"
                                                             : "anchor-size <= 32 is violated");
    }
    m->uploaded = m->have_result = false;
    const int G = int(m->ctx.size());
    // SeqBase::make_seqs (SeqI.hpp:45-57): size >= k, by size descending, ties by index
    struct S { uint32_t caller; uint64_t size; };
    std::vector<S> order;
    for (uint32_t s = 0; s < n_seqs; s++) {
        if (seq_offsets[s + 1] < seq_offsets[s]) {
            return fail(NPGE_ANCHOR_ERR_PARAMS, "seq_offsets must be non-decreasing");
        }
        const uint64_t size = seq_offsets[s + 1] - seq_offsets[s];
        if (size >= uint64_t(anchor_size)) order.push_back({s, size});
    }
    std::stable_sort(order.begin(), order.end(),
                     [](const S& a, const S& b) { return a.size > b.size; });
    const uint32_t n_global = uint32_t(order.size());
    std::vector<uint64_t> gstart(n_global + 1, 0);
    std::vector<uint32_t> gcaller(n_global + 1, 0);
    for (uint32_t i = 0; i < n_global; i++) {
        gstart[i + 1] = gstart[i] + order[i].size + 1;  // one pad after every sequence
        gcaller[i] = order[i].caller;
    }
    const uint64_t total = gstart[n_global];
    std::vector<uint64_t> cuts;
    dist_position_cuts(total, uint32_t(G), kPosAlign, cuts);
    RankError err;
    run_ranks(m, [&](int r) {
        const uint64_t g0 = cuts[r], n_own = cuts[r + 1] - cuts[r];
        const uint64_t have_end = std::min(cuts[r + 1] + uint64_t(anchor_size) - 1, total);
        // the parts of the sequences inside [g0, have_end): straight from the caller's buffer
        std::vector<uint64_t> src_off, dst_pos, len;
        size_t i = std::upper_bound(gstart.begin(), gstart.end(), g0) - gstart.begin();
        i = i ? i - 1 : 0;
        for (; i < n_global && gstart[i] < have_end; i++) {
            const uint64_t b = gstart[i], e = gstart[i + 1] - 1;  // bases of sequence i
            const uint64_t x = std::max(b, g0), y = std::min(e, have_end);
            if (y > x) {
                src_off.push_back(seq_offsets[order[i].caller] + (x - b));
                dst_pos.push_back(x - g0);
                len.push_back(y - x);
            }
        }
        err.set(npge_anchor_dist_upload_range(
            m->ctx[r], anchor_size, bases, uint32_t(len.size()), src_off.data(), dst_pos.data(),
            len.data(), g0, n_own, have_end - g0, gstart.data(), gcaller.data(), n_global));
    });
    if (err.failed()) return fail(err.code, err.text);
    for (int r = 0; r < G; r++) {
        int rc = npge_anchor_dist_set_chunks(m->ctx[r], m->chunks);
        if (rc) return rc;
    }
    // room for an uneven split of the job's windows; grown on demand
    const uint64_t need = total / uint64_t(G) * 5 / 4 + 65536;
    if (need > m->capacity) {
        int rc = peer_setup(m, need);
        if (rc) return rc;
    }
    m->k = anchor_size;
    m->uploaded = true;
    return NPGE_ANCHOR_OK;
}

// ---- find: the stages of every rank between in-process rendezvous ---------------------
extern "C" int npge_anchor_multi_find_device(npge_anchor_multi* m, const npge_anchor_params* p,
                                             const uint64_t* used_sorted, uint64_t n_used,
                                             npge_anchor_stats* job_stats) {
    if (!m || !p) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    if (!m->uploaded) return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_multi_upload has not been called");
    if (p->anchor_size != m->k) {
        return fail(NPGE_ANCHOR_ERR_STATE, "anchor-size differs from the uploaded one");
    }
    const int G = int(m->ctx.size());
    const bool similar = p->anchor_similar != 0;
    const uint64_t limit = uint64_t(p->max_anchor_fragments);
    m->have_result = false;
    for (int attempt = 0; attempt < 2; attempt++) {
        const uint32_t hist_cap = uint32_t(kMaxChunks) << kMaxDigitBits;
        std::vector<uint32_t> hist_all(size_t(G) * hist_cap, 0);
        std::vector<uint32_t> n_bins(G, 0);
        std::vector<uint64_t> n_recv(size_t(G) * G, 0), totals(size_t(G) * 2, 0);
        std::vector<const void*> found(G, nullptr);
        Barrier bar(G);
        RankError err;
        bool regrow = false;
        run_ranks(m, [&](int r) {
            npge_anchor_ctx* c = m->ctx[r];
            // count A, then pass A (asynchronous) while the others arrive
            std::vector<uint32_t> mine(hist_cap, 0);
            int rc = npge_anchor_dist_count_a(c, p, mine.data(), hist_cap, &n_bins[r]);
            if (!rc) rc = npge_anchor_dist_scatter_local(c, p);
            err.set(rc);
            bar.wait();  // (also: nobody writes into a buffer whose owner is a step behind)
            const uint32_t per = n_bins[0] * uint32_t(m->chunks);  // counts per rank
            if (!err.failed()) {
                std::copy(mine.begin(), mine.begin() + per, hist_all.begin() + size_t(r) * per);
            }
            bar.wait();
            if (!err.failed()) {
                rc = npge_anchor_dist_exchange(c, p, hist_all.data(), uint32_t(G), uint32_t(r),
                                               &n_recv[size_t(r) * G]);
                if (rc == NPGE_ANCHOR_ERR_CAPACITY && r == 0) regrow = true;
                err.set(rc);
            }
            bar.wait();  // every copy has landed
            uint64_t tot[2] = {0, 0};
            if (!err.failed()) err.set(npge_anchor_dist_sort_planned(c, p, 0, used_sorted, n_used, tot));
            if (similar) {
                found[r] = c->found_bits.ptr;
                bar.wait();  // every rank has set its found bits
                if (!err.failed()) err.set(npge_anchor_dist_found_or(c, uint32_t(G), found.data()));
                bar.wait();  // every rank has read them
                if (!err.failed()) {
                    err.set(npge_anchor_dist_found_swap(c));
                    err.set(npge_anchor_dist_sort_planned(c, p, 1, used_sorted, n_used, tot));
                }
            }
            totals[2 * r] = tot[0];
            totals[2 * r + 1] = tot[1];
            bar.wait();
            // share of max-anchor-fragments: the ranks own ascending hash ranges
            std::vector<uint64_t> frag_tot(G);
            for (int x = 0; x < G; x++) frag_tot[x] = totals[2 * x];
            if (!err.failed()) {
                err.set(npge_anchor_dist_emit(c, p, dist_local_limit(frag_tot.data(), uint32_t(r), limit),
                                              &m->stats[r]));
            }
        });
        if (regrow && attempt == 0) {  // the same answer on every rank: grow and redo the step
            uint64_t need = 0;
            for (int d = 0; d < G; d++) need = std::max(need, n_recv[d]);
            int rc = peer_setup(m, need * 9 / 8 + 65536);
            if (rc) return rc;
            continue;
        }
        if (err.failed()) return fail(err.code, err.text);
        break;
    }
    // the job's totals: the ranks' results are slices of the one-GPU result
    npge_anchor_stats& j = m->job;
    memset(&j, 0, sizeof(j));
    for (int r = 0; r < G; r++) {
        const npge_anchor_stats& s = m->stats[r];
        j.n_positions += s.n_positions;
        j.n_windows += s.n_windows;
        j.n_fragments += s.n_fragments;
        j.n_blocks += s.n_blocks;
        j.digest[0] += s.digest[0];
        j.digest[1] += s.digest[1];
        j.gpu_launches += s.gpu_launches;
        j.bytes_algorithmic += s.bytes_algorithmic;
        j.bytes_moved += s.bytes_moved;
        j.radix_passes = s.radix_passes;
        j.ms_hist = std::max(j.ms_hist, s.ms_hist);
        j.ms_sort = std::max(j.ms_sort, s.ms_sort);
        j.ms_group = std::max(j.ms_group, s.ms_group);
        j.ms_total = std::max(j.ms_total, s.ms_total);
        for (int i = 0; i < 16; i++) j.ms_pass[i] = std::max(j.ms_pass[i], s.ms_pass[i]);
    }
    m->last = *p;
    m->have_result = true;
    if (job_stats) *job_stats = j;
    return NPGE_ANCHOR_OK;
}

// ---- fetch: the ranks' slices (ascending hash ranges) joined into one result ---------
namespace {

struct MultiResult {
    std::vector<npge_anchor_result> parts;
    std::vector<uint64_t> block_offsets, new_used;
    // block_hash / fragment arrays: views into the largest part when there is one
    // rank, else joined copies
    std::vector<uint64_t> block_hash;
    std::vector<uint32_t> frag_seq, frag_min_pos;
    std::vector<int8_t> frag_ori;
    ~MultiResult() {
        for (npge_anchor_result& r : parts) npge_anchor_result_free(&r);
    }
};

}  // namespace

extern "C" int npge_anchor_multi_fetch(npge_anchor_multi* m, npge_anchor_result* out) {
    if (!m || !out) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    memset(out, 0, sizeof(*out));
    if (!m->have_result) return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_multi_find_device has not run");
    const int G = int(m->ctx.size());
    MultiResult* res = new (std::nothrow) MultiResult;
    if (!res) return fail(NPGE_ANCHOR_ERR_MEMORY, "out of host memory");
    res->parts.resize(G);
    for (auto& r : res->parts) memset(&r, 0, sizeof(r));
    RankError err;
    run_ranks(m, [&](int r) { err.set(npge_anchor_fetch(m->ctx[r], &res->parts[r])); });
    if (err.failed()) {
        delete res;
        return fail(err.code, err.text);
    }
    uint64_t nb = 0, nf = 0, nu = 0;
    for (const npge_anchor_result& r : res->parts) {
        nb += r.n_blocks;
        nf += r.n_fragments;
        nu += r.n_new_used;
    }
    try {
        res->block_offsets.resize(nb + 1);
        res->block_hash.resize(nb);
        res->frag_seq.resize(nf);
        res->frag_min_pos.resize(nf);
        res->frag_ori.resize(nf);
        res->new_used.reserve(nu);
    } catch (std::bad_alloc&) {
        delete res;
        return fail(NPGE_ANCHOR_ERR_MEMORY, "out of host memory for the result");
    }
    std::vector<uint64_t> b0(G + 1, 0), f0(G + 1, 0);
    for (int r = 0; r < G; r++) {
        b0[r + 1] = b0[r] + res->parts[r].n_blocks;
        f0[r + 1] = f0[r] + res->parts[r].n_fragments;
    }
    run_ranks(m, [&](int r) {
        const npge_anchor_result& p = res->parts[r];
        for (uint64_t b = 0; b < p.n_blocks; b++) {
            res->block_offsets[b0[r] + b] = p.block_offsets[b] + f0[r];
        }
        if (p.n_blocks) memcpy(&res->block_hash[b0[r]], p.block_hash, p.n_blocks * 8);
        if (p.n_fragments) {
            memcpy(&res->frag_seq[f0[r]], p.frag_seq, p.n_fragments * 4);
            memcpy(&res->frag_min_pos[f0[r]], p.frag_min_pos, p.n_fragments * 4);
            memcpy(&res->frag_ori[f0[r]], p.frag_ori, p.n_fragments);
        }
    });
    res->block_offsets[nb] = nf;
    // every rank lists EVERY hash of its range; the job's list drops its first one
    for (const npge_anchor_result& r : res->parts) {
        res->new_used.insert(res->new_used.end(), r.new_used, r.new_used + r.n_new_used);
    }
    if (!res->new_used.empty()) res->new_used.erase(res->new_used.begin());
    out->n_blocks = nb;
    out->n_fragments = nf;
    out->block_offsets = res->block_offsets.data();
    out->block_hash = res->block_hash.data();
    out->frag_seq = res->frag_seq.data();
    out->frag_min_pos = res->frag_min_pos.data();
    out->frag_ori = res->frag_ori.data();
    out->n_new_used = res->new_used.size();
    out->new_used = res->new_used.data();
    for (const npge_anchor_result& r : res->parts) {
        out->digest[0] += r.digest[0];
        out->digest[1] += r.digest[1];
    }
    out->owner_ = res;
    return NPGE_ANCHOR_OK;
}

extern "C" void npge_anchor_multi_result_free(npge_anchor_result* out) {
    if (!out) return;
    delete static_cast<MultiResult*>(out->owner_);
    memset(out, 0, sizeof(*out));
}

// ---- the compact layout over several GPUs ----------------------------------------------
extern "C" int npge_anchor_multi_set_result_layout(npge_anchor_multi* m, int32_t layout) {
    if (!m) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    for (npge_anchor_ctx* c : m->ctx) {
        const int rc = npge_anchor_set_result_layout(c, layout);
        if (rc) return rc;
    }
    return NPGE_ANCHOR_OK;
}

namespace {

struct MultiCompact {
    uint32_t* pos_ori = nullptr;
    uint64_t* block_off = nullptr;
    uint64_t* block_hash = nullptr;
    size_t bytes[3] = {0, 0, 0};
    std::vector<uint64_t> new_used, layout_start;
    std::vector<uint32_t> layout_seq;
    ~MultiCompact() {
        pinned_give(pos_ori, bytes[0]);
        pinned_give(block_off, bytes[1]);
        pinned_give(block_hash, bytes[2]);
    }
};

}  // namespace

// Every rank copies its slice straight into its place in the job's arrays (positions
// are job-wide: the slices need no translation, only the block offsets a shift).
extern "C" int npge_anchor_multi_fetch_compact(npge_anchor_multi* m,
                                               npge_anchor_compact_result* out) {
    if (!m || !out) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    memset(out, 0, sizeof(*out));
    if (!m->have_result) return fail(NPGE_ANCHOR_ERR_STATE, "npge_anchor_multi_find_device has not run");
    const int G = int(m->ctx.size());
    for (npge_anchor_ctx* c : m->ctx) {
        if (!c->have_pay) {
            return fail(NPGE_ANCHOR_ERR_STATE, "the resident result holds no payloads: "
                        "npge_anchor_multi_set_result_layout(NPGE_ANCHOR_LAYOUT_COMPACT) first");
        }
    }
    std::vector<uint64_t> b0(G + 1, 0), f0(G + 1, 0);
    for (int r = 0; r < G; r++) {
        b0[r + 1] = b0[r] + m->ctx[r]->stats.n_blocks;
        f0[r + 1] = f0[r] + m->ctx[r]->stats.n_fragments;
    }
    const uint64_t nb = b0[G], nf = f0[G];
    MultiCompact* res = new (std::nothrow) MultiCompact;
    if (!res) return fail(NPGE_ANCHOR_ERR_MEMORY, "out of host memory");
    try {
        res->bytes[0] = (nf + 1) * 4;
        res->pos_ori = static_cast<uint32_t*>(pinned_take(res->bytes[0]));
        res->bytes[1] = (nb + 1) * 8;
        res->block_off = static_cast<uint64_t*>(pinned_take(res->bytes[1]));
        res->bytes[2] = (nb + 1) * 8;
        res->block_hash = static_cast<uint64_t*>(pinned_take(res->bytes[2]));
    } catch (std::bad_alloc&) {
        delete res;
        return fail(NPGE_ANCHOR_ERR_MEMORY, "out of host memory for the result");
    }
    RankError err;
    run_ranks(m, [&](int r) {
        const int rc = fetch_compact_into(m->ctx[r], res->pos_ori + f0[r], res->block_off + b0[r],
                                          res->block_hash + b0[r]);
        err.set(rc);
        if (!rc && f0[r]) {
            for (uint64_t b = b0[r]; b < b0[r + 1]; b++) res->block_off[b] += f0[r];
        }
    });
    if (err.failed()) {
        delete res;
        return fail(err.code, err.text);
    }
    res->block_off[nb] = nf;
    // the job's list: every hash of every rank's range (+ a rank's lone cut tail,
    // AnchorFinder.cpp:385) but the job's first one
    bool lone = false;
    for (npge_anchor_ctx* c : m->ctx) lone = lone || c->lone_tail;
    if (lone) {
        try {
            for (int r = 0; r < G; r++) {
                res->new_used.insert(res->new_used.end(), res->block_hash + b0[r],
                                     res->block_hash + b0[r + 1]);
                if (m->ctx[r]->lone_tail) res->new_used.push_back(m->ctx[r]->lone_hash);
            }
        } catch (std::bad_alloc&) {
            delete res;
            return fail(NPGE_ANCHOR_ERR_MEMORY, "out of host memory for the result");
        }
        if (!res->new_used.empty()) res->new_used.erase(res->new_used.begin());
        out->new_used = res->new_used.data();
        out->n_new_used = res->new_used.size();
    } else if (nb > 1) {
        out->new_used = res->block_hash + 1;  // the list IS the tail of block_hash
        out->n_new_used = nb - 1;
    }
    const auto& seqs = m->ctx[0]->gseqs;
    res->layout_start.resize(seqs.size() + 1);
    res->layout_seq.resize(seqs.size() + 1);
    uint64_t end = 0;
    for (size_t i = 0; i < seqs.size(); i++) {
        res->layout_start[i] = seqs[i].start;
        res->layout_seq[i] = seqs[i].caller_index;
        end = seqs[i].start + seqs[i].size + 1;
    }
    res->layout_start[seqs.size()] = end;
    out->n_blocks = nb;
    out->n_fragments = nf;
    out->block_offsets = res->block_off;
    out->block_hash = res->block_hash;
    out->frag_pos_ori = res->pos_ori;
    out->n_layout = uint32_t(seqs.size());
    out->layout_start = res->layout_start.data();
    out->layout_seq = res->layout_seq.data();
    out->anchor_size = m->k;
    for (npge_anchor_ctx* c : m->ctx) {
        out->digest[0] += c->stats.digest[0];
        out->digest[1] += c->stats.digest[1];
    }
    out->owner_ = res;
    return NPGE_ANCHOR_OK;
}

extern "C" void npge_anchor_multi_compact_result_free(npge_anchor_compact_result* out) {
    if (!out) return;
    delete static_cast<MultiCompact*>(out->owner_);
    memset(out, 0, sizeof(*out));
}

extern "C" int npge_anchor_multi_find(npge_anchor_multi* m, const npge_anchor_params* p,
                                      const uint8_t* bases, const uint64_t* seq_offsets,
                                      uint32_t n_seqs, const uint64_t* used_sorted,
                                      uint64_t n_used, npge_anchor_result* out) {
    if (!out) return fail(NPGE_ANCHOR_ERR_PARAMS, "out is NULL");
    memset(out, 0, sizeof(*out));
    if (!p) return fail(NPGE_ANCHOR_ERR_PARAMS, "params is NULL");
    int rc = npge_anchor_multi_upload(m, p->anchor_size, bases, seq_offsets, n_seqs);
    if (!rc) rc = npge_anchor_multi_find_device(m, p, used_sorted, n_used, nullptr);
    if (!rc) rc = npge_anchor_multi_fetch(m, out);
    return rc;
}

extern "C" int npge_anchor_multi_get_stats(npge_anchor_multi* m, int rank,
                                           npge_anchor_stats* stats) {
    if (!m || !stats) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    if (!m->have_result) return fail(NPGE_ANCHOR_ERR_STATE, "no result yet");
    if (rank < 0) {
        *stats = m->job;
    } else if (rank < int(m->stats.size())) {
        *stats = m->stats[rank];
    } else {
        return fail(NPGE_ANCHOR_ERR_PARAMS, "rank out of range");
    }
    return NPGE_ANCHOR_OK;
}

// Guard bytes of every rank's device buffers (npge_anchor_debug_check), summed.
extern "C" int npge_anchor_multi_debug_check(npge_anchor_multi* m, uint32_t* n_damaged) {
    if (!m || !n_damaged) return fail(NPGE_ANCHOR_ERR_PARAMS, "NULL argument");
    *n_damaged = 0;
    for (npge_anchor_ctx* c : m->ctx) {
        uint32_t n = 0;
        const int rc = npge_anchor_debug_check(c, &n);
        if (rc) return rc;
        *n_damaged += n;
    }
    return NPGE_ANCHOR_OK;
}
