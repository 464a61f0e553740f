// gpu_anchor_finder.hpp -- npge::gpu::AnchorFinder, the drop-in for the
// reference's AnchorFinder processor (src/algo/AnchorFinder.hpp:20-51).
// Same options, same target blockset contract, same cross-run state
// (used_hashes_); the work happens in libnpge_anchor.so (include/npge_anchor.h).
#pragma once

#include <stdint.h>

#include <vector>

#ifdef NPGE_B200_REAL_HEADERS   // inside an NPGE checkout
#include "Processor.hpp"
#else
#include "npge_facade.hpp"
#endif

namespace npge {
namespace gpu {

class AnchorFinder : public Processor {
public:
    AnchorFinder();
    ~AnchorFinder();

    // the CUDA device of this instance (default 0); call before run()
    void set_device(int device);
    // several GPUs: positions are sharded over them, k-mers repartitioned by hash
    // range (npge_anchor_multi_*); one id may repeat (tests on one GPU)
    void set_devices(const std::vector<int>& devices);
    // how the last run() got its sequences onto the device(s): true = the packed
    // genome of the previous run was still valid (same sequences) and was reused
    bool last_run_reused_upload() const;
    // wall time of the parts of the last run_impl(), ms: [0] linearising + uploading the
    // sequences, [1] the device search + the copy of the result, [2] building the
    // Fragment / Block objects (and check_block), [3] the device search alone (in [1])
    const double* last_run_ms() const;

protected:
    void run_impl() const;          // Processor.hpp:482
    const char* name_impl() const;  // Processor.hpp:492

private:
    struct Impl;  // run_impl is const: the state lives behind a pointer, as in
    Impl* impl_;  // the reference (AnchorFinder.hpp:48-50)
};

// Drop-in for the reference's FragmentFinder (src/algo/FragmentFinder.hpp:
// "Locate fragment by its sequence"): options `pattern`, `max-matches`; adds one
// single-fragment block per occurrence of the pattern, either strand.
class FragmentFinder : public Processor {
public:
    FragmentFinder();
    ~FragmentFinder();
    void set_device(int device);

protected:
    void run_impl() const;
    const char* name_impl() const;

private:
    struct Impl;
    Impl* impl_;
};

// Drop-in for the reference's MoveUnchanged (src/algo/MoveUnchanged.hpp:
// "Move unchanged blocks from other to target"): remembers block_hash
// (src/model/block_hash.cpp:29-54) of every block of `other` it has seen; a
// block whose hash was seen in an earlier run moves to `target`.  AnchorLoop
// runs it right after AnchorFinder (src/algo/lua_lib.lua:719-721).  The hashes
// of all blocks are computed in one call on the device (npge_block_hashes).
class MoveUnchanged : public Processor {
public:
    MoveUnchanged();
    ~MoveUnchanged();
    void set_device(int device);
    void clear();  // forget the hashes (MoveUnchanged.cpp:24-26)
    // block_hash of the blocks of the last run, in the order of `other` before
    // the run (for tests and diagnostics)
    const std::vector<uint64_t>& last_hashes() const;

protected:
    void run_impl() const;
    const char* name_impl() const;

private:
    struct Impl;
    Impl* impl_;
};

// Meta registration: replaces the built-in factory (Meta.hpp:83-87), so every
// pipeline, Lua script and the CLI that asks for "AnchorFinder" gets this one.
template <typename MetaT>
inline void register_processors(MetaT* meta) {
    meta->template set_processor<AnchorFinder>("AnchorFinder", true);
    meta->template set_processor<FragmentFinder>("FragmentFinder", true);
    meta->template set_processor<MoveUnchanged>("MoveUnchanged", true);
}

}  // namespace gpu
}  // namespace npge
