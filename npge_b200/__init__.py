"""npge_b200 -- B200-native AnchorFinder behind NPGE's Processor API.

Only what the hot path needs: csrc/ (CUDA kernels + the C ABI of
include/npge_anchor.h), a ctypes driver (engine), the host-side mirror of the
reference's Processor/BlockSet interface for this path (processor, model) and
the synthetic genome generator used by the benchmarks (synth)."""
