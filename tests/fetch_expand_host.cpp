// Test harness: npge_b200/csrc/fetch_expand.h (host side of the compact fetch)
// behind a C entry point for tests/test_zz_fetch.py.  Test infrastructure only.
#include "../npge_b200/csrc/fetch_expand.h"

extern "C" void fetch_expand_host(const uint32_t* pay, uint64_t lo, uint64_t hi,
                                  const uint32_t* lut, int shift, const uint32_t* start,
                                  const uint32_t* caller, uint32_t* seq, uint32_t* pos,
                                  int8_t* ori) {
    npge_b200::expand_payloads(pay, lo, hi, lut, shift, start, caller, seq, pos, ori);
}
