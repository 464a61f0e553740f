"""The digit plan (make_msd_plan, npge_b200/csrc/msd_kernels.cu) through its device-less
test hook: how the key bits are split between the two partition passes and the
shared-memory sort.  The bucket sort costs a fixed time per unit of 4096 records and
units are whole (A, B) bins, so the plan keeps the bins small (profiles/r02_summary.md,
"The digit plan"); these cases pin the rule the measurements chose."""
import ctypes as C

import pytest

from npge_b200 import _abi


def plan(k, n):
    lib = _abi.load()
    a, b, p = C.c_int32(), C.c_int32(), C.c_int32()
    rc = lib.npge_anchor_debug_plan_of(k, n, C.byref(a), C.byref(b), C.byref(p))
    assert rc == _abi.OK, lib.npge_anchor_last_error().decode()
    return a.value, b.value, p.value


@pytest.mark.parametrize("n, want", [
    (9_900_000, (8, 8)),          # C1 brucella3
    (56_140_046, (9, 9)),         # C2 brucella17: bins of ~214 records
    (112_293_299, (9, 9)),        # 2 GPUs x brucella17
    (224_585_656, (10, 9)),       # 4 GPUs: bins below 512 records need 19 bits
    (449_169_209, (10, 10)),      # 8 GPUs
    (1_006_000_000, (10, 10)),    # C5: rows are capped at 1024 bins
])
def test_plan_of_the_benchmarked_jobs(n, want):
    wa, wb, passes = plan(20, n)
    assert (wa, wb) == want or (wa, wb) == want[::-1] and wa <= 9
    assert passes == (40 - wa + 7) // 8


def test_larger_half_goes_to_pass_a_while_its_rows_stay_at_512_bins():
    assert plan(20, 30_000_000)[:2] == (9, 8)      # 17 bits
    assert plan(20, 224_585_656)[:2] == (9, 10)    # 19 bits: 10 + 9 would cost pass A more


def test_plan_invariants():
    for k in range(1, 33):
        for n in (0, 1, 100, 5000, 10 ** 6, 10 ** 8, 2 ** 30 - 1):
            wa, wb, passes = plan(k, n)
            assert 1 <= wa <= 10 and 1 <= wb <= 10
            assert wa + wb <= max(2 * k, 2)
            if 0 < 2 * k - 32 <= 10:   # k = 17..21: pass A takes what an 8-byte record
                assert wa >= 2 * k - 32  # cannot hold (32 key bits); above: 12-byte records
            assert passes == (2 * k - wa + 7) // 8


def test_plan_rejects_a_bad_anchor_size():
    lib = _abi.load()
    assert lib.npge_anchor_debug_plan_of(0, 1000, None, None, None) == _abi.ERR_PARAMS
    assert lib.npge_anchor_debug_plan_of(33, 1000, None, None, None) == _abi.ERR_PARAMS
