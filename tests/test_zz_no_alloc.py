"""npge_fragment_find_device_noalloc (no counterpart in the reference):
the FragmentFinder search with its scratch in a fixed
file-scope device array and caller-owned device outputs, so that the library allocates
nothing.  Its output must equal npge_fragment_find_device's, exactly (integer work), on
calls in a row that share the scratch array.  Collected last on purpose (test_zz_n...)."""
import numpy as np
import pytest

from npge_b200 import _abi, synth
from npge_b200.engine import AnchorEngine, AnchorError

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    e = AnchorEngine(0)
    yield e
    assert e.debug_check() == 0   # no kernel stored past the end of a device buffer
    e.close()


def run_noalloc(eng, pattern, max_matches, capacity):
    import torch
    dev = torch.device("cuda", 0)
    # the CALLER's buffers; poisoned so that stale contents cannot pass for results
    seq = torch.full((max(capacity, 1),), 0x7fffffff, dtype=torch.int32, device=dev)
    pos = torch.full((max(capacity, 1),), 0x7fffffff, dtype=torch.int32, device=dev)
    ori = torch.full((max(capacity, 1),), 99, dtype=torch.int8, device=dev)
    torch.cuda.synchronize()
    n, total = eng.find_fragments_device_noalloc(
        pattern, max_matches, seq.data_ptr(), pos.data_ptr(), ori.data_ptr(), capacity)
    s, p, o = seq.cpu().numpy(), pos.cpu().numpy(), ori.cpu().numpy()
    assert np.all(s[n:] == 0x7fffffff) and np.all(o[n:] == 99)   # nothing past n written
    return list(zip(s[:n].tolist(), p[:n].tolist(), o[:n].tolist())), total


def test_noalloc_equals_the_allocating_entry_point_on_calls_in_a_row(eng):
    names, seqs = synth.generate("small")
    text = bytes(seqs[0]).decode()
    eng.upload(seqs, 12)
    patterns = [text[5000:5020],            # a 20-mer of strain 0: hits in every strain
                text[100:112],              # shorter pattern, other tile counts
                text[70000:70000 + 75],     # longer than 32 bases (three chunks)
                "ACGTACGTACGTA",            # probably absent
                text[5000:5020]]            # the first one again, after other contents
    for round_ in range(2):                 # two passes: the scratch keeps old contents
        for pat in patterns:
            want, want_total = eng.find_fragments_device(pat, max_matches=1000, with_total=True)
            got, total = run_noalloc(eng, pat, 1000, 1000)
            assert total == want_total
            assert got == want
    # the cut: max_matches below the number of matches, and capacity below max_matches
    pat = text[100:112]
    want_all, total = eng.find_fragments_device(pat, max_matches=10 ** 6, with_total=True)
    assert total >= 3
    got, t = run_noalloc(eng, pat, 2, 1000)
    assert got == want_all[:2] and t == total
    got, t = run_noalloc(eng, pat, 1000, 1)
    assert got == want_all[:1] and t == total
    # a pattern with N matches N only, as in the allocating entry point
    with_n = synth.generate("tiny")[1]
    eng.upload(with_n, 8)
    flat = bytes(np.concatenate(with_n)).decode()
    at = flat.index("N")
    pat = flat[max(at - 6, 0):at + 4]
    want = eng.find_fragments_device(pat, max_matches=100)
    got, _ = run_noalloc(eng, pat, 100, 100)
    assert got == want and len(want) >= 1


def test_noalloc_limits_and_errors(eng):
    names, seqs = synth.generate("tiny")
    eng.upload(seqs, 8)
    with pytest.raises(AnchorError) as e:   # pattern beyond the fixed maximum
        run_noalloc(eng, "ACGT" * 257, 10, 10)
    assert e.value.code == _abi.ERR_CAPACITY and "MAX_PATTERN" in str(e.value)
    with pytest.raises(AnchorError) as e:
        run_noalloc(eng, "", 10, 10)
    assert e.value.code == _abi.ERR_PARAMS
    with pytest.raises(AnchorError) as e:
        run_noalloc(eng, "ACGTRACGT", 10, 10)       # not to_atgcn-clean
    assert e.value.code == _abi.ERR_INPUT
    # more positions than the scratch array covers: refused, clearly
    big = np.full((1 << 26) + 5000, ord("A"), dtype=np.uint8)
    eng.upload([big], 8)
    with pytest.raises(AnchorError) as e:
        run_noalloc(eng, "ACGTACGTAC", 10, 10)
    assert e.value.code == _abi.ERR_CAPACITY and "MAX_POSITIONS" in str(e.value)
    # and the engine still works afterwards
    eng.upload(seqs, 8)
    pat = bytes(seqs[0][40:60]).decode()
    want = eng.find_fragments_device(pat, max_matches=50)
    got, _ = run_noalloc(eng, pat, 50, 50)
    assert got == want and len(want) >= 1
