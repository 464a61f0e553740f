"""Device ingest: raw FASTA text -> to_atgcn -> packed genome on
the GPU, against the oracle's to_atgcn (Sequence.cpp:151-179) + the clean path."""
import random

import numpy as np
import pytest

import oracle
from npge_b200 import fasta, synth

BIG = 2**62


def test_split_records():
    text = b">s1 first one\nACGT\nacgt\n>s2\nNNNN\n>empty\n>s4 x y\nAC-GT"
    names, desc, raw, offsets = fasta.split_records(text)
    assert names == ["s1", "s2", "empty", "s4"]
    assert desc == ["first one", "", "", "x y"]
    bodies = [bytes(raw[int(offsets[i]):int(offsets[i + 1])]) for i in range(4)]
    assert bodies == [b"ACGT\nacgt\n", b"NNNN\n", b"", b"AC-GT"]
    assert fasta.split_records(b"")[0] == []


def test_oracle_to_atgcn_golden():  # test-lua/sequence.lua:1-40
    assert oracle.to_atgcn("AT-G-C-A#T$N .C") == "ATGCATNC"
    assert oracle.to_atgcn("acgtRYMKWSBVHDn") == "ACGT" + "N" * 11


def dirty(rng, clean):
    """FASTA-like text whose to_atgcn is `clean`: line breaks, lower case, gaps."""
    out, col = [], 0
    for ch in clean:
        if rng.random() < 0.03:
            out.append(rng.choice("-. *0123456789#$xzXZ\r\t"))
        out.append(ch.lower() if rng.random() < 0.3 else ch)
        col += 1
        if col % 60 == 0:
            out.append("\n")
    return "".join(out) + rng.choice(["", "\n", "\n\n"])


@pytest.fixture(scope="module")
def eng():
    from npge_b200.engine import AnchorEngine
    e = AnchorEngine(0)
    yield e
    assert e.debug_check() == 0   # no kernel stored past the end of a device buffer
    e.close()


@pytest.mark.gpu
def test_gpu_upload_raw_equals_clean_upload(eng):
    rng = random.Random(3)
    for trial in range(25):
        clean = ["".join(rng.choice("ATGCATGCATGCN") for _ in range(rng.choice(
            [0, 3, 17, 100, 700, 4096, 9001]))) for _ in range(rng.randint(1, 7))]
        raws = [dirty(rng, c) for c in clean]
        if trial % 4 == 0:  # ambiguity codes become N
            raws = [r.replace("N", rng.choice("RYMKWSBVHDn")) for r in raws]
        assert [oracle.to_atgcn(r) for r in raws] == clean
        k = rng.choice([3, 8, 13, 20])
        sizes = eng.upload_raw(raws, k)
        assert [int(x) for x in sizes] == [len(c) for c in clean]
        eng.find_device(k, similar=False, max_frags=BIG)
        got = eng.fetch()
        want = oracle.find_ideal(clean, k, similar=False, max_frags=BIG)
        assert got.blocks() == want.blocks()
        assert list(got.block_hash) == list(want.block_hash)
        assert got.digest == want.digest


@pytest.mark.gpu
def test_gpu_fasta_file_at_scale(eng):
    names, seqs = synth.generate("small")
    rng = random.Random(9)
    lines = []
    for name, s in zip(names, seqs):
        text = bytes(s).decode()
        lines.append(">%s synthetic\n" % name)
        lines.extend(text[i:i + 70].lower() + "\n" for i in range(0, len(text), 70))
    data = "".join(lines).encode()
    got_names, _, raw, offsets = fasta.split_records(data)
    assert got_names == names
    sizes = eng.upload_raw((raw, offsets), 20)
    assert [int(x) for x in sizes] == [len(s) for s in seqs]
    st = eng.find_device(20, similar=True, max_frags=100000)
    got = eng.fetch()
    want = oracle.find_ideal(seqs, 20, similar=True, max_frags=100000)
    assert got.blocks() == want.blocks() and got.digest == want.digest
    # the device copy stays resident: FragmentFinder on the same upload
    pat = bytes(seqs[0][1000:1040]).decode()
    assert eng.find_fragments_device(pat, 50) == oracle.fragment_find(
        [bytes(s).decode() for s in seqs], pat, 50)
