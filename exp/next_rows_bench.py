"""Timing of the SURVEY 8f rows: device ingest and FragmentFinder (brucella17)."""
import sys, time
import numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "oracle")
import torch
from npge_b200 import synth, fasta
from npge_b200.engine import AnchorEngine, concat
import oracle

names, seqs = synth.generate("brucella17")
bases, offsets = concat(seqs)
# FASTA text: 70-column lines
parts = []
for n, s in zip(names, seqs):
    body = np.insert(s, np.arange(70, len(s), 70), ord("\n"))
    parts.append(np.frombuffer((">%s\n" % n).encode(), dtype=np.uint8))
    parts.append(body)
    parts.append(np.frombuffer(b"\n", dtype=np.uint8))
text = np.concatenate(parts)
t0 = time.perf_counter(); nm, _, raw, roff = fasta.split_records(text); t_split = time.perf_counter() - t0
praw = torch.from_numpy(raw).pin_memory().numpy()
pclean = torch.from_numpy(bases).pin_memory().numpy()
eng = AnchorEngine(0)
for it in range(3):
    t0 = time.perf_counter(); eng.upload((pclean, offsets), 20); t_up = time.perf_counter() - t0
    t0 = time.perf_counter(); sizes = eng.upload_raw((praw, roff), 12); t_raw = time.perf_counter() - t0
assert [int(x) for x in sizes] == [len(s) for s in seqs]
print("FASTA %.1f MB: split_records %.1f ms (host), upload_raw %.2f ms (%.1f GB/s of text), clean upload %.2f ms"
      % (len(text) / 1e6, 1e3 * t_split, 1e3 * t_raw, len(raw) / t_raw / 1e9, 1e3 * t_up))
t0 = time.perf_counter(); c = oracle.to_atgcn(bytes(raw[:20_000_000]).decode()); t_cpu = time.perf_counter() - t0
print("CPU to_atgcn (oracle, 1 thread): %.1f ms per 20 MB = %.2f GB/s" % (1e3 * t_cpu, 20e6 / t_cpu / 1e9))
texts = None
for L in (12, 20, 40, 200, 2000):
    pat = bytes(seqs[3][100000:100000 + L]).decode()
    for it in range(3):
        t0 = time.perf_counter(); got, total = eng.find_fragments_device(pat, 1000, with_total=True); t = time.perf_counter() - t0
    print("FragmentFinder L=%d: %.3f ms per search of 56.2 Mbp (%.1f Gbp/s), %d matches" % (L, 1e3 * t, 56.2e6 / t / 1e9, total))
small = [bytes(s).decode() for s in seqs[:4]]
pat = small[3][100000:100020]
t0 = time.perf_counter(); w = oracle.fragment_find(small, pat, 1000); t = time.perf_counter() - t0
print("CPU FragmentFinder restatement (1 thread): %.1f ms per %.1f Mbp = %.4f Gbp/s" % (1e3 * t, sum(map(len, small)) / 1e6, sum(map(len, small)) / t / 1e9))
