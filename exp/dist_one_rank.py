"""Stage times of the multi-GPU code path with ONE rank against the one-GPU path on the
same input: separates what the path itself costs from what several GPUs cost."""
import sys
import numpy as np
from npge_b200 import synth
from npge_b200.engine import AnchorEngine, MultiEngine, concat

BIG = 2 ** 62
names, seqs = synth.generate("brucella17")
bo = concat(seqs)
e = AnchorEngine(0)
e.upload(bo, 20)
for _ in range(3):
    st = e.find_device(20, similar=False, max_frags=BIG)
print("one GPU   hist %.3f pass %s total %.3f" % (st["ms_hist"], [round(x, 3) for x in st["ms_pass"][:6]], st["ms_total"]))
e.close()
for ranks in (1, 2):
    m = MultiEngine([0] * ranks)
    m.upload(bo, 20)
    for _ in range(3):
        m.find_device(20, similar=False, max_frags=BIG)
    for r in range(ranks):
        s = m.rank_stats(r)
        print("multi %d/%d hist %.3f pass %s total %.3f records %d" % (
            r, ranks, s["ms_hist"], [round(x, 3) for x in s["ms_pass"][:7]], s["ms_total"], s["n_windows"]))
    m.close()
