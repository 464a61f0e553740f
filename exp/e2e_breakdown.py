"""Where the end-to-end time of npge_anchor_find goes (host buffers in and out)."""
import sys, time
import numpy as np
import torch
sys.path.insert(0, ".")
from npge_b200 import synth
from npge_b200.engine import AnchorEngine, concat

names, seqs = synth.generate("brucella17")
bases, offsets = concat(seqs)
pinned = torch.from_numpy(bases).pin_memory().numpy()
eng = AnchorEngine(0)
BIG = 2**62
for similar, mf in ((False, BIG), (True, 100000)):
    for it in range(4):
        t0 = time.perf_counter(); eng.upload((pinned, offsets), 20)
        t1 = time.perf_counter(); st = eng.find_device(20, similar=similar, max_frags=mf)
        t2 = time.perf_counter(); r = eng.fetch_counts()
        t3 = time.perf_counter(); c = eng.find_counts((pinned, offsets), 20, similar=similar, max_frags=mf)
        t4 = time.perf_counter()
    print("similar=%s upload %.2f ms, find_device %.2f ms (kernels %.2f), fetch %.2f ms (%.0f MB), one-call find %.2f ms"
          % (similar, 1e3*(t1-t0), 1e3*(t2-t1), st["ms_total"], 1e3*(t3-t2), r["result_bytes"]/1e6, 1e3*(t4-t3)))
