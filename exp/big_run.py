"""pangenome1g (C5) on one B200: throughput and self-consistency at full size."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
import torch
from npge_b200 import synth
from npge_b200.engine import AnchorEngine, concat

scale = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
ks = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [20]
t0 = time.perf_counter()
names, seqs = synth.generate("pangenome1g", scale=scale)
bases, offsets = concat(seqs)
print("generated %.1f Mbp, %d sequences in %.1f s" % (len(bases) / 1e6, len(seqs), time.perf_counter() - t0), flush=True)
eng = AnchorEngine(0)
BIG = 2**62
for k in ks:
    t0 = time.perf_counter(); eng.upload((bases, offsets), k); t_up = time.perf_counter() - t0
    for similar, mf in ((False, BIG), (True, 100000)):
        st = None
        for it in range(3):
            st = eng.find_device(k, similar=similar, max_frags=mf)
        print("k=%d similar=%s: %.2f ms/step = %.1f Gbp/s; windows %d fragments %d blocks %d digest %016x; stages %s (upload %.0f ms)"
              % (k, similar, st["ms_total"], st["n_positions"] / st["ms_total"] / 1e6, st["n_windows"],
                 st["n_fragments"], st["n_blocks"], st["digest"][0],
                 [round(st["ms_hist"], 2)] + [round(x, 2) for x in st["ms_pass"]], 1e3 * t_up), flush=True)
print("peak device memory: %.1f GB" % ((torch.cuda.mem_get_info()[1] - torch.cuda.mem_get_info()[0]) / 1e9))
