"""Stage times of the one-GPU step under forced digit plans (npge_anchor_debug_plan):
does a finer (A, B) split pay through better filled sort units?"""
import sys
from npge_b200 import synth
from npge_b200.engine import AnchorEngine, concat

BIG = 2 ** 62
name = sys.argv[1] if len(sys.argv) > 1 else "brucella17"
names, seqs = synth.generate(name)
bo = concat(seqs)
e = AnchorEngine(0)
PLANS = [tuple(int(v) for v in a.split(',')) for a in sys.argv[2:]] or [(0, 0), (17, 9), (18, 9)]
for f, wa in PLANS:
    e.debug_plan(f, wa)
    e.upload(bo, 20)
    best = None
    for _ in range(5):
        st = e.find_device(20, similar=False, max_frags=BIG)
        if best is None or st["ms_total"] < best["ms_total"]:
            best = st
    print("f=%2d wa=%d  hist %.3f pass %s total %.3f  digest %x" % (
        f, wa, best["ms_hist"], [round(x, 3) for x in best["ms_pass"][:5]], best["ms_total"],
        best["digest"][0]), flush=True)
e.close()
