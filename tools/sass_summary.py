"""Per-kernel SASS evidence from the built library, without a GPU.

usage: sass_summary.py [LIB] > profiles/rNN_sass_summary.md

For every kernel of libnpge_anchor.so: registers, shared memory and spill bytes
(cuobjdump -res-usage) and the count of the SASS mnemonics that show how the
kernel moves data (TMA bulk copies, mbarrier waits, 128-bit global accesses,
match.any, funnel shifts, ballots, shuffles, shared-memory atomics)."""
import collections
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "..", "npge_b200", "libnpge_anchor.so")

GROUPS = [
    ("UBLKCP (TMA bulk)", r"^UBLKCP"),
    ("SYNCS (mbarrier)", r"^SYNCS"),
    ("LDG.128", r"^LDG\.E(\.\w+)*\.128"),
    ("LDG.64", r"^LDG\.E(\.\w+)*\.64"),
    ("LDG other", r"^LDG"),
    ("STG.128", r"^STG\.E(\.\w+)*\.128"),
    ("STG.64", r"^STG\.E(\.\w+)*\.64"),
    ("STG other", r"^STG"),
    ("LDS", r"^LDS"),
    ("STS", r"^STS"),
    ("ATOMS", r"^ATOMS"),
    ("ATOMG/RED", r"^(ATOMG|RED)"),
    ("MATCH.ANY", r"^MATCH"),
    ("VOTE", r"^VOTE"),
    ("SHFL", r"^SHFL"),
    ("SHF (funnel)", r"^SHF"),
    ("BREV", r"^BREV"),
    ("POPC", r"^POPC"),
    ("LDL/STL (spill)", r"^(LDL|STL)"),
    ("BAR", r"^BAR"),
]


def demangle(names):
    out = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout
    return out.strip().split("\n")


def short(name):
    name = re.sub(r"npge_b200::", "", name)
    name = re.sub(r"\(anonymous namespace\)::", "", name)
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"<unnamed>::", "", name)
    name = re.sub(r"\((bool|int|unsigned int)\)", "", name)
    return re.sub(r"\(.*$", "", name)


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else LIB
    res = subprocess.run(["cuobjdump", "-res-usage", lib], capture_output=True,
                         text=True).stdout
    usage = {}
    cur = None
    for line in res.split("\n"):
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            cur = m.group(1)
            continue
        if cur and "REG:" in line:
            f = dict(re.findall(r"(\w+):(\d+)", line))
            usage[cur] = f
            cur = None
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True,
                          text=True).stdout
    counts = collections.OrderedDict()
    total = {}
    cur = None
    for line in sass.split("\n"):
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            total[cur] = 0
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur:
            op = m.group(1)
            total[cur] += 1
            for label, pat in GROUPS:
                if re.match(pat, op):
                    counts[cur][label] += 1
                    break
    names = list(counts)
    pretty = dict(zip(names, demangle(names)))
    print("# SASS summary of libnpge_anchor.so (sm_100a), `tools/sass_summary.py`")
    print()
    print("Static instruction counts per kernel (not executed counts): what each kernel is")
    print("built from.  REG / SHARED (static bytes) / LOCAL (spill bytes per thread) from")
    print("`cuobjdump -res-usage`; dynamic shared memory is set at launch (DESIGN.md section 4).")
    print()
    labels = [g[0] for g in GROUPS]
    print("| kernel | SASS lines | REG | SHARED | LOCAL | " + " | ".join(labels) + " |")
    print("|---|---|---|---|---|" + "---|" * len(labels))
    for n in sorted(names, key=lambda x: short(pretty[x])):
        u = usage.get(n, {})
        row = [short(pretty[n]), str(total[n]), u.get("REG", "?"), u.get("SHARED", "?"),
               u.get("LOCAL", "?")] + [str(counts[n].get(l, 0) or "") for l in labels]
        print("| " + " | ".join(row) + " |")


if __name__ == "__main__":
    main()
