"""Hot source lines of one kernel from an ncu report captured with
--import-source on (and -lineinfo at compile time).

usage: ncu_hot_lines.py REPORT.ncu-rep KERNEL_REGEX [N]"""
import csv
import io
import subprocess
import sys


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
    text = subprocess.run(
        ["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass",
         "--kernel-name", "regex:" + kern], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(text)))
    hdr, path, out = None, "", []
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            path = r[1].split("/")[-1]
        elif r[0] == "Line No":
            hdr = r
        elif hdr and r[0].isdigit():
            ii, si = hdr.index("Instructions Executed"), hdr.index("# Samples")
            num = lambda x: int(x) if x.isdigit() else 0
            out.append((num(r[ii]), num(r[si]), path, int(r[0]), r[1].strip()))
    tot_i = sum(o[0] for o in out) or 1
    tot_s = sum(o[1] for o in out) or 1
    print("%s: %d warp instructions, %d samples" % (kern, tot_i, tot_s))
    for n, s, p, ln, src in sorted(out, reverse=True)[:top]:
        print("%5.1f%% inst %5.1f%% stall  %s:%d  %s"
              % (100.0 * n / tot_i, 100.0 * s / tot_s, p, ln, src[:100]))


if __name__ == "__main__":
    main()
