"""Writes the SASS of the hot kernels (the instances brucella17 at k=20 runs)
to profiles/sass/*.sass: cuobjdump -sass of the built library, mnemonic lines
only.  No GPU needed.  usage: sass_extract.py [LIB]"""
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "..", "npge_b200", "libnpge_anchor.so")
OUT = os.path.join(HERE, "..", "profiles", "sass")
WANT = {
    "count_a_kernel": "count_a",
    "pass_a_kernelILi8ELb0ELb0": "pass_a_8_narrow_local",
    "pass_b_kernelILi8ELb0": "pass_b_8_narrow",
    "bucket_sort_kernelILb0ELi0": "bucket_sort_narrow_exact",
    "bucket_emit_kernelILb0ELi0": "bucket_emit_narrow_exact",
}


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else LIB
    txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    os.makedirs(OUT, exist_ok=True)
    for part in re.split(r"(?m)^\s*Function : ", txt)[1:]:
        name, body = part.split("\n", 1)
        for key, out in WANT.items():
            if key in name:
                lines = [re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l.rstrip())
                         for l in body.split("\n") if re.match(r"\s*/\*[0-9a-f]{4}\*/", l)]
                with open(os.path.join(OUT, out + ".sass"), "w") as f:
                    f.write("// %s\n" % name.strip() + "\n".join(lines) + "\n")
                print(out, len(lines))


if __name__ == "__main__":
    main()
