"""Static SASS opcode histogram per kernel of the built library (instruction-diet bookkeeping).
Usage: python -m tools.sass_hist <kernel-substring> [--top N]"""
import collections
import re
import subprocess
import sys

LIB = "lowlight_image_enhancement_b200/_lib/libllie_b200.so"


def main():
    pat = sys.argv[1]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 25
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    name, hist = None, collections.OrderedDict()
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            hist[name] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", line)
        if m and name:
            hist[name][m.group(1).split(".")[0]] += 1
    for k, h in hist.items():
        if pat in k:
            tot = sum(h.values())
            print(f"== {k[:150]}\n   total {tot}")
            print("   " + "  ".join(f"{op}:{n}" for op, n in h.most_common(top)))


if __name__ == "__main__":
    main()
