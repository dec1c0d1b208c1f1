"""Static SASS opcode histogram of the kernels in libb200attn.so (cuobjdump, no GPU needed): the proof that the tcgen05 / TMEM /
TMA instructions are really in the binary.    python tests/tools/sass_histogram.py [kernel-name-regex] > profiles/rNN_sass_histogram.md"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
SO = os.path.join(ROOT, "ht_vllm_omni_b200", "libb200attn.so")
KEY = ["UTCQMMA", "UTCHMMA", "UTCBAR", "UTCATOMSWS", "LDTM", "STTM", "UTMALDG", "UTMAPF", "UBLKCP", "SYNCS", "HMMA", "LDSM", "MOVM", "F2FP", "MUFU",
       "ACQBULK", "UCGABAR_ARV", "UCGABAR_WAIT", "FENCE", "MEMBAR", "STAS", "ATOMS", "REDG", "STS", "LDS", "LDG", "STG", "BAR", "NANOSLEEP"]


def main():
    pat = re.compile(sys.argv[1] if len(sys.argv) > 1 else ".")
    out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    fn, hist = None, {}
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
            hist[fn] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and fn:
            hist[fn][m.group(1)] += 1
    print("# Static SASS opcode histogram (cuobjdump -sass libb200attn.so; instruction COUNTS in the binary, not executions)\n")
    for fn, c in hist.items():
        if not pat.search(fn):
            continue
        tot = sum(c.values())
        print(f"## `{fn}` -- {tot} instructions")
        keys = [k for k in KEY if c.get(k)]
        print("key opcodes: " + ", ".join(f"{k} {c[k]}" for k in keys))
        print("top: " + ", ".join(f"{k} {v}" for k, v in c.most_common(14)) + "\n")


if __name__ == "__main__":
    main()
