"""Per-kernel SASS opcode histogram of a built library (cuobjdump -sass), the Blackwell evidence kept in the tree:
UTMALDG / UTMASTG (TMA tile loads / stores), SYNCS (mbarrier), DFMA / DMUL / DADD (FP64 pipe), LDS / STS, LDG / STG,
LDGSTS (cp.async), plus the total instruction count.  usage: python profiles/sass_hist.py <lib.so> <out.txt>"""
import collections
import re
import subprocess
import sys

WATCH = ("UTMALDG", "UTMASTG", "UTMAPF", "SYNCS", "DFMA", "DMUL", "DADD", "LDS", "STS", "LDG", "STG", "LDGSTS", "SHFL", "BAR", "ATOM", "RED", "MUFU")


def main():
    lib, out = sys.argv[1], sys.argv[2]
    txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    arch = sorted(set(re.findall(r"arch = (sm_\w+)", txt)))
    kernels, cur = collections.OrderedDict(), None
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)((?:\.[A-Z0-9_]+)*)", line)
        if m and cur:
            kernels[cur][m.group(1)] += 1
            kernels[cur]["_total"] += 1
    demangle = subprocess.run(["c++filt"] + list(kernels), capture_output=True, text=True).stdout.splitlines()
    with open(out, "w") as f:
        f.write(f"# {lib}: arch {', '.join(arch)}; SASS opcode counts per kernel (static instruction counts, cuobjdump -sass)\n")
        f.write("kernel\ttotal\t" + "\t".join(WATCH) + "\n")
        for (name, c), dn in zip(kernels.items(), demangle):
            short = re.sub(r"\(.*", "", dn)
            f.write(short + "\t" + str(c["_total"]) + "\t" + "\t".join(str(c[w]) for w in WATCH) + "\n")
    print(open(out).read())


if __name__ == "__main__":
    main()
