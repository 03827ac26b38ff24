"""Static loop analysis of a built library (no GPU needed): per kernel the instruction count, the largest backward-branch
loops and the opcode mix of the main loop, from `cuobjdump -sass`.

    python profiles/sass_loops.py codesign-kernels_b200/lib/libmpdata_b200.so [name filter] > profiles/<tag>_static.txt
"""
import collections
import re
import subprocess
import sys


def main():
    lib = sys.argv[1]
    flt = sys.argv[2] if len(sys.argv) > 2 else ""
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    funcs, cur = {}, None
    for ln in sass.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            funcs[cur] = []
            continue
        if cur is not None:
            funcs[cur].append(ln)
    for name, lines in funcs.items():
        if flt not in name:
            continue
        ins = []
        for ln in lines:
            m = re.match(r"\s*/\*([0-9a-f]{4,5})\*/\s+(.*?);", ln)
            if m:
                ins.append((int(m.group(1), 16), m.group(2).strip()))
        loops = []
        for a, t in ins:
            m = re.search(r"BRA(?:\.\S+)?\s+(?:\S+,\s*)?0x([0-9a-f]+)", t)
            if m and int(m.group(1), 16) <= a:
                tgt = int(m.group(1), 16)
                loops.append((tgt, a, (a - tgt) // 16 + 1))
        demangled = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip() or name
        print(f"{demangled}: {len(ins)} instructions")
        for lo, hi, n in sorted(loops, key=lambda x: -x[2])[:5]:
            print(f"    loop {lo:#x}..{hi:#x}: {n} instructions")
        if loops:
            lo, hi, n = max(loops, key=lambda x: x[2])
            body = [t for a, t in ins if lo <= a <= hi]
            ops = collections.Counter((t.split()[1] if t.startswith("@") else t.split()[0]).split(".")[0] for t in body)
            fp64 = sum(v for k, v in ops.items() if k in ("DADD", "DMUL", "DFMA", "DSETP", "MUFU"))
            integer = sum(v for k, v in ops.items() if k in ("IMAD", "IADD3", "VIADD", "LEA", "LOP3", "ISETP", "SHF", "IABS", "SEL", "FSEL", "MOV", "PRMT", "CS2R"))
            print(f"    largest loop: FP64-pipe {fp64} ({100 * fp64 / n:.0f} %), integer/select/move {integer} ({100 * integer / n:.0f} %), "
                  f"LDS+STS {ops['LDS'] + ops['STS']}, LDG+STG {ops['LDG'] + ops['STG']}, BAR {ops['BAR']}, WARPSYNC {ops['WARPSYNC']}")
            print("    mix: " + ", ".join(f"{k} {v}" for k, v in ops.most_common(16)))


if __name__ == "__main__":
    main()
