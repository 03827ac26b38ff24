"""Attributes the per-instruction counters of an ncu report to source lines (run where ncu / nvdisasm / cuobjdump are).

    python profiles/ncu_by_line.py <report.ncu-rep> <lib.so> <kernel-name-substring> [--top 40]

ncu's `--page source --csv` lists one row per SASS instruction of the kernel (executed instructions, stall samples ...) but
not the source line; `nvdisasm -g` of the same cubin lists the same instructions in the same order with `//## File .. line N`
markers (the library is compiled with -lineinfo).  The two are joined by position.  Prints the executed warp instructions,
the average active threads and the stall samples per source line, and per `// ---- ` section of the file when it has them.
"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile


def sass_lines(lib, kernel):
    out = []
    with tempfile.TemporaryDirectory() as d:
        subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=d, check=True, stdout=subprocess.DEVNULL)
        for cubin in sorted(os.listdir(d)):
            txt = subprocess.run(["nvdisasm", "-g", os.path.join(d, cubin)], capture_output=True, text=True).stdout
            if kernel not in txt:
                continue
            on, cur = False, (None, 0)
            for l in txt.splitlines():
                if l.startswith("\t.section\t.text."):
                    on = kernel in l
                    continue
                if l.startswith("\t.section") or l.startswith("//-----"):
                    if not l.startswith("//-----"):
                        on = False
                    continue
                if not on:
                    continue
                m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
                if m:
                    cur = (os.path.basename(m.group(1)), int(m.group(2)))
                    continue
                if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", l):
                    out.append((cur, l.split("*/", 1)[1].strip().rstrip(";")))
            if out:
                break
    return out


def main():
    rep, lib, kernel = sys.argv[1:4]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 40
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    start = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[start]
    body = [r for r in rows[start + 1:] if len(r) == len(hdr)]
    sass = sass_lines(lib, kernel)
    if len(sass) != len(body):
        sys.exit(f"instruction counts differ: ncu {len(body)}, nvdisasm {len(sass)} -- is {lib} the binary that was profiled?")
    ci, ct, cs = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
    by = {}
    for (loc, _), r in zip(sass, body):
        a = by.setdefault(loc, [0, 0, 0])
        a[0] += int(r[ci]); a[1] += int(r[ct]); a[2] += int(r[cs])
    tot = [sum(v[i] for v in by.values()) for i in range(3)]
    print(f"kernel {kernel}: {tot[0]} warp instructions, {tot[1] / max(tot[0], 1):.1f} threads active on average, {tot[2]} stall samples")
    srcs = {}
    for (fn, ln), v in sorted(by.items(), key=lambda kv: -kv[1][0])[:top]:
        path = next((os.path.join(dp, fn) for dp, _, fs in os.walk(os.path.dirname(os.path.abspath(lib)) + "/..") if fn in fs), None) if fn else None
        if path and path not in srcs:
            srcs[path] = open(path).read().splitlines()
        text = srcs[path][ln - 1].strip()[:90] if path and 0 < ln <= len(srcs[path]) else ""
        print(f"{100 * v[0] / tot[0]:5.1f}% inst {v[1] / max(v[0], 1):5.1f} thr {100 * v[2] / max(tot[2], 1):5.1f}% stall  {fn}:{ln}  {text}")
    # per section of the file that holds most of the instructions
    if srcs:
        path = max(srcs, key=lambda p: sum(v[0] for (fn, _), v in by.items() if fn == os.path.basename(p)))
        marks = [(i + 1, l.strip()[:70]) for i, l in enumerate(srcs[path]) if l.lstrip().startswith("// ----") or re.match(r"\s*(template|__global__|auto \w+ = \[)", l)]
        sec = {}
        for (fn, ln), v in by.items():
            if fn != os.path.basename(path):
                continue
            name = next((m for m in reversed(marks) if m[0] <= ln), (0, "(top)"))
            a = sec.setdefault(name, [0, 0, 0])
            for i in range(3):
                a[i] += v[i]
        print(f"-- by section of {os.path.basename(path)}")
        for (ln, name), v in sorted(sec.items()):
            print(f"{100 * v[0] / tot[0]:5.1f}% inst {v[1] / max(v[0], 1):5.1f} thr {100 * v[2] / max(tot[2], 1):5.1f}% stall  line {ln}: {name}")


if __name__ == "__main__":
    main()
