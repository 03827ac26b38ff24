"""Per-opcode and per-instruction view of one kernel of an .ncu-rep (ncu --set full --import-source on).
usage: python profiles/source_hotspots.py gpurun_out/X.ncu-rep <kernel regex> <tiles per launch> [top N]
Prints instructions / shared-memory wavefronts per 32-slab tile by opcode, the stall-reason totals and the
most-sampled instructions.  Runs in the build container (ncu -i needs no GPU)."""
import collections
import csv
import io
import subprocess
import sys


def main():
    rep, kern, tiles = sys.argv[1], sys.argv[2], float(sys.argv[3])
    topn = int(sys.argv[4]) if len(sys.argv) > 4 else 30
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kern}"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr = rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    seen, d = set(), []
    for r in rows[2:]:
        if len(r) < len(hdr) or r[0] in seen:  # the CSV lists every instruction twice
            continue
        seen.add(r[0])
        d.append(r)

    def f(r, k):
        try:
            return float(r[ix[k]])
        except ValueError:
            return 0.0

    tot = sum(f(r, "# Samples") for r in d)
    print(f"samples {tot:.0f}  instructions/tile {sum(f(r, 'Instructions Executed') for r in d) / tiles:.1f}  "
          f"shared wavefronts/tile {sum(f(r, 'L1 Wavefronts Shared') for r in d) / tiles:.1f}")
    agg, cnt, wav, exc = (collections.Counter() for _ in range(4))
    for r in d:
        t = r[ix["Source"]].split()
        op = t[1] if t[0].startswith("@") else t[0]
        agg[op] += f(r, "# Samples"); cnt[op] += f(r, "Instructions Executed")
        wav[op] += f(r, "L1 Wavefronts Shared"); exc[op] += f(r, "L1 Wavefronts Shared Excessive")
    for op, s in cnt.most_common(24):
        print(f"  {op:32s} inst/tile {s / tiles:7.1f}  samples {100 * agg[op] / tot:5.1f}%  wavefronts/tile {wav[op] / tiles:6.1f} (excessive {exc[op] / tiles:5.1f})")
    stalls = [k for k in hdr if k.startswith("stall_") and "Not Issued" not in k]
    print("stall reasons (% of samples):", ", ".join(f"{k[6:]} {100 * sum(f(r, k) for r in d) / tot:.1f}" for k in stalls))
    top = sorted(range(len(d)), key=lambda i: -f(d[i], "# Samples"))[:topn]
    for i in sorted(top):
        r = d[i]
        st = sorted(((f(r, k), k) for k in stalls), reverse=True)[:2]
        print(f"  [{i:5d}] {100 * f(r, '# Samples') / tot:5.2f}%  {r[ix['Source']].strip()[:64]:64s} {[(k[6:], int(v)) for v, k in st]}")


if __name__ == "__main__":
    main()
