#!/bin/bash
tag=$1
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_zzz_mpdata.py -m gpu -x -q 2>&1 | tail -3
for pf in 1 0; do
MPDATA_B200_PREFETCH=$pf timeout 200 python profiles/prof_mpdata.py --slices 1184,9472 > gpurun_out/${tag}_mpdata_prof_pf$pf.json 2> gpurun_out/${tag}_mpdata_prof.err
python - <<PY
import json
d=json.load(open("gpurun_out/${tag}_mpdata_prof_pf$pf.json"))
for r in d["runs"]:
    print("prefetch $pf", r["nslices"], "ms", round(r["ms_per_step"],5), r["launch"], "frac", round(r["frac_of_hbm"],4), "bitwise", r["bitwise_equal_to_oracle"], {k: round(v,4) for k,v in r["ms_per_step_by_shape"].items()})
PY
done
