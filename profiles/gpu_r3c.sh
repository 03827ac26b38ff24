#!/bin/bash
# MPDATA after the size-specialised instantiation: GPU tests, timing at 48 / 1184 / 9472 slices (specialised and generic)
tag=$1
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_zzz_mpdata.py -m gpu -x -q 2>&1 | tail -3
timeout 200 python profiles/prof_mpdata.py > gpurun_out/${tag}_mpdata_prof.json 2> gpurun_out/${tag}_mpdata_prof.err
MPDATA_B200_GENERIC=1 timeout 200 python profiles/prof_mpdata.py --slices 9472 > gpurun_out/${tag}_mpdata_prof_generic.json 2>> gpurun_out/${tag}_mpdata_prof.err
python - <<'PY'
import json
for f in ("mpdata_prof","mpdata_prof_generic"):
    d=json.load(open(f"gpurun_out/r3c_{f}.json"))
    for r in d["runs"]:
        print(f, r["nslices"], "ms", round(r["ms_per_step"],5), r["launch"], "frac", round(r["frac_of_hbm"],4), "bitwise", r["bitwise_equal_to_oracle"], {k: round(v,4) for k,v in r["ms_per_step_by_shape"].items()})
PY
