for a in 0 300 444 600 0 444; do
  echo "== ahead $a"
  BWK_SWEEP2_TPU=15 BWK_PREFETCH_AHEAD=$a ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum --clock-control none -k regex:strip -s 2 -c 2 python profiles/prof_run.py 60 72 40 2>&1 | grep -E "gpu__time|dram__" | tr '\n' ' '; echo
done
for a in 0 444 0 444; do echo "== ne120 ahead $a"; BWK_SWEEP2_TPU=15 BWK_PREFETCH_AHEAD=$a python profiles/prof_run.py 120 72 40 2>&1 | grep sweep2 | awk '{printf "%s ", $4}'; echo; done
