for t in 3 5 9 15 30 90; do
  echo "== tpu $t"
  BWK_SWEEP2_TPU=$t ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,gpu__time_duration.sum --clock-control none -k regex:strip -s 1 -c 1 python profiles/prof_run.py ${1:-30} 72 40 2>&1 | grep -E "dram__|lts__|gpu__time"
done
