for ne in 30 60 84 100 120; do for t in 9 15; do echo "== ne $ne tpu $t"; BWK_SWEEP2_TPU=$t python profiles/prof_run.py $ne 72 40 2>&1 | grep "sweep2" | tr '\n' ' '; echo; done; done
