python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for ne in 84 100 120; do python profiles/prof_run.py $ne 72 40 2>&1 | grep sweep2 | tr '\n' ' '; echo; done
BWK_FAKE_RANKS=0,2 BWK_TRACE=1 python scratch/fake_rank.py 2>&1 | grep "independent strips"
