timeout 300 python -m pytest tests/test_hypervis.py -m gpu -x -q -s 2>&1 | tail -8
BWK_HV_UNFUSED=1 timeout 300 python -m pytest tests/test_hypervis.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python profiles/prof_hv.py 2>&1 | tail -3
BWK_HV_UNFUSED=1 timeout 300 python profiles/prof_hv.py 2>&1 | tail -3
