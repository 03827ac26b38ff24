for d in 0 1 3 7; do echo "== BWK_COLSWZ=$d"; BWK_COLSWZ=$d timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "small_cubed or ne30_one" 2>&1 | tail -3; done
