python tools/e2e_sweep.py default
FBN_NO_RAMP=1 python tools/e2e_sweep.py default
python -m pytest tests/test_gpu_parity.py -x -q -m gpu | tail -1
