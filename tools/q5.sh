timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sm_affine or variants" 2>&1 | tail -3
AB_LEGS=host,host_sm,host_sm5,host_sm7 AB_PASSES=1 timeout 300 python tools/ab_frame.py 16777216 1000000 2>&1 | tail -2
