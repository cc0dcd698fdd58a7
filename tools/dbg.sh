python tools/debug_slab.py 1 > gpurun_out/dbg_slab1.log 2>&1; tail -40 gpurun_out/dbg_slab1.log
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_all.log 2>&1; tail -15 gpurun_out/pytest_gpu_all.log
