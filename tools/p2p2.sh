# 2 GPUs: peer-memory halo vs NCCL halo — parity (one process per GPU and one thread per GPU) and speed
timeout 600 python -m pytest tests/test_gpu_slab.py tests/test_gpu_host.py -m gpu -q -x -k "nccl or slab_engines" 2>&1 | grep -v "^NCCL\|Scene Loading" | tail -6
for p in 1 0; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2954$p \
    bench.py --gpus 2 --steps 20 --warmup 3 --scaling strong --particles 4194304 --e2e-steps 0 --p2p $p > gpurun_out/bench_n2_p2p$p.json 2> gpurun_out/bench_n2_p2p$p.err
python - <<PY
import json
d=json.load(open("gpurun_out/bench_n2_p2p$p.json"))
print("p2p=$p", d["ms_per_step"], d["config"].get("p2p_halo"), d["ranks"][0]["host_phase_ms_per_step"])
PY
done
