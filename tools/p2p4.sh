# 4 GPUs, 16.7 M strong: peer-memory halo vs NCCL halo
for p in 1 0; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 2955$p \
    bench.py --gpus 4 --steps 20 --warmup 3 --scaling strong --e2e-steps 0 --p2p $p > gpurun_out/bench_n4_strong_p2p$p.json 2> gpurun_out/bench_n4_strong_p2p$p.err
python - <<PY
import json
d=json.load(open("gpurun_out/bench_n4_strong_p2p$p.json"))
print("p2p=$p", d["ms_per_step"], d["value"], d["config"].get("p2p_halo"), [r["host_phase_ms_per_step"]["pack_exchange_unpack"] for r in d["ranks"]])
PY
done
