#!/bin/bash
# N = 1, 2, 4, 8 on ONE box, back to back (what the driver's SCALE step does), short form: weak
# scaling only, no parity check / secondary legs / end-to-end leg.  Needs an 8-GPU box.
o=gpurun_out; mkdir -p $o
python bench.py --steps 10 --warmup 3 --no-cpu --no-secondary --e2e-steps 0 > $o/scale_n1.json 2> $o/scale_n1.err; echo "n1 rc=$?"
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) \
    bench.py --gpus $n --steps 10 --warmup 3 --no-secondary --e2e-steps 0 --no-parity > $o/scale_n$n.json 2> $o/scale_n$n.err; echo "n$n rc=$?"
done
python - <<'PY'
import json
base = None
for n in (1, 2, 4, 8):
    d = json.load(open(f"gpurun_out/scale_n{n}.json"))
    if n == 1: base = d["value"]
    print(f"N={n}: {d['ms_per_step']:.3f} ms per step, {d['value']:.4g} particle-steps/s, efficiency {d['value'] / (n * base):.3f}")
PY
