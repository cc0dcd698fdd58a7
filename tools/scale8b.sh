#!/bin/bash
# N = 8: weak and strong scaling legs + the C++ multi-device frame loop (run with gpurun --gpus 8)
run() { # name, extra args
  name=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 \
    bench.py --gpus 8 --steps 10 --warmup 3 "$@" > gpurun_out/bench_n8_$name.json 2> gpurun_out/bench_n8_$name.err
  echo "$name rc=$? $(cut -c1-200 gpurun_out/bench_n8_$name.json)"
}
run weak --e2e-steps 2
run strong --scaling strong --e2e-steps 0
timeout 300 python tools/host_devices.py 8 4096000 12 > gpurun_out/host_devices_8.log 2>&1; tail -5 gpurun_out/host_devices_8.log
