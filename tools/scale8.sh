#!/bin/bash
# the N = 8 legs recorded under profiles/ (run with gpurun --gpus 8)
run() { # name, extra args
  name=$1; shift
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 \
    bench.py --gpus 8 --steps 10 --warmup 3 "$@" > gpurun_out/bench_n8_$name.json 2> gpurun_out/bench_n8_$name.err
  echo "$name rc=$? $(cut -c1-160 gpurun_out/bench_n8_$name.json)"
}
run weak --e2e-steps 2
run strong --scaling strong --e2e-steps 0
run strong_labyrinth --scaling strong --particles 4096000 --scene labyrinth --e2e-steps 0
run strong_river --scaling strong --particles 4096000 --scene river --e2e-steps 0
