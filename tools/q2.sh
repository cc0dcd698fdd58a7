o=gpurun_out; tag=$1
timeout 900 python -m pytest tests -m gpu -x -q -s > $o/pytest_$tag.log 2>&1; tail -4 $o/pytest_$tag.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > $o/bench_$tag.json 2> $o/bench_$tag.err; python - <<PY
import json
d=json.load(open("$o/bench_$tag.json"))
print("ms/step", d["ms_per_step"], {k:round(v["ms_per_step"],3) for k,v in d["passes"].items() if v["ms_per_step"]})
print("1M", d["config"]["secondary"]["ms_per_step"], d["config"]["secondary"]["pass_ms_per_step"])
print("e2e", d["e2e"]["value"], d["frame_loop"]["ms_per_step"])
PY
AB_LEGS=host,host_noside timeout 300 python tools/ab_frame.py 16777216 2>&1 | tail -2
