o=gpurun_out; tag=$1
timeout 900 python -m pytest tests -m gpu -x -q > $o/pytest_$tag.log 2>&1; tail -4 $o/pytest_$tag.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > $o/bench_$tag.json 2> $o/bench_$tag.err; tail -3 $o/bench_$tag.err; python - <<PY
import json
d=json.load(open("$o/bench_$tag.json"))
print("ms/step", d["ms_per_step"], {k:round(v["ms_per_step"],3) for k,v in d["passes"].items() if v["ms_per_step"]})
print("e2e", d["e2e"]["value"], d["e2e"]["serial"], d["e2e"]["simulate_frames"]["value"])
PY
for lib in libsphb.so build/libsphb_f7.so; do echo $lib; SPHB_LIB=$PWD/libclsph_b200/$lib AB_LEGS=host AB_PASSES=1 timeout 300 python tools/ab_frame.py 16777216 2>&1 | tail -1; done
