#!/bin/bash
# AddressSanitizer + UndefinedBehaviorSanitizer over the CPU-side native code: the C++ host classes
# (libclsph_b200/host/ -> libclsph_host.so) under tests/test_host.py, and the plain-C oracle
# (oracle/sph_oracle.c) under tests/test_oracle_golden.py.  Instrumented copies are built under
# libclsph_b200/build/asan/ and selected with CLSPH_HOST_LIB / SPHB_ORACLE_LIB; the product
# libraries are not touched.  Then ThreadSanitizer over the threaded host code.  Prints the number of sanitizer reports (0 = clean).
#   tools/sanitize_cpu.sh            (needs libclsph_b200/libsphb.so: python -m libclsph_b200.build)
set -e
cd "$(dirname "$0")/.."
OUT=libclsph_b200/build/asan
mkdir -p "$OUT" gpurun_out
SAN="-fsanitize=address,undefined -fno-omit-frame-pointer -g -O1"
python - "$OUT" $SAN <<'PY'
import os, subprocess, sys
from libclsph_b200 import build as B
out, san = sys.argv[1], sys.argv[2:]
flags = [f for f in B.HOST_FLAGS if f != "-O2"] + san
objs = []
for src in B.HOST_SOURCES:
    o = os.path.join(out, src.replace("/", "_").replace(".cpp", ".o"))
    objs.append(o)
    subprocess.check_call(["g++"] + flags + ["-c", os.path.join(B.HOST, src), "-o", o])
subprocess.check_call(["g++", "-shared"] + san + ["-o", os.path.join(out, "libclsph_host.so")] + objs +
                      ["-L", B.HERE, "-lsphb", "-Wl,-rpath," + B.HERE, "-pthread"])
PY
gcc -std=c11 $SAN -ffp-contract=off -fno-fast-math -fopenmp -fPIC -shared -o "$OUT/liboracle.so" \
    oracle/sph_oracle.c -lm
export LD_PRELOAD="$(gcc -print-file-name=libasan.so) $(gcc -print-file-name=libubsan.so)"
export ASAN_OPTIONS=detect_leaks=0:halt_on_error=0 UBSAN_OPTIONS=print_stacktrace=1
CLSPH_HOST_LIB="$PWD/$OUT/libclsph_host.so" python -m pytest tests/test_host.py -q -s -p no:cacheprovider \
    > gpurun_out/asan_host.log 2>&1 || true
SPHB_ORACLE_LIB="$PWD/$OUT/liboracle.so" python -m pytest tests/test_oracle_golden.py -q -s -p no:cacheprovider \
    > gpurun_out/asan_oracle.log 2>&1 || true
unset LD_PRELOAD
# ThreadSanitizer over the threaded host code (frame gather, deal-out, .geo / .bgeo formatting):
# a stand-alone driver, tools/tsan_host.cpp, linked with the host sources
mkdir -p libclsph_b200/build/tsan
python - <<'PY'
import os, subprocess
from libclsph_b200 import build as B
out = os.path.join(B.OBJDIR, "tsan")
flags = [f for f in B.HOST_FLAGS if f not in ("-O2", "-fPIC")] + ["-fsanitize=thread", "-g", "-O1", "-fPIE"]
srcs = [os.path.join(B.HOST, s) for s in B.HOST_SOURCES] + [os.path.join(B.HERE, "..", "tools", "tsan_host.cpp")]
subprocess.check_call(["g++"] + flags + srcs + ["-o", os.path.join(out, "tsan_host"), "-L", B.HERE, "-lsphb",
                                                "-Wl,-rpath," + B.HERE, "-pthread"])
PY
TSAN_OPTIONS=halt_on_error=0 libclsph_b200/build/tsan/tsan_host > gpurun_out/tsan_host.log 2>&1 || true
echo "gpurun_out/tsan_host.log: $(tail -1 gpurun_out/tsan_host.log) | data races: $(grep -c 'WARNING: ThreadSanitizer' gpurun_out/tsan_host.log || true)"
for f in gpurun_out/asan_host.log gpurun_out/asan_oracle.log; do
  echo "$f: $(tail -1 $f) | sanitizer reports: $(grep -c 'AddressSanitizer\|runtime error' $f || true)"
done
