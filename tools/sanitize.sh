#!/bin/bash
# compute-sanitizer over the kernel-level GPU tests (hand-rolled mbarrier / TMEM / TMA pipelines are what memcheck and
# racecheck are for).  Usage, on a GPU box:  bash tools/sanitize.sh [outdir]      (about 6 minutes)
# Writes <outdir>/sanitize_<tool>.log and prints each tool's summary lines; profiles/r02_sanitizer.md holds a committed copy.
out=${1:-gpurun_out}
mkdir -p "$out"
tests="tests/test_gpu_kernels.py"
sel='gemm or attention or layernorm or fsmn or fbank or lfr or cif or argmax or ctc'
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 0 \
    python -m pytest $tests -m gpu -x -q -k "$sel" -p no:cacheprovider > "$out/sanitize_$tool.log" 2>&1
  echo "== $tool: exit $?"
  grep -a -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|Hazard|Invalid|error" "$out/sanitize_$tool.log" | sort | uniq -c | sort -rn | head -n 12
done
