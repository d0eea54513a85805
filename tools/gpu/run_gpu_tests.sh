#!/bin/bash
# GPU test suite, one pytest process per file (a CUDA fault in one file cannot hide the results of the others);
# summary lines in gpurun_out/gputests_<tag>.log
tag=${1:-run}
out=gpurun_out/gputests_${tag}.log
: > $out
for f in tests/test_*.py; do
  echo "=== $f" >> $out
  python -X faulthandler -m pytest $f -m gpu -q -x --tb=short 2>&1 | grep -v "^  File\|^Extension\|Warning\|^$\|warnings.warn\|^  /\|^    \(assert\|return\|loss\)" | tail -${2:-25} >> $out
done
grep -c "passed" $out
