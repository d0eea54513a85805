#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_ops_gpu.py -q -m gpu --tb=short -k "bilstm or losses" > gpurun_out/lstm_tests.log 2>&1; echo "lstm tests rc=$?"
tail -15 gpurun_out/lstm_tests.log
timeout 900 python -m pytest tests -q -m gpu --tb=short -x > gpurun_out/tests.log 2>&1; echo "all tests rc=$?"
tail -5 gpurun_out/tests.log
for v in 0 1; do
SVPC_LSTM_STEPS=$v timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-kernel-timing > gpurun_out/bench_lstm$v.json 2> gpurun_out/bench_lstm$v.err; echo "bench steps=$v rc=$?"
tail -2 gpurun_out/bench_lstm$v.err | cut -c1-300; cut -c1-200 gpurun_out/bench_lstm$v.json
done
