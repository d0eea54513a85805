timeout 600 python -m pytest tests/test_model_gpu.py -q -m gpu -k "prefetch or graph" --tb=short 2>&1 | tail -5
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-kernel-timing 2>/dev/null | cut -c1-1400
