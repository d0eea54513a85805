#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/oracle_on_gpu.py > gpurun_out/r1_e_oracle_on_gpu.json 2> gpurun_out/oracle_gpu.err; echo "rc=$?"; tail -3 gpurun_out/oracle_gpu.err; cat gpurun_out/r1_e_oracle_on_gpu.json
