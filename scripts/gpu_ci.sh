#!/bin/bash
# GPU check run under gpurun (one GPU): tests, smoke, driver binary, benches, per-phase times, the filter
# kernel's wait profile and the host -> device copy rate.  Logs -> gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
echo "== pytest gpu" ; timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -40 | tee gpurun_out/pytest_gpu.log
echo "== smoke"; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -5 | tee gpurun_out/smoke.log
echo "== engine binary"; (timeout 300 ./engine --json --synth 20000 1000 64 -100 100 1 16 10 42 | sort | md5sum; DMLP_BACKEND=cpu ./engine --synth 20000 1000 64 -100 100 1 16 10 42 | sort | md5sum) 2>&1 | tee gpurun_out/engine.log
echo "== bench small"; timeout 600 python bench.py --steps 3 --warmup 3 --num-attrs 1024 --num-queries 8192 --num-data 16384 --verify 64 2>&1 | tail -1 | tee gpurun_out/bench_small.log
echo "== bench full"; timeout 900 python bench.py 2>&1 | tail -1 | tee gpurun_out/bench_1gpu.json
echo "== phase times"; timeout 300 python scripts/phase_times.py --reps 3 2>&1 | tail -1 | tee gpurun_out/phase_full.json
echo "== filter kernel wait profile"; timeout 300 python scripts/topk_prof.py 2>&1 | tail -1 | tee gpurun_out/topk_prof.json
echo "== h2d rate"; timeout 300 python scripts/h2d_bw.py 2>&1 | tail -1 | tee gpurun_out/h2d_bw.json
