#!/bin/bash
# Developer tool (run on the GPU box through gpurun): the round's records -- tests, one bench line per BASELINE.json config,
# the reference arm, the ncu launch list of the bench command and one full ncu capture of the two sweeps.
python -m pytest tests -x -q -m gpu 2>&1 | tail -3
bash tools/run_configs.sh 1 2>&1 | cut -c1-150
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_reference_cpu.json 2> gpurun_out/r02_bench_reference_cpu.err; echo "reference rc=$? $(head -c 200 gpurun_out/r02_bench_reference_cpu.json)"
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r02_launch_list_c2_batch16.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-cloud-e2e > gpurun_out/ncu_launch.log 2>&1; echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:wave_kernel -c 2 -o gpurun_out/r02_wave_final python tools/wave_bench.py --reps 1 > gpurun_out/ncu_wave_final.log 2>&1; echo "wave ncu rc=$?"
ncu --set full --clock-control none --import-source on -k regex:cost_kernel -s 3 -c 1 -o gpurun_out/r02_cost_final python tools/wave_bench.py --reps 1 > gpurun_out/ncu_cost_final.log 2>&1; echo "cost ncu rc=$?"
