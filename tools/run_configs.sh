#!/bin/bash
# Developer tool (run on the GPU box through gpurun): one bench line per BASELINE.json config into gpurun_out/.
#   tools/run_configs.sh [n_gpus]
N=${1:-1}
run() { # name, args...
  name=$1; shift
  if [ "$N" = 1 ]; then timeout 600 python bench.py --gpus 1 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; fi
  echo "$name rc=$? $(head -c 300 gpurun_out/$name.json)"
}
if [ "$N" = 1 ]; then
  run r02_bench_c2_n1 --steps 8 --warmup 3
  run r02_bench_c1_n1 --workload c1 --steps 8 --warmup 3
  run r02_bench_c4_n1 --workload c4 --steps 8 --warmup 3 --no-cpu-baseline
  run r02_bench_c2_batch1_n1 --batch 1 --steps 8 --warmup 3 --no-cpu-baseline --no-cloud-e2e
  # configs[4]: the corners of the resolution / disparity sweep
  run r02_bench_c5_vga64_n1 --workload vga64 --steps 8 --warmup 3 --no-cpu-baseline --no-cloud-e2e
  run r02_bench_c5_720p128_n1 --workload 720p128 --steps 8 --warmup 3 --no-cpu-baseline --no-cloud-e2e
  run r02_bench_c5_1440p256_n1 --workload 1440p256 --batch 8 --steps 6 --warmup 3 --no-cpu-baseline --no-cloud-e2e
  run r02_bench_c5_4k256_n1 --workload 4k256 --batch 4 --steps 6 --warmup 3 --no-cpu-baseline --no-cloud-e2e
else
  run r02_bench_c2_n$N --steps 8 --warmup 3
  run r02_bench_c3_n$N --workload c3 --steps 8 --warmup 3
  run r02_bench_c4_n$N --workload c4 --steps 8 --warmup 3
fi
