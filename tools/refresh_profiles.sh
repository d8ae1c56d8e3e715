#!/bin/bash
# builder tool, run ON the GPU box (gpurun): bench lines and ncu captures of the shipped library for profiles/r02_*
# usage: tools/refresh_profiles.sh [bench|ncu|all]
set -u
what=${1:-all}
out=gpurun_out/prof
mkdir -p $out
if [ "$what" = bench ] || [ "$what" = all ]; then
  python bench.py > $out/r02_bench_single_food_1m.json 2> $out/bench_sf.err
  python bench.py --workload defaults_1m > $out/r02_bench_defaults_1m.json 2> $out/bench_def.err
  python bench.py --workload long_horizon_64k > $out/r02_bench_long_horizon_64k.json 2>> $out/bench_def.err
  python bench.py --workload parity_4k > $out/r02_bench_parity_4k.json 2>> $out/bench_def.err
  python bench.py --workload sac_rollout_256k > $out/r02_bench_sac_rollout_256k.json 2>> $out/bench_def.err
  python bench.py --workload sac_rollout_256k --append-kernel > $out/r02_bench_sac_rollout_256k_append_kernel.json 2>> $out/bench_def.err
  python bench.py --workload single_food_1m --dtype float64 --no-extras > $out/r02_bench_single_food_1m_f64.json 2>> $out/bench_def.err
  python bench.py --workload defaults_1m --dtype float64 --no-extras > $out/r02_bench_defaults_1m_f64.json 2>> $out/bench_def.err
  python bench.py --impl reference > $out/r02_bench_reference_arm.json 2>> $out/bench_def.err
fi
if [ "$what" = ncu ] || [ "$what" = all ]; then
  NCU="ncu --set full --clock-control none --cache-control none --import-source on -k regex:salp_step_fast -s 2052"
  python tools/run_fast_once.py single_food 1048576 fast/philox > /dev/null 2>&1 || exit 1
  $NCU -c 2 -o $out/sf_philox python tools/run_fast_once.py single_food 1048576 fast/philox > $out/ncu1.log 2>&1
  $NCU -c 2 -o $out/def_philox python tools/run_fast_once.py defaults 1048576 fast/philox > $out/ncu2.log 2>&1
  $NCU -c 2 -o $out/sf_table python tools/run_fast_once.py single_food 1048576 fast/table > $out/ncu3.log 2>&1
  $NCU -c 1 -o $out/sf_philox_128k python tools/run_fast_once.py single_food 131072 fast/philox > $out/ncu4.log 2>&1
  $NCU -c 1 -o $out/def_philox_128k python tools/run_fast_once.py defaults 131072 fast/philox > $out/ncu5.log 2>&1
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $out/r02_launches_single_food_1m.csv \
      python bench.py --steps 40 --warmup 5 --no-cpu-baseline --no-extras > $out/launches_bench.json 2> $out/launches.err
fi
ls -la $out
