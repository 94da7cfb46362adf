#!/bin/bash
# ncu evidence of round 2: launch list of the bench command (C4), dram traffic of the C4 Gram launch, --set full captures
# of the Gram at d = 1024, the fp64 Gram, and both tensor-core projections.  Each capture only after the same command has
# exited 0 without ncu.
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --also none --no-cpu-baseline --e2e-steps 0"
timeout 600 $B > gpurun_out/p2_bench_plain.json 2> gpurun_out/p2_bench_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:dr4g -c 400 --csv --log-file gpurun_out/r02_launches_c4.csv $B > gpurun_out/p2_bench_ncu.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/r02_launches_c4.csv
timeout 300 python tools/ncu_target.py gram 33554432 512 > gpurun_out/p2_plain_c4gram.log 2>&1 && \
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct,sm__cycles_elapsed.avg.per_second,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed --clock-control none -k regex:gram_pair_kernel -s 1 -c 1 --csv --log-file gpurun_out/r02_c4_gram_traffic.csv python tools/ncu_target.py gram 33554432 512 > gpurun_out/p2_ncu_c4gram.log 2>&1
echo "c4 traffic rc=$?"; tail -8 gpurun_out/r02_c4_gram_traffic.csv | cut -c1-200
cap() {  # name kernel-regex target-args...
  local name=$1 rx=$2; shift 2
  timeout 300 python tools/ncu_target.py "$@" > gpurun_out/p2_plain_$name.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:$rx -s 1 -c 1 -f -o gpurun_out/r02_$name python tools/ncu_target.py "$@" > gpurun_out/p2_ncu_$name.log 2>&1
  echo "$name rc=$?"
}
cap gram_pair_bf16x3_d1024 gram_pair_kernel gram 2097152 1024
cap gram_pair_bf16x3_d512_final gram_pair_kernel gram 4194304 512
cap gram_pair_bf16x3_d256 gram_pair_kernel gram 1000000 256
cap gram_f64_d4096 gram_f64_kernel gram64 65536 4096
cap project_tc_c2 project_tc_kernel project 1000000 256 32
cap project_f64_c5 project_f64_kernel project64 262144 4096 100
ls -la gpurun_out/*.ncu-rep
