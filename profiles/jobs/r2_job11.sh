#!/bin/bash
# Round 2, GPU job 11: whole GPU suite + smoke after the series / engine changes; bench lines of configs 4 and 5 and of
# the default N = 1 run (both arms); ncu of the lattice-gas and hard-disc kernels.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2_pytest11.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2_pytest11.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/r2_smoke11.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2_smoke11.log
for W in "latgas" "latgas --kernel two_pass" "latgas --kernel pbatch" "latgas --layout bits" "discs" "rx --rx-model hardcore --rx-replicas-total 64 --rx-kT-lo 0.5 --rx-kT-hi 2.0"; do
  tag=$(echo $W | tr ' -' '__')
  timeout 600 python bench.py --workload $W --steps 20 --warmup 3 > gpurun_out/r2_bench11_$tag.json 2> gpurun_out/r2_bench11_$tag.err; echo "bench $W rc=$?"
done
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench11_default.json 2> gpurun_out/r2_bench11_default.err; echo "bench default rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2_bench11_reference.json 2> gpurun_out/r2_bench11_reference.err; echo "bench reference rc=$?"
A="python bench.py --workload latgas --steps 20 --warmup 3 --no-cpu"
$A > gpurun_out/r2_plain11a.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:resident -s 4 -c 2 -o gpurun_out/r2_latgas_resident $A > gpurun_out/r2_ncu11a.log 2>&1; echo "ncu a rc=$?"
B="python bench.py --workload latgas --kernel two_pass --steps 20 --warmup 3 --no-cpu"
$B > gpurun_out/r2_plain11b.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:pair_class -s 64 -c 3 -o gpurun_out/r2_latgas_pair $B > gpurun_out/r2_ncu11b.log 2>&1; echo "ncu b rc=$?"
C="python bench.py --workload discs --steps 30 --no-cpu"
$C > gpurun_out/r2_plain11c.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:discs_run -s 1 -c 2 -o gpurun_out/r2_discs $C > gpurun_out/r2_ncu11c.log 2>&1; echo "ncu c rc=$?"
