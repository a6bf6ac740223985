#!/bin/bash
# Everything under profiles/ that is measured on the GPU box, for the code as it is: run under gpurun, then copy
# gpurun_out/<tag>_* into profiles/ (tools/profile_collect.sh).   Usage: tools/profile_pack.sh <tag>
cd "$(dirname "$0")/.."
T=${1:-r2}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm --format=csv > $O/${T}_gpu.txt; nproc >> $O/${T}_gpu.txt
# 1. bench lines (own arm, reference arm)
timeout 900 python bench.py > $O/${T}_bench_n1.json 2> $O/${T}_bench_n1.err; echo "bench rc $?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/${T}_bench_reference_arm.json 2>/dev/null
# 2. launch list of a short bench run (per-launch times are cold-cache and serialised: shares, not absolutes)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-secondary --config5-robots 65536 --config5-ticks 10 > $O/${T}_ncu_launches.log 2>&1
# 3. full captures of the dominant kernel: headline shape (128 slots, shared memory) and throughput shape (256 slots, TMEM)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_mpc_slots -s 4 -c 1 -o $O/${T}_full_slots \
    python bench.py --steps 3 --warmup 3 --no-cpu --no-secondary --no-config5 > $O/${T}_ncu_full.log 2>&1
B2C_SLOTS_CFG=256,16 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_mpc_slots -s 4 -c 1 -o $O/${T}_full_slots_tmem \
    python bench.py --steps 3 --warmup 3 --no-cpu --no-secondary --no-config5 > $O/${T}_ncu_full_tmem.log 2>&1
timeout 600 ncu --set full --clock-control none -k regex:k_mpc_prepare -s 4 -c 1 -o $O/${T}_full_prepare \
    python bench.py --steps 3 --warmup 3 --no-cpu --no-secondary --no-config5 > $O/${T}_ncu_full_prep.log 2>&1
# 4. per-warp traces and batch sweep
for cfg in 128,0 256,16; do B2C_SLOTS_CFG=$cfg timeout 200 python tools/slot_trace.py > $O/${T}_slot_trace_${cfg/,/_}.json 2>/dev/null; done
timeout 300 python tools/slot_trace.py --B 1048576 > $O/${T}_slot_trace_1M.json 2>/dev/null
timeout 600 python tools/sweep_batch.py --sizes 1,256,4096,65536,262144,1048576 > $O/${T}_batch_sweep.jsonl 2>/dev/null
# 5. the reference's unmodified run_simulation.py on the drop-in (SURVEY 8b acceptance run)
for mode in lqr mpc; do
  rm -rf /tmp/sim_$mode; mkdir -p /tmp/sim_$mode
  (cd /tmp/sim_$mode && timeout 600 python $OLDPWD/tools/run_reference_sim.py --mode $mode --no-plot > $OLDPWD/$O/${T}_refsim_$mode.log 2>&1; echo "rc $?" >> $OLDPWD/$O/${T}_refsim_$mode.log
   cp logs/states_*.csv $OLDPWD/$O/${T}_refsim_${mode}_states.csv; cp logs/controls_*.csv $OLDPWD/$O/${T}_refsim_${mode}_controls.csv)
done
ls -la $O | grep ${T}_ | awk '{print $5, $9}'
