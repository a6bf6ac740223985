#!/bin/bash
# Copy one evidence pack (tools/profile_pack.sh <tag>, merged back into gpurun_out/) into profiles/r2_* and derive the
# summaries from the captures.   Usage: tools/profile_collect.sh <tag>
cd "$(dirname "$0")/.."
T=${1:-r2b}; O=gpurun_out; P=profiles
for f in bench_n1.json bench_reference_arm.json launches.csv slot_trace_128_0.json slot_trace_256_16.json slot_trace_1M.json batch_sweep.jsonl gpu.txt refsim_lqr.log refsim_mpc.log; do cp $O/${T}_$f $P/r2_$f; done
python tools/launch_summary.py $P/r2_launches.csv > $P/r2_launches_summary.md
python tools/ncu_summary.py $O/${T}_full_slots.ncu-rep "Round 2 -- k_mpc_slots<128,5,0> (headline shape: 128 slots per SM, strips in shared memory): ncu --set full --clock-control none --import-source on -k regex:k_mpc_slots -s 4 -c 1 python bench.py --steps 3 --warmup 3 --no-cpu --no-secondary --no-config5 (B200, 65,536 instances, N=20, M=5)" > $P/r2_ncu_full_summary.md
python tools/ncu_summary.py $O/${T}_full_slots_tmem.ncu-rep "Round 2 -- k_mpc_slots<256,5,16> (throughput shape forced with B2C_SLOTS_CFG=256,16: 256 slots per SM, first 16 stages of U in tensor memory), same command" > $P/r2_ncu_full_tmem_summary.md
python tools/ncu_summary.py $O/${T}_full_prepare.ncu-rep "Round 2 -- k_mpc_prepare<5> (one thread per instance: unwrap, sincos, records, classes), same command with -k regex:k_mpc_prepare" > $P/r2_ncu_full_prepare_summary.md 2>/dev/null
ncu -i $O/${T}_full_slots.ncu-rep --page source --csv --print-source cuda,sass > /tmp/srcs_$T.csv 2>/dev/null
python tools/ncu_top_lines.py 40 /tmp/srcs_$T.csv > $P/r2_ncu_top_lines.md
python tools/ncu_traffic.py $O/${T}_full_slots.ncu-rep $P/r2_slots_traffic.json 65536 > /dev/null
grep -E "kernel_src_sha|ratio_to_algorithmic|dram_bytes_per_launch" $P/r2_slots_traffic.json
