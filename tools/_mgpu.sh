N=$1
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 200 --warmup 10 > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err; echo "rc $?"
tail -3 gpurun_out/r2_bench_n$N.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r2_bench_n$N.json").read().strip().splitlines()[-1])
print("N=$N value %.1f M ms %.4f | e2e %.1f M (sync %.1f, mat %.1f) | cfg5 %.3f s %.2f G steps/s robots %d reduce: %s" % (d["value"]/1e6, d["ms_per_step"], d["e2e"]["value"]/1e6, d["e2e"]["sync_value"]/1e6, d["e2e_materialised"]["value"]/1e6, d["config5"]["elapsed_s_max_over_ranks"], d["config5"]["robot_steps_per_s"]/1e9, d["config5"]["robots_global"], d["config5"]["stats_reduce"]))
PY
