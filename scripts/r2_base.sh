set -x
cd $GRAFT_REPO_ROOT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
NCU="ncu --set full --metrics smsp__inst_executed_op_local_ld.sum,smsp__inst_executed_op_local_st.sum --clock-control none --import-source on"
python scripts/ab_probe.py > gpurun_out/r2_base_ab.log 2>&1
python scripts/small_probe.py > gpurun_out/r2_base_small.log 2>&1
python scripts/run_one.py cpx64 8 3 > gpurun_out/r2_base_c3_plain.log 2>&1 && \
$NCU -k regex:tp_exhaustive_split -s 1 -c 1 -o gpurun_out/r2_base_c3 python scripts/run_one.py cpx64 8 3 > gpurun_out/r2_base_c3_ncu.log 2>&1
python scripts/shard_one.py 3 8 > gpurun_out/r2_base_shard_plain.log 2>&1 && \
$NCU -k regex:tp_exhaustive_split -s 1 -c 1 -o gpurun_out/r2_base_shard python scripts/shard_one.py 3 8 > gpurun_out/r2_base_shard_ncu.log 2>&1
python scripts/latency_one.py 20 > gpurun_out/r2_base_lat_plain.log 2>&1 && \
$NCU -k regex:tp_exhaustive_lex_small -s 5 -c 1 -o gpurun_out/r2_base_lexsmall python scripts/latency_one.py 20 > gpurun_out/r2_base_lat_ncu.log 2>&1
TP_BATCH_TRACE=1 python scripts/pod_storm.py > gpurun_out/r2_base_storm_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_base_storm_launches.csv python scripts/pod_storm.py 10000 > gpurun_out/r2_base_storm_ncu.log 2>&1
echo done
