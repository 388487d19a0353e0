# Round-2 ncu captures of the committed build (summaries go to profiles/ through profiles/summarize.py)
cd $GRAFT_REPO_ROOT
bash scripts/prof_one.sh r2_final_tri64_c3 tp_exhaustive_tri cpx64 8 3
bash scripts/prof_one.sh r2_final_tri80_c4 tp_exhaustive_tri c4 8 3
NCU="ncu --set full --metrics smsp__inst_executed_op_local_ld.sum,smsp__inst_executed_op_local_st.sum --clock-control none --import-source on"
python scripts/shard_one.py 3 8 > gpurun_out/r2_final_shard_plain.log 2>&1 && \
$NCU -k regex:tp_exhaustive_tri -s 1 -c 1 -o gpurun_out/r2_final_tri80_shard3of8 python scripts/shard_one.py 3 8 > gpurun_out/r2_final_shard_ncu.log 2>&1
python scripts/run_one.py cpx64 8 3 2 > gpurun_out/r2_final_split64_plain.log 2>&1 && \
$NCU -k regex:tp_exhaustive_split -s 1 -c 1 -o gpurun_out/r2_final_split64_c3 python scripts/run_one.py cpx64 8 3 2 > gpurun_out/r2_final_split64_ncu.log 2>&1
TP_BATCH_TRACE=1 python scripts/pod_storm.py > gpurun_out/r2_final_storm_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_final_storm_launches.csv python scripts/pod_storm.py 10000 > gpurun_out/r2_final_storm_ncu.log 2>&1
python bench.py --steps 3 --warmup 3 --no-cpu --no-extra > gpurun_out/r2_final_bench_short.json 2>/dev/null && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_final_bench_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-extra > /dev/null 2>&1
cat gpurun_out/r2_final_*plain.log
