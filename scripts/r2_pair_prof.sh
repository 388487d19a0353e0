# Round-2 ncu captures of the committed build, second form of kernel TRI (summaries go to profiles/ through
# profiles/summarize_all.sh); r2_pair_c3 comes from `bash scripts/prof_one.sh r2_pair_c3 tp_exhaustive_tri cpx64 8 3`
cd $GRAFT_REPO_ROOT
bash scripts/prof_one.sh r2_pair_c4 tp_exhaustive_tri c4 8 3
NCU="ncu --set full --metrics smsp__inst_executed_op_local_ld.sum,smsp__inst_executed_op_local_st.sum --clock-control none --import-source on"
python scripts/shard_one.py 3 8 > gpurun_out/r2_pair_shard_plain.log 2>&1 && \
$NCU -k regex:tp_exhaustive_tri -s 1 -c 1 -o gpurun_out/r2_pair_shard3of8 python scripts/shard_one.py 3 8 > gpurun_out/r2_pair_shard_ncu.log 2>&1
TP_BATCH_TRACE=1 python scripts/pod_storm.py > gpurun_out/r2_pair_storm_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_pair_storm_launches.csv python scripts/pod_storm.py 10000 > gpurun_out/r2_pair_storm_ncu.log 2>&1
python bench.py --steps 3 --warmup 3 --no-cpu --no-extra > gpurun_out/r2_pair_bench_short.json 2>/dev/null && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_pair_bench_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-extra > /dev/null 2>&1
cat gpurun_out/r2_pair_*plain.log
