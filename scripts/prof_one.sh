# ncu --set full of one search: bash scripts/prof_one.sh <out-name> <kernel-regex> <run_one args...>
cd $GRAFT_REPO_ROOT
name=$1; rx=$2; shift 2
python scripts/run_one.py "$@" > gpurun_out/${name}_plain.log 2>&1 && \
ncu --set full --metrics smsp__inst_executed_op_local_ld.sum,smsp__inst_executed_op_local_st.sum --clock-control none --import-source on -k regex:$rx -s 1 -c 1 -o gpurun_out/$name python scripts/run_one.py "$@" > gpurun_out/${name}_ncu.log 2>&1
cat gpurun_out/${name}_plain.log
