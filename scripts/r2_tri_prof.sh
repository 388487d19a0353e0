set -x
cd $GRAFT_REPO_ROOT
python scripts/tri_sweep.py > gpurun_out/tri_sweep.log 2>&1
NCU="ncu --set full --metrics smsp__inst_executed_op_local_ld.sum,smsp__inst_executed_op_local_st.sum --clock-control none --import-source on"
python scripts/run_one.py c4 8 3 > gpurun_out/tri_c4_plain.log 2>&1 && \
$NCU -k regex:tp_exhaustive_tri -s 1 -c 1 -o gpurun_out/r2_tri_c4 python scripts/run_one.py c4 8 3 > gpurun_out/tri_c4_ncu.log 2>&1
python scripts/run_one.py cpx64 8 3 > gpurun_out/tri_c3_plain.log 2>&1 && \
$NCU -k regex:tp_exhaustive_tri -s 1 -c 1 -o gpurun_out/r2_tri_c3 python scripts/run_one.py cpx64 8 3 > gpurun_out/tri_c3_ncu.log 2>&1
