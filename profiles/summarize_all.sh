# profiles/<name>_ncu_full.txt from gpurun_out/<name>.ncu-rep: the quoted counters + the per-opcode instruction mix
for n in "$@"; do
  python profiles/summarize.py gpurun_out/$n.ncu-rep > profiles/${n}_ncu_full.txt
  ncu -i gpurun_out/$n.ncu-rep --page raw --print-metric-instances details --metrics sass__inst_executed_per_opcode 2>/dev/null | tr -d '\n' | sed 's/  */ /g' | sed 's/.*sass__inst_executed_per_opcode/sass__inst_executed_per_opcode/' >> profiles/${n}_ncu_full.txt
  echo >> profiles/${n}_ncu_full.txt
  ncu -i gpurun_out/$n.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); hdr=rows[0]
for k in ['smsp__inst_executed_op_local_ld.sum','smsp__inst_executed_op_local_st.sum','l1tex__t_sector_pipe_lsu_mem_local_op_ld_hit_rate.pct','smsp__sass_inst_executed_op_shared_ld.sum','launch__shared_mem_per_block_dynamic','launch__shared_mem_per_block_static']:
    if k in hdr: print('  %s = %s' % (k, rows[2][hdr.index(k)]))
" >> profiles/${n}_ncu_full.txt
done
