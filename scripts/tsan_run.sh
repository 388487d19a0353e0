# ThreadSanitizer over the native harness and a library whose host code is instrumented (tuning/verification aid):
#   TP_BUILD_TAG=tsan TP_BUILD_DEFS="-Xcompiler -fsanitize=thread -Xcompiler -g" TP_BUILD_LDFLAGS="-Xcompiler -fsanitize=thread" python rocm-k8s-device-plugin_b200/build.py
#   make -C tests/harness tsan && bash scripts/tsan_run.sh
cd $GRAFT_REPO_ROOT
python - <<PY
import sys; sys.path.insert(0,"tests"); sys.path.insert(0,"tests/harness")
import topo, run_harness
devs,pairs=topo.fixture_topology("cpx")
run_harness.write_input("/tmp/tsan_in.txt","c2_mesh_4of8",devs,pairs,["test%d"%i for i in range(1,9)],[],4)
ids=[d.Id for d in devs]
run_harness.write_input("/tmp/tsan_in2.txt","cpx_6of40",devs,pairs,ids[:40],[ids[3]],6)
run_harness.write_input("/tmp/tsan_in3.txt","cpx_8of40",devs,pairs,ids[:40],[],8)
PY
for f in /tmp/tsan_in.txt /tmp/tsan_in2.txt; do
  TSAN_OPTIONS="report_signal_unsafe=0 history_size=4" tests/harness/abi_harness_tsan $f 300 8 100 3 10 2>&1 | head -120
done
# kernel TRI, single and in concurrent batches (two tp_allocate_batch calls in flight, a third waiting for a slot)
TSAN_OPTIONS="report_signal_unsafe=0 history_size=4" tests/harness/abi_harness_tsan /tmp/tsan_in3.txt 100 8 30 3 10 2>&1 | head -120
