# A/B of tagged builds: bash scripts/ab_q.sh tag1 tag2 ...   ("base" = the default build)
cd $GRAFT_REPO_ROOT
for tag in "$@"; do
  lib=rocm-k8s-device-plugin_b200/libtpalloc_$tag.so
  [ "$tag" = base ] && lib=rocm-k8s-device-plugin_b200/libtpalloc.so
  echo "=== $tag"
  export TP_ALLOC_LIB=$PWD/$lib
  for rep in 1 2; do
    timeout 100 python scripts/run_one.py cpx64 8 10 | cut -c1-70
    timeout 100 python scripts/run_one.py c4 8 10 | cut -c1-70
  done
  timeout 200 python scripts/batch_probe.py | grep -E "8 of (48|56)" | cut -c1-140
done
