# A/B of tagged builds: bash scripts/ab_tri.sh tag1 tag2 ...   ("base" = the default build)
cd $GRAFT_REPO_ROOT
for tag in "$@"; do
  lib=rocm-k8s-device-plugin_b200/libtpalloc_$tag.so
  [ "$tag" = base ] && lib=rocm-k8s-device-plugin_b200/libtpalloc.so
  echo "=== $tag"
  TP_ALLOC_LIB=$PWD/$lib timeout 300 python scripts/tri_check.py ${CASES:-40} 2>&1 | grep -E "mismatch|MISMATCH|WRONG|kernel=0|Error|error" 
done
