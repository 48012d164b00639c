#!/bin/bash
# dev tool (runs on the GPU box): rebuild render_bwd.cu with a few launch-bound / batch variants
# and print the per-kernel timing of each.  Restores the file afterwards.
F=monogs_dev_b200/csrc/render_bwd.cu
cp $F /tmp/rb_orig.cu
run() {
  python monogs_dev_b200/build.py > /dev/null 2>&1 || { echo "build failed: $1"; return; }
  echo "== $1: $(cuobjdump -res-usage monogs_dev_b200/libb2gs.so 2>/dev/null | grep -A1 render_bwd | grep -o 'REG:[0-9]*\|STACK:[0-9]*' | paste - -)"
  timeout 120 python tools/quick_step.py 2>/dev/null | grep "render_bwd\|kernel ms"
}
for v in "128 8" "64 8" "64 0" "128 9"; do
  set -- $v
  cp /tmp/rb_orig.cu $F
  sed -i "s/constexpr int RB_BATCH = 256;/constexpr int RB_BATCH = $1;/" $F
  if [ "$2" != "0" ]; then sed -i "s/__global__ void __launch_bounds__(RB_THREADS)\$/__global__ void __launch_bounds__(RB_THREADS, $2)/" $F; fi
  run "batch=$1 minblocks=$2"
done
cp /tmp/rb_orig.cu $F
python monogs_dev_b200/build.py > /dev/null 2>&1
