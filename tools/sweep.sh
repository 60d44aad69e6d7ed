#!/bin/bash
# experiment sweep on the GPU box: tools/sweep.sh  (writes gpurun_out/sweep.log)
V=self-supervised-anomaly-detection_b200/build/variants
L=gpurun_out/sweep.log; : > $L
run() { n="$1"; shift; echo "== $n" >> $L; env "$@" >> $L 2>/dev/null; }
for o in f32 bf16; do
run "base $o" X=1 python bench.py --steps 300 --warmup 10 --out $o --no-cpu
run "normcs $o" SSAD_LIB=$V/libssad_normcs.so python bench.py --steps 300 --warmup 10 --out $o --no-cpu
run "base $o" X=1 python bench.py --steps 300 --warmup 10 --out $o --no-cpu
run "normcs $o" SSAD_LIB=$V/libssad_normcs.so python bench.py --steps 300 --warmup 10 --out $o --no-cpu
done
python tools/bench_log.py $L
# planes kernel with deeper row prefetch: kernel times from ncu
for l in "" $V/libssad_omrd.so; do
  SSAD_LIB=$l ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"canny_nms|mask_planes" --launch-skip 4 -c 4 --csv python bench.py --config 4 --graph 0 --inflight 1 --steps 2 --warmup 3 --no-cpu 2>/dev/null | grep -o '"[a-z_]*_kernel[^"]*".*' | awk -F'","' '{print substr($1,2,18), $NF}'
done
SSAD_LIB=$V/libssad_omrd.so python -m pytest tests/test_gpu_objmask.py -m gpu -x -q 2>&1 | tail -1
