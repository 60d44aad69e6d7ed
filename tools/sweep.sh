#!/bin/bash
# experiment sweep on the GPU box: tools/sweep.sh  (writes gpurun_out/sweep.log)
V=self-supervised-anomaly-detection_b200/build/variants
L=gpurun_out/sweep.log; : > $L
run() { n="$1"; b="$2"; shift 2; echo "== $n b$b" >> $L; env "$@" python bench.py --steps 300 --warmup 10 --batch $b --no-cpu 2>/dev/null >> $L; }
for b in 512 4096; do
run base $b X=1
for v in n11 n12 n21 n22 n33 n24 n42; do run $v $b SSAD_LIB=$V/libssad_$v.so; done
run base $b X=1
done
python tools/bench_log.py $L
