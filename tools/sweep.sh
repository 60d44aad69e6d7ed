#!/bin/bash
# experiment sweep on the GPU box: tools/sweep.sh  (writes gpurun_out/sweep.log)
V=self-supervised-anomaly-detection_b200/build/variants
L=gpurun_out/sweep.log; : > $L
python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
run() { n="$1"; b="$2"; shift 2; echo "== $n b$b" >> $L; env "$@" python bench.py --steps 300 --warmup 10 --batch $b --no-cpu 2>/dev/null >> $L; }
for b in 512 4096; do
run deep $b X=1
run nodeep $b SSAD_LIB=$V/libssad_nodeep.so
run deep $b X=1
run nodeep $b SSAD_LIB=$V/libssad_nodeep.so
done
python tools/bench_log.py $L
