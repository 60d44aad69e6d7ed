# Round-end evidence: tests, default bench line, whole-step traffic, launch list and ncu full captures of the obj_mask kernels.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -1 gpurun_out/gpu_tests.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference rc=$?"
timeout 600 python tools/traffic.py 4 > gpurun_out/traffic4.txt 2>&1; tail -11 gpurun_out/traffic4.txt
timeout 300 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --csv --log-file gpurun_out/config4_launches.csv python bench.py --config 4 --steps 2 --warmup 1 --no-cpu --per-config 0 --verify 0 --e2e-steps 0 --graph 0 > /dev/null 2>&1
for k in canny32 canny_resolve mask_planes; do
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$k -s 2 -c 1 -o gpurun_out/om_$k -f python tools/time_objmask.py 512 1 > gpurun_out/ncu_$k.log 2>&1; echo "ncu $k rc=$?"
done
