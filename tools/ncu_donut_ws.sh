set -x
for k in donut_f32 donut_f64; do
  python tools/ncu_one.py $k > gpurun_out/r2q_plain_$k.log 2>&1 && \
  timeout 500 ncu --set full --clock-control none --import-source on -k regex:donut_ws -s 1 -c 1 -o gpurun_out/r2q_$k -f python tools/ncu_one.py $k > gpurun_out/r2q_ncu_$k.log 2>&1
  cat gpurun_out/r2q_plain_$k.log; tail -2 gpurun_out/r2q_ncu_$k.log
done
python bench.py --steps 10 --warmup 3 > gpurun_out/r2q_bench_1gpu.json 2> gpurun_out/r2q_bench.err; tail -c 600 gpurun_out/r2q_bench_1gpu.json; tail -3 gpurun_out/r2q_bench.err
