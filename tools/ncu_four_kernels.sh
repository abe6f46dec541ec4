set -x
for k in cepstrum pcm16 donut_f32 donut_f64; do
  case $k in cepstrum|pcm16) pat=cepstrum_kernel;; donut_f32|donut_f64) pat=donut_ws;; esac
  python tools/ncu_one.py $k > gpurun_out/r2f_plain_$k.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:$pat -s 1 -c 1 -o gpurun_out/r2f_$k -f python tools/ncu_one.py $k > gpurun_out/r2f_ncu_$k.log 2>&1
  cat gpurun_out/r2f_plain_$k.log; tail -2 gpurun_out/r2f_ncu_$k.log
done
