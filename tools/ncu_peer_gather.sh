# NVLink bytes (nvltx / nvlrx unit counters of GPU 0) of the fused compute + gather kernel: one process, two GPUs, the kernel on GPU 0 with gather
# targets on GPU 0 only (launches 0, 1) and on GPU 0 + GPU 1 (launches 2, 3); needs `gpurun --gpus 2`
set -x
python tools/nvlink_gather_evidence.py 200000 1 > gpurun_out/r2r_plain_peer.json 2> gpurun_out/r2r_plain_peer.err && \
ncu --metrics nvltx__bytes.sum,nvltx__bytes_data_user.sum,nvltx__bytes_packet_request_data_user.sum,nvlrx__bytes.sum,nvlrx__bytes_data_user.sum,gpu__time_duration.sum \
    --clock-control none -k regex:cepstrum_kernel -c 8 --csv --log-file gpurun_out/r2r_ncu_peer.csv python tools/nvlink_gather_evidence.py 200000 1 > gpurun_out/r2r_ncu_peer.log 2>&1
tail -12 gpurun_out/r2r_ncu_peer.csv
