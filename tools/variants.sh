python tools/frame_times.py 120 > gpurun_out/s23_ft120.txt 2>&1
EAV1_RECON_CTAS_PER_SM=1 python tools/frame_times.py 120 > gpurun_out/s23_ft120_c1.txt 2>&1
for k in 0 9 68; do timeout 100 python tools/unit_trace.py $k 0 > gpurun_out/s23_trace$k.txt 2>&1; done
