for v in nofence spin p40 nofence spin p40; do
  cp gpurun_variants/$v.so easyav1_b200/libeasyav1_cuda.so
  python tools/frame_times.py 120 > gpurun_out/s26_ft_$v.txt 2>&1
  echo "$v: $(awk '{for(i=1;i<=NF;i++) if($i=="recon") s+=$(i+1)} END{print s}' gpurun_out/s26_ft_$v.txt) us recon; $(timeout 100 python tools/unit_trace.py 0 0 2>&1 | grep critical)"
done
