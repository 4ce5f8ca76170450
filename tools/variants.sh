for v in gpurun_variants/*.so; do
  cp $v easyav1_b200/libeasyav1_cuda.so
  n=$(basename $v .so)
  timeout 200 python tools/frame_times.py > gpurun_out/s14_ft_$n.txt 2>&1
  echo "$n: $(awk '{for(i=1;i<=NF;i++) if($i=="recon") s+=$(i+1)} END{print s}' gpurun_out/s14_ft_$n.txt) us recon for 30 frames; key $(head -1 gpurun_out/s14_ft_$n.txt | awk '{for(i=1;i<=NF;i++) if($i=="recon") print $(i+1)}')"
done
