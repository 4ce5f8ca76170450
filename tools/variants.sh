for m in 0 1 2; do
  EAV1_WARP_MODE=$m timeout 200 python tools/frame_times.py > gpurun_out/s20_ft_w$m.txt 2>&1
  echo "warp mode $m: $(awk '{for(i=1;i<=NF;i++) if($i=="recon") s+=$(i+1)} END{print s}' gpurun_out/s20_ft_w$m.txt) us recon for 30 frames; key $(head -1 gpurun_out/s20_ft_w$m.txt | awk '{for(i=1;i<=NF;i++) if($i=="recon") print $(i+1)}')"
done
EAV1_WARP_MODE=2 timeout 100 python tools/unit_trace.py 9 0 2>&1 | tail -14
