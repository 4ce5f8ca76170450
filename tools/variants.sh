for w in 8 16 1 8 16 1; do
  EAV1_EST_WEIGHT=$w python tools/frame_times.py 120 > gpurun_out/s29_ft_$w.txt 2>&1
  echo "est weight $w: $(awk '{for(i=1;i<=NF;i++) if($i=="recon") s+=$(i+1)} END{print s}' gpurun_out/s29_ft_$w.txt) us recon; top: $(awk '{for(i=1;i<=NF;i++) if($i=="recon") print $(i+1)}' gpurun_out/s29_ft_$w.txt | sort -rn | head -4 | tr '\n' ' ')"
done
