for n in 1 2 3; do
  EAV1_RECON_CTAS_PER_SM=$n timeout 200 python tools/frame_times.py > gpurun_out/s17_ft_$n.txt 2>&1
  echo "ctas/sm $n: $(awk '{for(i=1;i<=NF;i++) if($i=="recon") s+=$(i+1)} END{print s}' gpurun_out/s17_ft_$n.txt) us recon for 30 frames; key $(head -1 gpurun_out/s17_ft_$n.txt | awk '{for(i=1;i<=NF;i++) if($i=="recon") print $(i+1)}')"
done
EAV1_RECON_CTAS_PER_SM=1 timeout 120 python tools/unit_trace.py 0 0 2>&1 | tail -5
