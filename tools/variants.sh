cp easyav1_b200/libeasyav1_cuda.so /tmp/base.so
for v in base occ3 occ4 base occ3; do
  if [ $v = base ]; then cp /tmp/base.so easyav1_b200/libeasyav1_cuda.so; else cp gpurun_variants/$v.so easyav1_b200/libeasyav1_cuda.so; fi
  python bench.py --steps 2 --warmup 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$v', round(d['value'],1), round(d['ms_per_step'],2), {k:v['ms_per_step'] for k,v in d['roofline']['stages'].items()})"
done
