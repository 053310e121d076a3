# A/B of the tcgen05 network kernel variants on the benched frame (device-timed, per-kernel CUDA events)
for o in "eval_texsplit=0" "eval_texsplit=1" "eval_texsplit=2" "eval_texsplit=3" "eval_texsplit=4" "eval_texsplit=5"; do
  args=""; for kv in $o; do args="$args --opt $kv"; done
  python bench.py --steps 10 --warmup 3 --no-cpu-baseline $args 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); k=d['roofline']['kernel_ms_per_frame']
print('$o', 'ms/frame %.3f'%d['ms_per_step'], 'eval %.3f march %.3f comp %.3f mesh %.3f raygen %.3f'%(k['ms_eval'],k['ms_march'],k['ms_composite'],k['ms_mesh'],k['ms_raygen']))"
done
