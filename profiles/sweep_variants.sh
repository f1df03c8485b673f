for v in "" c128 c512 c1024 r3 r5; do
  if [ -z "$v" ]; then lib=""; else lib="$PWD/k8s-gpu-scheduler_b200/libgpusched_$v.so"; fi
  GS_LIB_PATH=$lib python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v', d['ms_per_step'], d['roofline']['kernel_ms'])"
done
