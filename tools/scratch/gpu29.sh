mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest29.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest29.log | cut -c1-250
export PROF_ONLY=strided0+field,rows+field+add+dot
for v in 1 0; do echo "== PROF_SYM=$v"; PROF_SYM=$v timeout 300 python tools/prof_passes.py f32 1024 5 2>&1 | tail -2; done | tee gpurun_out/passes29.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench29.json 2> gpurun_out/bench29.err; echo "bench rc=$?"; python - <<'PY'
import json; d=json.load(open('gpurun_out/bench29.json')); print(d['value'], d['ms_per_step'], d['e2e']['value']); print(json.dumps(d['extra']['kernel_ms'], indent=0)); print(d['extra']['fwd_applies_per_s'], d['extra']['adj_applies_per_s'])
PY
