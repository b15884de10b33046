mkdir -p gpurun_out
for pf in 0 64 512 0 256 2048; do
  MB200_BODY_PREFETCH=$pf timeout 200 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ab_pf$pf.json 2> gpurun_out/ab_pf$pf.err
  python - <<PY
import json
d=json.load(open("gpurun_out/ab_pf$pf.json"))
print("prefetch $pf: ms/step", round(d["ms_per_step"],3), "body avg_us", round(d["roofline"]["avg_us"],2), "rate", round(d["stages"]["lcp_rate"]["ms_total"]*1e3/d["stages"]["lcp_rate"]["regions"],2), "iters", sum(d["lcp_iters_per_step"]))
PY
done
MB200_BODY_PREFETCH=512 timeout 300 python -m pytest tests/test_gpu_collision.py tests/test_gpu_step.py tests/test_gpu_primitives.py -x -q 2>&1 | tail -3
timeout 100 python profiles/bench_primitives.py 2>/dev/null | head -7 | cut -c1-200 | tee gpurun_out/bench_primitives2.jsonl
