for f in "" "--no-pdl" "--no-graphs" "--no-graphs --no-pdl" "--attn-splits 4" "--attn-splits 1"; do
  echo "== $f"
  timeout 200 python bench.py --steps 256 --warmup 8 --no-cpu-baseline $f 2>&1 | tail -1 > /tmp/o.json
  python -c "import json; d=json.load(open('/tmp/o.json')); print(d['ms_per_step'], d['value'], d['roofline']['frac'], d['ttft_ms'], d['e2e']['value'])" || tail -3 /tmp/o.json
done
