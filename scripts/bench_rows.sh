for m in llama3.2 duckdb-nsql; do for r in 64 128 256; do
  echo "== $m max_rows $r"
  timeout 300 python bench.py --model $m --max-rows $r --steps 64 --warmup 4 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value'],1), 'tok/s', 'ttft', round(d['ttft_ms'],2), 'e2e', round(d['e2e']['value'],1))"
done; done
