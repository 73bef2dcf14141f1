# decode numbers of every model / batch the docs quote (bs1 = persistent kernel, bs>1 = tcgen05 batch path)
for cfg in "llama3.2 1" "llama3.2:1b 1" "duckdb-nsql 1" "llama3.2 8" "llama3.2 64" "llama3.2:1b 64"; do
  set -- $cfg
  echo "== $1 bs$2"
  timeout 400 python bench.py --model $1 --bs $2 --steps 128 --warmup 4 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],4), 'ms', round(d['value'],1), 'tok/s', 'frac', round(d['roofline']['frac'],3), 'ttft', round(d['ttft_ms'],1), 'e2e', round(d['e2e']['value'],1))"
done
