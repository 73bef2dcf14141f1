for env in "" "B200LLM_TC_NO_PDL=1"; do
  for cfg in "llama3.2:1b 64" "llama3.2 64" "llama3.2 8"; do
    set -- $cfg
    echo "== $env $1 bs$2"
    env $env timeout 300 python bench.py --model $1 --bs $2 --steps 64 --warmup 4 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],3), round(d['value']), round(d['roofline']['frac'],3), 'ttft', round(d['ttft_ms'],1))"
  done
done
