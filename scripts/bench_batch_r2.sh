# batch-64 decode under the round-2 switches of the tcgen05 GEMM path: weight tiles as 16 KB bulk copies out of the tile-major copy
# (B200LLM_TC_TM, default 1), weight tiles requested before griddepcontrol.wait
# (B200LLM_TC_W_EARLY, default 1) and the split count of the batched attention (B200LLM_BATCH_SPLITS, default: by grid size)
for env in "" "B200LLM_TC_TM=0" "B200LLM_TC_W_EARLY=0" "B200LLM_TC_TM=0 B200LLM_TC_W_EARLY=0" "B200LLM_BATCH_SPLITS=2" "B200LLM_BATCH_SPLITS=3" "B200LLM_TC_NO_PDL=1"; do
  for cfg in "llama3.2:1b 64" "llama3.2 64"; do
    set -- $cfg
    echo "== ${env:-default} $1 bs$2"
    env $env timeout 300 python bench.py --model $1 --bs $2 --steps 64 --warmup 4 --no-cpu-baseline --no-batch 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],3), round(d['value']), round(d['roofline']['frac'],3), 'ttft', round(d['ttft_ms'],2))"
  done
done
