q() { python bench.py --steps 3 --warmup 2 --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['value']), round(d['e2e']['value']), {k:round(v,1) for k,v in d['stage_ms_per_step'].items()}, d['e2e']['matches_device_path'])"; }
q base
for S in 1024 2048; do for rb in 3 4; do for tsb in 1 2 3; do
  export CLIMA_RT_PIPELINE_COLS=$S CLIMA_RT_RORR_BLOCKS_PER_SM=$rb CLIMA_RT_PIPELINE_TS_BLOCKS=$tsb
  q "S=$S rb=$rb ts=$tsb"
done; done; done
export CLIMA_RT_PIPELINE_COLS=1024 CLIMA_RT_PIPELINE_TS_BLOCKS=2; unset CLIMA_RT_RORR_BLOCKS_PER_SM; q "S=1024 rb=- ts=2"
export CLIMA_RT_PIPELINE_COLS=512 CLIMA_RT_RORR_BLOCKS_PER_SM=3 CLIMA_RT_PIPELINE_TS_BLOCKS=2 CLIMA_RT_RORR_CHAIN_LEN=256; q "S=512 rb=3 ts=2 L=256"
