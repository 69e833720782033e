mkdir -p gpurun_out; N=$1; shift
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --workload bigscene_10m_4k_4spp --no-ab --no-cpu-baseline"
for cfg in "$@"; do
  IFS=, read d c f k <<< "$cfg"
  RT_FRAMES_IN_FLIGHT=$d RT_CTAS_PER_SM=$c RT_DYN_FRACTION=$f RT_DYN_CHUNK=${k:-8} timeout 300 $T 2>/dev/null | python -c "
import sys, json
for line in sys.stdin:
    if line.startswith('{'):
        j = json.loads(line); print('depth,ctas,dyn,chunk=$cfg', 'ms', round(j['ms_per_step'], 3), 'e2e_ms', round(j['e2e'].get('ms_per_frame', 0) or 0, 3), 'single', [round(x, 2) for x in j['roofline']['kernel_ms_per_rank']], j['parity']['ok'])
"
done
