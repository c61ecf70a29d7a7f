N=$1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload c4 --steps ${STEPS:-5} --warmup 3 2>gpurun_out/c4_$N.err > gpurun_out/c4_$N.json; tail -2 gpurun_out/c4_$N.err
python -c "
import json
d=json.loads(open('gpurun_out/c4_$N.json').read()); r=d['roofline']; print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], r['frac'], r['probe']['ms'], r['list_scan_ms'], r['achieved'], d['config']['recall_at_100'], d.get('stages_ms'))"
