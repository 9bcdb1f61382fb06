mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02zd_scale_n$N.json 2> gpurun_out/r02zd_scale_n$N.err; echo "bench N=$N exit $?"; tail -3 gpurun_out/r02zd_scale_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02zd_scale_n$N.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'frac',d['roofline']['frac'])
for k,v in d.get('extra',{}).items(): print(k, v['value'], v['ms_per_step'], v['roofline']['frac'])
e=d['e2e']; print('e2e',e['value'],'without_file',e['without_file']['value'], e['pipeline'].get('file_GBps_while_placing'))
PY
