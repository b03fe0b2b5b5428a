#!/bin/bash
# round 2, call 5 (2 GPUs): multi-GPU parity on hardware (peer-store and NCCL halo, strips and METIS) + config 5 split 2 ways
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r02_topo_2gpu.txt 2>&1
timeout 900 python -m pytest tests/test_multigpu_nccl.py -m gpu -q -s -rs 2>&1 | grep -v "^$" | tail -40 > gpurun_out/r02_multigpu_2gpu_pytest.log
tail -4 gpurun_out/r02_multigpu_2gpu_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02_scale_n2.json 2> gpurun_out/r02_scale_n2.err
tail -c 1500 gpurun_out/r02_scale_n2.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/r02_scale_n2.json').read().strip().split('\n')[-1])
    print('N=2', round(d['value']/1e9,1), 'G  frac', round(d['roofline']['frac'],3), 'parity_multi', d['parity_multi'], 'secondary', d['secondary'] and d['secondary'].get('value'), 'setup', d['host_setup_s'], 'e2e', d['e2e'] and d['e2e']['value'])
    print(d['config']['partition_stats'])
except Exception as e:
    print('no line', e)
PY
