#!/bin/bash
# round 2, call 6 (8 GPUs): 4- and 8-rank parity (strips + METIS, peer + NCCL halo), config 5 on 8 and 4 GPUs
mkdir -p gpurun_out
(nvidia-smi topo -m; lscpu | grep -E "Model name|Socket|NUMA|^CPU\(s\)") > gpurun_out/r02_topo_8gpu.txt 2>&1
timeout 600 python -m pytest tests/test_multigpu_nccl.py -m gpu -q -s -k "w4 or w8" 2>&1 | grep -v "^$" | cut -c1-1200 | tail -30 > gpurun_out/r02_multigpu_8gpu_pytest.log
tail -3 gpurun_out/r02_multigpu_8gpu_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r02_scale_n8.json 2> gpurun_out/r02_scale_n8.err
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 4 --steps 5 --warmup 3 --no-secondary > gpurun_out/r02_scale_n4.json 2> gpurun_out/r02_scale_n4.err
python - <<'PY'
import json
for n in (8, 4):
    try:
        d=json.loads(open('gpurun_out/r02_scale_n%d.json'%n).read().strip().split('\n')[-1])
        print('N=%d'%n, round(d['value']/1e9,1), 'G frac', round(d['roofline']['frac'],3), 'pm', d['parity_multi'] and d['parity_multi'].get('max_rel_err'), 'sec', d['secondary'] and d['secondary'].get('value'), 'setup', round(d['host_setup_s']), 'e2e', d['e2e'] and round(d['e2e']['value']/1e9,1), d['e2e'] and round(d['e2e']['host_gbs_per_direction_per_gpu'],1))
        print(d['config']['partition_stats'])
    except Exception as e:
        print('N=%d no line'%n, e); print(open('gpurun_out/r02_scale_n%d.err'%n).read()[-1500:])
PY
