#!/bin/bash
# 2-GPU diagnosis of the transition gather: bench at N=2 with the gather off / pack only / full, NCCL vs copy-engine pulls
O=gpurun_out; mkdir -p $O
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 --steps 20 --warmup 3 "${@:2}" 2>$O/scale2_err.log | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('${*:2}', 'value=%.1fM ms=%.3f e2e=%.1fM' % (d['value']/1e6, d['ms_per_step'], d['e2e']['value']/1e6))" || tail -5 $O/scale2_err.log; }
timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
CMBPO_TRANSPORT=p2p timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/nccl_gather_check.py 2>&1 | grep -E "equal_to|Error|error" | head
python tools/pack_bench.py; python tools/pack_bench.py humanoidsafe_iv_gae_H12_125k
timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('N=1 value=%.1fM ms=%.3f e2e=%.1fM' % (d['value']/1e6, d['ms_per_step'], d['e2e']['value']/1e6))"
timeout 200 python bench.py --workload humanoidsafe_iv_gae_H12_125k --steps 5 --warmup 3 --no-cpu 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('iv N=1 value=%.1fM ms=%.3f' % (d['value']/1e6, d['ms_per_step']))"
run 29602 --gather pack
run 29603 --gather full --nccl-max-ctas 8
run 29605 --gather full --transport p2p
run 29606 --gather full --transport p2p --workload halfcheetah_schedule_L10_1M --steps 5
run 29608 --gather none --workload halfcheetah_schedule_L10_1M --steps 5
