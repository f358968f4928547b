# One 8-GPU box: DP parity at 8 ranks, ResNet-50 at N=8, VGG-16 at N=1,2,4,8 (back to back, same box).
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 8 --master-port 29601 tools/dp_parity.py gpurun_out/r02_dp_parity_n8.txt > gpurun_out/dp8.log 2>&1; echo dp8_rc=$?
timeout 200 $TR --nproc-per-node 8 --master-port 29602 bench.py --gpus 8 --steps 10 --warmup 5 > gpurun_out/r02_bench_n8.json 2> gpurun_out/r02_bench_n8.err
timeout 200 python bench.py --config vgg16 --gpus 1 --steps 10 --warmup 5 --no-cpu-baseline > gpurun_out/r02_vgg16_n1.json 2>/dev/null
for n in 2 4 8; do
  timeout 200 $TR --nproc-per-node $n --master-port $((29610+n)) bench.py --config vgg16 --gpus $n --steps 10 --warmup 5 > gpurun_out/r02_vgg16_n$n.json 2> gpurun_out/r02_vgg16_n$n.err
done
timeout 200 python bench.py --gpus 1 --steps 10 --warmup 5 --no-cpu-baseline > gpurun_out/r02_bench_n1_samebox.json 2>/dev/null
for f in gpurun_out/r02_bench_n1_samebox.json gpurun_out/r02_bench_n8.json gpurun_out/r02_vgg16_n1.json gpurun_out/r02_vgg16_n2.json gpurun_out/r02_vgg16_n4.json gpurun_out/r02_vgg16_n8.json; do
  python -c "
import json,sys
try:
    d=json.load(open('$f')); print('$f', d['n_gpus'], round(d['value']), round(d['ms_per_step'],3))
except Exception as e: print('$f', 'FAILED', e)"
done
grep -v "^\[rank\|^$\|^\*\*\*\|OMP_NUM" gpurun_out/dp8.log | tail -14
