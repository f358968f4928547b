TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
run() { # name, config, env...
  name=$1; cfg=$2; shift 2
  env "$@" timeout 200 $TR --master-port $((29700 + RANDOM % 200)) bench.py --config $cfg --gpus 2 --steps 10 --warmup 5 > gpurun_out/dps_$name.json 2>/dev/null
  python -c "
import json
try:
    d=json.load(open('gpurun_out/dps_$name.json')); print('$name', round(d['value']), round(d['ms_per_step'],3))
except Exception as e: print('$name FAILED')"
}
timeout 200 python bench.py --config vgg16 --steps 10 --warmup 5 --no-cpu-baseline > gpurun_out/dps_vgg_n1.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/dps_vgg_n1.json')); print('vgg_n1', round(d['value']), round(d['ms_per_step'],3))"
run vgg_default vgg16 X=1
run vgg_ctas8 vgg16 NCCL_MAX_CTAS=8
run vgg_ctas4 vgg16 NCCL_MAX_CTAS=4
run vgg_ctas2 vgg16 NCCL_MAX_CTAS=2
run vgg_bucket100 vgg16 RNN_DP_BUCKET_MB=100
run vgg_bucket8 vgg16 RNN_DP_BUCKET_MB=8
run r50_default resnet50 X=1
run r50_ctas4 resnet50 NCCL_MAX_CTAS=4
run r50_tail0 resnet50 RNN_DP_TAIL_MB=0
