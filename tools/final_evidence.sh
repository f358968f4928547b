#!/bin/bash
# Round-2 evidence on ONE B200 (everything under its own timeout; output in gpurun_out/final/):
#   part "a": the whole GPU test-suite, smoke(), every bench.py configuration, the reference arm, micro-benchmarks
#   part "b": ncu passes -- launch list + DRAM bytes of one eager training step, --set full of the C64 3x3 convolution,
#             of the fused attention kernels and of the stem pool kernels (each only after the plain run exited 0)
O=gpurun_out/final
mkdir -p $O
part=${1:-a}
if [ "$part" = "a" ]; then
  bash tools/run_gpu_tests.sh tests/test_*.py > $O/gpu_tests_summary.txt 2>&1
  cp gpurun_out/gpu_tests.log $O/gpu_tests.log
  tail -25 $O/gpu_tests_summary.txt
  timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
  timeout 400 python bench.py --steps 20 --warmup 5 --detail $O/per_layer_detail.tsv > $O/bench_resnet50_n1.json 2> $O/bench_resnet50_n1.err
  timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err
  for c in vgg16 alexnet lenet5; do
    timeout 300 python bench.py --config $c --steps 20 --warmup 5 > $O/bench_$c.json 2> $O/bench_$c.err
  done
  timeout 400 python bench.py --config conv_sweep --steps 5 --warmup 3 --detail $O/conv_sweep.tsv > $O/bench_conv_sweep.json 2> $O/bench_conv_sweep.err
  timeout 200 python tools/bench_attention.py $O/attention_fused_vs_unfused.tsv > $O/bench_attention.log 2>&1
  timeout 200 python tools/bench_bn_bwd.py $O/bn_bwd.tsv > $O/bench_bn_bwd.log 2>&1
  timeout 100 python tools/bench_pool.py $O/pool.tsv > $O/bench_pool.log 2>&1
  timeout 300 python tools/bench_models.py $O/other_models.tsv > $O/bench_models.log 2>&1
  for f in $O/bench_*.json; do python - "$f" <<'EOF'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], d.get("metric"), round(d.get("value", 0), 1), d.get("unit"), "ms/step", round(d.get("ms_per_step", 0), 3),
          "e2e", (d.get("e2e") or {}).get("value"))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
EOF
  done
else
  timeout 200 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > $O/plain_step.log 2>&1 && \
  timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 8000 --csv \
      --log-file $O/step_capture.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > $O/ncu_step.log 2>&1
  python tools/step_profile.py $O/step_capture.csv $O/r02 > $O/step_profile.log 2>&1; tail -3 $O/step_profile.log
  rm -f $O/step_capture.csv.keep; gzip -f -9 $O/step_capture.csv
  timeout 100 python tools/prof_conv.py fprop 256 56 56 64 64 3 1 1 > /dev/null 2>&1 && \
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:tc_kernel -s 3 -c 1 -f -o $O/r02_tc_c64 \
      python tools/prof_conv.py fprop 256 56 56 64 64 3 1 1 > $O/ncu_c64.log 2>&1
  timeout 100 python tools/bench_attention.py --once > /dev/null 2>&1 && \
  timeout 300 ncu --set full --clock-control none --import-source on -k "regex:^(fwd|dq|dkdv)_kernel" -c 6 -f \
      -o $O/r02_fa python tools/bench_attention.py --once > $O/ncu_fa.log 2>&1
  timeout 100 python tools/bench_pool.py > /dev/null 2>&1 && \
  timeout 300 ncu --set full --clock-control none --import-source on -k "regex:maxpool_.*march" -s 6 -c 3 -f \
      -o $O/r02_pool python tools/bench_pool.py > $O/ncu_pool.log 2>&1
  ls -la $O | tail -15
fi
