O=gpurun_out/final
mkdir -p $O
timeout 400 python bench.py --steps 20 --warmup 5 --detail $O/per_layer_detail.tsv > $O/bench_resnet50_n1.json 2> $O/bench_resnet50_n1.err
for c in vgg16 alexnet lenet5; do
  timeout 300 python bench.py --config $c --steps 20 --warmup 5 > $O/bench_$c.json 2> $O/bench_$c.err
done
timeout 400 python bench.py --config conv_sweep --steps 5 --warmup 3 --detail $O/conv_sweep.tsv > $O/bench_conv_sweep.json 2> $O/bench_conv_sweep.err
timeout 300 python tools/bench_models.py $O/other_models.tsv > $O/bench_models.log 2>&1
timeout 200 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > $O/plain_step.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 8000 --csv \
    --log-file $O/step_capture.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > $O/ncu_step.log 2>&1
python tools/step_profile.py $O/step_capture.csv $O/r02 > $O/step_profile.log 2>&1
gzip -f -9 $O/step_capture.csv
for f in $O/bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d.get("value", 0), 1), d.get("unit"), "ms/step", round(d.get("ms_per_step", 0), 3), "e2e", (d.get("e2e") or {}).get("value"))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done
head -3 $O/r02_launches_step.txt
