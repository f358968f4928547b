set -x
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/plain_step.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 8000 --csv --log-file gpurun_out/step_capture.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/ncu_step.log 2>&1
python tools/step_profile.py gpurun_out/step_capture.csv gpurun_out/r02 > gpurun_out/step_profile.log 2>&1; tail -5 gpurun_out/step_profile.log
python tools/prof_conv.py fprop 256 56 56 64 64 3 1 1 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:tc_kernel -s 3 -c 1 -o gpurun_out/r02_tc_c64 python tools/prof_conv.py fprop 256 56 56 64 64 3 1 1 > gpurun_out/ncu_c64.log 2>&1
python tools/bench_attention.py --once > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:fa:: -s 3 -c 3 -o gpurun_out/r02_fa python tools/bench_attention.py --once > gpurun_out/ncu_fa.log 2>&1
ls -la gpurun_out | tail -12
