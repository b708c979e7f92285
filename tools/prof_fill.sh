set -x
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 1 --cpu-pairs 1 > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fill_narrow -s 1 -c 1 -o gpurun_out/fill_narrow -f python bench.py --steps 1 --warmup 1 --cpu-pairs 1 > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out
