set -x
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_final.log 2>&1; tail -3 gpurun_out/pytest_gpu_final.log
python bench.py > gpurun_out/bench_n1_final.json 2> gpurun_out/bench_n1_final.err; tail -2 gpurun_out/bench_n1_final.err
python bench.py --impl reference > gpurun_out/bench_reference_arm.json 2> gpurun_out/bench_reference_arm.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_final.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_bench.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:distmat_kernel -s 1 -c 1 -o gpurun_out/distmat python bench.py --steps 2 --warmup 3 --no-em --no-recon --no-codon > gpurun_out/ncu_distmat.log 2>&1
ncu -i gpurun_out/distmat.ncu-rep --page details > gpurun_out/distmat_ncu_details.txt 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:branch_viterbi_kernel -s 3 -c 1 -o gpurun_out/branch python bench.py --steps 2 --warmup 3 --no-em --no-recon --no-codon > gpurun_out/ncu_branch.log 2>&1
ncu -i gpurun_out/branch.ncu-rep --page details > gpurun_out/branch_ncu_details.txt 2>&1
ls -la gpurun_out | tail -12
