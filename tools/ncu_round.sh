set -x
mkdir -p gpurun_out
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__throughput.avg.pct_of_peak_sustained_elapsed
python profiles/run_register_once.py > /dev/null 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_${TAG:-r02f}.csv python profiles/run_register_once.py > gpurun_out/ncu1.log 2>&1
ncu --metrics $M --clock-control none --profile-from-start off --csv --log-file gpurun_out/kmetrics_${TAG:-r02f}.csv python profiles/run_register_once.py > gpurun_out/ncu2.log 2>&1
python tools/gemm_bench.py --modes 3 --layers conv6 --reps 5
ncu --set full --clock-control none --import-source on -k regex:tc2_kernel -s 2 -c 1 -o gpurun_out/conv6_${TAG:-r02f} python tools/gemm_bench.py --modes 3 --layers conv6 --reps 2 > gpurun_out/ncu3.log 2>&1
ls -la gpurun_out
