set -x
python bench.py > gpurun_out/r01_bench_v3.json 2> gpurun_out/bench_v3.err || exit 1
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r01_bench_ref_v3.json 2> gpurun_out/bench_ref_v3.err
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"warp_|rotate_|encode_|decode_|fliplr|transform_" -c 200 --csv --log-file gpurun_out/r01_launches_v3.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_v3.log 2>&1
tail -n 3 gpurun_out/r01_bench_v3.json gpurun_out/r01_bench_ref_v3.json; wc -l gpurun_out/r01_launches_v3.csv
