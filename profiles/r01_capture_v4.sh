# Round-1 capture of the bench command (run on the B200 box through gpurun); outputs under gpurun_out/.
set -x
python bench.py > gpurun_out/r01_bench_v4.json 2> gpurun_out/bench_v4.err || exit 1
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r01_bench_ref_v4.json 2> gpurun_out/bench_ref_v4.err
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"warp_|rotate_|encode_|decode_|fliplr|transform_" -c 200 --csv --log-file gpurun_out/r01_launches_v4.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-fused > gpurun_out/ncu_v4.log 2>&1
tail -n 2 gpurun_out/r01_bench_v4.json gpurun_out/r01_bench_ref_v4.json; wc -l gpurun_out/r01_launches_v4.csv
