set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 8 > gpurun_out/gpu_tests_v12.txt; cat gpurun_out/gpu_tests_v12.txt
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
python bench.py > gpurun_out/r01_bench_v12.json 2> gpurun_out/bench_v12.err || tail -20 gpurun_out/bench_v12.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r01_bench_ref_v12.json 2> gpurun_out/bench_ref_v12.err
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"warp_|rotate_|encode_|decode_|fliplr|transform_|rsn_|oks_|coco_|pck_|aid_" -c 200 --csv --log-file gpurun_out/r01_launches_v12.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-fused > gpurun_out/ncu_v12.log 2>&1
tail -n 1 gpurun_out/r01_bench_v12.json gpurun_out/r01_bench_ref_v12.json; wc -l gpurun_out/r01_launches_v12.csv
rm -f gpurun_out/r01_configs_v12.jsonl
for c in 1 2 3 4 5; do timeout 300 python bench.py --config $c --steps 20 --warmup 3 --no-cpu --no-e2e 2>/dev/null | tail -n 1 >> gpurun_out/r01_configs_v12.jsonl; done
