#!/bin/bash
# Round-2 evidence capture (one B200): tests, both bench arms, ncu launch list of the bench command, full ncu
# captures of the kernels listed below.  Run from the repository root under gpurun; outputs land in gpurun_out/.
set -x
V=${1:-v3}
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 6 > gpurun_out/r02_gpu_tests_$V.txt; cat gpurun_out/r02_gpu_tests_$V.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/r02_bench_$V.json 2> gpurun_out/bench_$V.err || tail -20 gpurun_out/bench_$V.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_ref_$V.json 2> gpurun_out/bench_ref_$V.err
# launch list of the bench command (kernel times are cold-cache and serialised under ncu: shares, not absolutes)
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    -k regex:"warp_|rotate_|encode_|decode_" -c 200 --csv --log-file gpurun_out/r02_launches_$V.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-fused --no-configs --no-sweep > gpurun_out/ncu_launch_$V.log 2>&1
python profiles/make_traffic.py gpurun_out/r02_launches_$V.csv 2 16384 > gpurun_out/traffic_$V.log 2>&1; cp profiles/traffic.json gpurun_out/traffic_$V.json
# full captures (one launch each, after warm-up launches); FULL=0 skips them
[ "${FULL:-1}" = "0" ] && exit 0
cap() { # name kernel-regex skip command...
  n=$1; k=$2; s=$3; shift 3
  ncu --set full --import-source on --clock-control none -k regex:$k --launch-skip $s -c 1 -o gpurun_out/r02_${n}_$V -f "$@" > gpurun_out/ncu_$n.log 2>&1
  tail -1 gpurun_out/ncu_$n.log
}
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-configs --no-sweep"
cap crop_tma warp_affine_u8_tma_kernel 3 $B --no-fused
cap encode_offset encode_offset_kernel 3 $B --no-fused
cap decode_offset decode_offset_team_kernel 3 $B --no-fused
cap crop_f32 "warp_affine_u8_tma_kernel.*CropOutF32" 1 $B
cap pck_dist pck_dist_kernel 1 $B
cap decode_quarter decode_quarter_kernel 1 $B
cap encode_gaussian encode_gaussian_kernel 3 $B --no-fused --config 1
cap decode_gauss_96x72 decode_gaussian_pipe_kernel 3 $B --no-fused --config 3
cap crop_tma_384 warp_affine_u8_tma_kernel 3 $B --no-fused --config 3
ls -la gpurun_out/*.ncu-rep | wc -l
