set -x
mkdir -p gpurun_out/r2
python bench.py --steps 200 --grid rows --no-cpu-baseline --no-rpn > gpurun_out/r2/bench_car16_rows.json 2>/dev/null
python bench.py --steps 200 --in-flight 1 --no-cpu-baseline --no-rpn > gpurun_out/r2/bench_car16_p1.json 2>/dev/null
python bench.py --steps 100 --data kitti --no-cpu-baseline --no-rpn > gpurun_out/r2/bench_car16_kitti.json 2> gpurun_out/r2/bench_kitti.err
python bench.py --steps 100 --config ped16 --no-cpu-baseline --no-rpn > gpurun_out/r2/bench_ped16.json 2>/dev/null
python bench.py --steps 30 --config stress64 --no-cpu-baseline > gpurun_out/r2/bench_stress64.json 2>/dev/null
python bench.py --steps 200 --output nhwc_bf16 --no-cpu-baseline --no-rpn > gpurun_out/r2/bench_car16_output_nhwc_bf16.json 2>/dev/null
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 420 --csv --log-file gpurun_out/r2/launches.csv python bench.py --steps 2 --warmup 3 --in-flight 1 --no-cpu-baseline --no-rpn > gpurun_out/r2/ncu_launches.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"k_conv3d_bn_relu|k_vox_svfe_packed" --launch-skip 12 -c 4 -f -o gpurun_out/r2/full python bench.py --steps 2 --warmup 3 --in-flight 1 --no-cpu-baseline --no-rpn > gpurun_out/r2/ncu_full.log 2>&1
ls -la gpurun_out/r2
