# Round-2 evidence run on one B200 (gpurun): GPU tests, benches, ncu launch list, ncu --set full of the four big kernels.
set -x
mkdir -p gpurun_out/r2
python -m pytest tests -m gpu -q > gpurun_out/r2/pytest_gpu.log 2>&1; tail -3 gpurun_out/r2/pytest_gpu.log
python bench.py --steps 200 > gpurun_out/r2/bench_car16.json 2> gpurun_out/r2/bench_car16.err; tail -c 300 gpurun_out/r2/bench_car16.json
python bench.py --steps 200 --grid rows --no-cpu-baseline > gpurun_out/r2/bench_car16_rows.json 2>/dev/null
python bench.py --steps 200 --in-flight 1 --no-cpu-baseline > gpurun_out/r2/bench_car16_p1.json 2>/dev/null
python bench.py --steps 100 --data kitti --no-cpu-baseline > gpurun_out/r2/bench_car16_kitti.json 2> gpurun_out/r2/bench_kitti.err
python bench.py --steps 100 --config ped16 --no-cpu-baseline > gpurun_out/r2/bench_ped16.json 2>/dev/null
python bench.py --steps 30 --config train16 > gpurun_out/r2/bench_train16.json 2>/dev/null
python bench.py --steps 30 --config stress64 --no-cpu-baseline > gpurun_out/r2/bench_stress64.json 2>/dev/null
python bench.py --impl reference --steps 4 --warmup 1 > gpurun_out/r2/bench_reference_arm.json 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none -c 420 --csv --log-file gpurun_out/r2/launches.csv python bench.py --steps 2 --warmup 3 --in-flight 1 --no-cpu-baseline > gpurun_out/r2/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_conv3d_bn_relu|k_vox_svfe_packed" --launch-skip 12 -c 4 -f -o gpurun_out/r2/full python bench.py --steps 2 --warmup 3 --in-flight 1 --no-cpu-baseline > gpurun_out/r2/ncu_full.log 2>&1
ls -la gpurun_out/r2
