set -x
mkdir -p gpurun_out/multi2
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node 8 --master-port 29701 bench.py --gpus 8 --steps 200 --warmup 5 > gpurun_out/multi2/car16_n8_200.json 2> gpurun_out/multi2/car16_n8_200.err
tail -c 300 gpurun_out/multi2/car16_n8_200.json
