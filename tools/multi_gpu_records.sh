set -x
mkdir -p gpurun_out/multi
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 2 4 8; do
  $TR --nproc-per-node $n --master-port $((29500+n)) bench.py --gpus $n --config stress64 --steps 60 --warmup 3 > gpurun_out/multi/stress64_n$n.json 2> gpurun_out/multi/stress64_n$n.err
  tail -c 300 gpurun_out/multi/stress64_n$n.json
done
$TR --nproc-per-node 8 --master-port 29600 bench.py --gpus 8 --config train16 --steps 30 --warmup 3 > gpurun_out/multi/train16_n8.json 2> gpurun_out/multi/train16_n8.err
tail -c 300 gpurun_out/multi/train16_n8.json
$TR --nproc-per-node 8 --master-port 29601 bench.py --gpus 8 --steps 200 --warmup 5 > gpurun_out/multi/car16_n8_200.json 2> gpurun_out/multi/car16_n8_200.err
tail -c 300 gpurun_out/multi/car16_n8_200.json
