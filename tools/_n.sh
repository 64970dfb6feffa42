N=$1
mkdir -p gpurun_out
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $N --steps 20 --warmup 3 ) > gpurun_out/n${N}_bench.json 2> gpurun_out/n${N}_bench.err
echo "bench rc=$?" >> gpurun_out/n${N}_bench.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29556 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/n${N}_ref.json 2> gpurun_out/n${N}_ref.err
echo "ref rc=$?" >> gpurun_out/n${N}_ref.err
