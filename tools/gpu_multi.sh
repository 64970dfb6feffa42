#!/bin/bash
# multi-GPU evidence: P2P test, probe matrix, NVLink counters under ncu, replica scaling, NVML re-enumeration
N=${1:-2}
T=m${N}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/${T}_topo.txt 2>&1
( time timeout 900 python -m pytest tests/test_gpu_probes.py -m gpu -x -q ) > gpurun_out/${T}_pytest_probes.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${T}_pytest_probes.log
timeout 900 python tools/probe.py --gpus $N --reps 10 > gpurun_out/${T}_probe.json 2> gpurun_out/${T}_probe.err
timeout 300 ./tools/bin/nvml_reenum 30 > gpurun_out/${T}_nvml_reenum.json 2> gpurun_out/${T}_nvml_reenum.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 \
  bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
echo "bench rc=$?" >> gpurun_out/${T}_bench.err
timeout 900 ncu --metrics nvlrx__bytes.sum,nvltx__bytes.sum,nvlrx__bytes_data_user.sum,nvltx__bytes_data_user.sum,gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_srcunit_ltcfabric.sum \
  --clock-control none -k regex:stream_kernel --csv --log-file gpurun_out/${T}_p2p_nvlink.csv \
  python tools/probe.py --gpus 2 --reps 4 --hbm-bytes 1073741824 --p2p-bytes 1073741824 > gpurun_out/${T}_ncu_probe.log 2>&1
