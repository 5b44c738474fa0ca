# usage: bash tools/gpu_replicas.sh N   (inside gpurun --gpus N): the driver's own scaling command
N=${1:-2}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/scale_replica_n$N.json 2> gpurun_out/scale_replica_n$N.err
echo "rc=$?"; grep '^{"metric"' gpurun_out/scale_replica_n$N.json | cut -c1-160
