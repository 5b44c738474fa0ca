# usage: bash tools/gpu_multi.sh N   (inside gpurun --gpus N)
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611"
nvidia-smi -L
timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/scale_replica_n$N.json 2> gpurun_out/scale_replica_n$N.err; echo "replica rc=$?"; tail -c 300 gpurun_out/scale_replica_n$N.err
cut -c1-330 gpurun_out/scale_replica_n$N.json
timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 --shard > gpurun_out/scale_shard_ultra22_n$N.json 2> gpurun_out/scale_shard_ultra22_n$N.err; echo "shard ultra rc=$?"; tail -c 300 gpurun_out/scale_shard_ultra22_n$N.err
cut -c1-330 gpurun_out/scale_shard_ultra22_n$N.json
timeout 900 $TR bench.py --gpus $N --steps 3 --warmup 3 --shard --workload groth16 --logn 24 > gpurun_out/scale_shard_g16_24_n$N.json 2> gpurun_out/scale_shard_g16_24_n$N.err; echo "shard g16 2^24 rc=$?"; tail -c 300 gpurun_out/scale_shard_g16_24_n$N.err
cut -c1-330 gpurun_out/scale_shard_g16_24_n$N.json
