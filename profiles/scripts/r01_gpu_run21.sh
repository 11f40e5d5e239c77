# 2-GPU bench line (one rank per GPU under torchrun, batch-sharded replicas, NCCL gather of the codes only)
set -x
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 2 --warmup 3 --skip-cpu --e2e-steps 1 > gpurun_out/r21_n2.json 2> gpurun_out/r21_n2.err; echo "n2 exit $?"
tail -1 gpurun_out/r21_n2.json | cut -c1-700
tail -5 gpurun_out/r21_n2.err | cut -c1-300
