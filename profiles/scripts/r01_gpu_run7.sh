# 2-GPU torchrun bench (one rank per GPU, NCCL gather of codes only) + the reference arm on the box's host cores
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 2 --warmup 1 --skip-cpu > gpurun_out/r7_n2.json 2> gpurun_out/r7_n2.err; echo "n2 exit $?"
tail -3 gpurun_out/r7_n2.err; cat gpurun_out/r7_n2.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r7_ref.json 2> gpurun_out/r7_ref.err; echo "ref exit $?"
cat gpurun_out/r7_ref.json; tail -3 gpurun_out/r7_ref.err
