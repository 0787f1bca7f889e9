# usage: bash tools/n2_sweep.sh [N]   (gradient-exchange settings at N GPUs; prints ms/step, tokens/s, e2e tokens/s)
N=${1:-2}
run() { echo "== $*"; env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 2>&1 | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['ms_per_step'], d['value'], d['e2e']['value'])"; }
run OWL_COMM_CTAS=16
run OWL_COMM_DTYPE=bf16 OWL_COMM_CTAS=16
run OWL_COMM_DTYPE=bf16 OWL_COMM_CTAS=8 OWL_NCCL_CTAS=8
run OWL_COMM_DTYPE=bf16 OWL_COMM_CTAS=4 OWL_NCCL_CTAS=4
run OWL_COMM_DTYPE=bf16 OWL_COMM_CTAS=8 OWL_NCCL_CTAS=8 NCCL_ALGO=NVLS
run OWL_COMM_CTAS=8 OWL_NCCL_CTAS=8
