# usage: bash tools/n8_sweep.sh N   (sweep of NCCL settings at N GPUs)
N=${1:-8}
run() { echo "== $*"; env "$@" timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 2>&1 | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['ms_per_step'], d['value'], d['e2e']['value'])"; }
run NCCL_ALGO=NVLS
run NCCL_ALGO=Ring
