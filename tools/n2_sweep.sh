run() { echo "== $*"; env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 2>&1 | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['ms_per_step'], d['value'], d['e2e']['value'])"; }
run OWL_COMM_CTAS=0
run OWL_COMM_CTAS=16
run OWL_COMM_CTAS=16 OWL_NCCL_CTAS=16
run OWL_COMM_CTAS=32 OWL_NCCL_CTAS=32
run OWL_COMM_CTAS=24
