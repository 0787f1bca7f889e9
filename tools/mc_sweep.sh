# usage: bash tools/mc_sweep.sh [N]   (in-switch vs copy-engine gradient exchange at N GPUs; prints ms/step, tokens/s)
N=${1:-2}
run() { echo "== $*"; env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 2>&1 | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['ms_per_step'], d['value'], d['e2e']['value'])"; }
shift
if [ $# -gt 0 ]; then for cfg in "$@"; do run $(echo $cfg | tr ',' ' '); done; exit 0; fi
run OWL_EXCHANGE=ce
run OWL_EXCHANGE=mc OWL_MC_CTAS=32
run OWL_EXCHANGE=mc OWL_MC_CTAS=64
run OWL_EXCHANGE=mc OWL_MC_CTAS=148
run OWL_EXCHANGE=mc OWL_MC_CTAS=64 OWL_CE_LANES=2
