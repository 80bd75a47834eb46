cd $GRAFT_REPO_ROOT
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { tag=$1; port=$2; shift; shift; timeout 400 env "$@" $TR --master-port $port bench.py --gpus $N --steps 6 --warmup 3 --no-e2e $EXTRA > gpurun_out/s8_$tag.json 2> gpurun_out/s8_$tag.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/s8_$tag.json").read().strip().splitlines()[-1])
    print("$tag", round(d["ms_per_step"],3), {k:round(v["us"]) for k,v in d["stages"].items()}, "nvl", round(d["roofline_nvlink"]["achieved"]))
except Exception as e:
    print("$tag FAILED", open("gpurun_out/s8_$tag.err").read()[-600:])
PY
}
EXTRA="" run A_overlap 29601 X=1
EXTRA="--overlap 0" run B_serial 29602 X=1
