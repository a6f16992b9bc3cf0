# Round-2 scaling session on one 8-GPU box: headline config at 1/2/4/8 GPUs (driver protocol 20/5 and a longer 300/20 run),
# 1 / 4 / 16 clients per GPU at 8 GPUs, the sklearn-avg and grid configs at 1 and 8 GPUs, the wide config at 1 and 8.
#   bash tools/scale_r2.sh            (everything)        bash tools/scale_r2.sh quick   (headline only)      bash tools/scale_r2.sh final
OUT=gpurun_out/scale_r2
mkdir -p $OUT
run() {  # run N name args...
  N=$1; name=$2; shift 2
  if [ $N -eq 1 ]; then
    timeout 600 python bench.py --gpus 1 "$@" > $OUT/$name.json 2> $OUT/$name.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2970$N bench.py --gpus $N "$@" 2> $OUT/$name.err | grep "^{" > $OUT/$name.json
  fi
}
if [ "$1" = "final" ]; then   # the short session: 300-step headline at every N, driver protocol at 8, grid at 8, sklearn at 4
  for N in 1 2 4 8; do run $N income_${N} --steps 300 --warmup 20; done
  run 8 income_8_driver --steps 20 --warmup 5
  run 8 grid_8 --config grid --steps 2 --warmup 1; run 4 sklearn_4 --config sklearn
else
for N in 1 2 4 8; do run $N income_${N}_driver --steps 20 --warmup 5; run $N income_${N} --steps 300 --warmup 20; done
fi
if [ "$1" != "quick" ] && [ "$1" != "final" ]; then
  for C in 4 16; do run 8 income_8_cpg$C --steps 200 --warmup 10 --clients-per-gpu $C; run 1 income_1_cpg$C --steps 200 --warmup 10 --clients-per-gpu $C; done
  run 8 sklearn_8 --config sklearn; run 4 sklearn_4 --config sklearn
  run 8 grid_8 --config grid; run 4 grid_4 --config grid
  run 1 wide_1 --config wide --steps 10 --warmup 3; run 8 wide_8 --config wide --steps 10 --warmup 3
  run 8 nccl_8 --impl nccl --steps 300 --warmup 20; run 1 nccl_1 --impl nccl --steps 300 --warmup 20
fi
python - <<'PY'
import json, glob, os
for f in sorted(glob.glob("gpurun_out/scale_r2/*.json")):
    name = os.path.basename(f)[:-5]
    try:
        r = json.load(open(f))
        w = r.get("warm_no_flush", {}).get("ms_per_step")
        print(name.ljust(22), "ms/step", round(r["ms_per_step"], 4), r["metric"], round(r["value"], 2), "warm", None if w is None else round(w, 4),
              "e2e", round(r["e2e"]["ms_per_step"], 4) if "e2e" in r else None, r["clocks"]["sm_mhz"], r["clocks"]["reasons"])
    except Exception as e:
        print(name, "FAILED", e)
PY
