# ours only: headline config at the GPU counts in $1 (default "1 2 4 8") and BASELINE config 5 (wide) at the counts in $2 (default "1 8")
# e.g. on a 2-GPU box:  bash tools/scale_ours.sh "1 2" "1"      on an 8-GPU box:  bash tools/scale_ours.sh "8" "8"
NS=${1:-"1 2 4 8"}
WIDE=${2-"1 8"}
mkdir -p gpurun_out/scale3
run() {  # run N outfile args...
  N=$1; out=$2; shift 2
  if [ $N -eq 1 ]; then
    python bench.py --gpus 1 "$@" > gpurun_out/scale3/$out.json 2> gpurun_out/scale3/$out.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2970$N bench.py --gpus $N "$@" 2> gpurun_out/scale3/$out.err | grep "^{" > gpurun_out/scale3/$out.json
  fi
}
for N in $NS; do run $N ours_$N --steps 2000 --warmup 20; done
for N in $WIDE; do run $N wide_$N --config wide --steps 10 --warmup 3; done
python - <<'PY'
import json, glob, os
for f in sorted(glob.glob("gpurun_out/scale3/*.json")):
    name = os.path.basename(f)[:-5]
    try:
        r = json.load(open(f))
        print(name, "ms/step", round(r["ms_per_step"], 4), "M samples/s", round(r["value"] / 1e6, 2), "warm", round(r["warm_no_flush"]["ms_per_step"], 4),
              "e2e", round(r["e2e"]["ms_per_step"], 4), "acc", r["final_global_accuracy_train_shards"], r["clocks"]["sm_mhz"], r["clocks"]["reasons"])
    except Exception as e:
        print(name, "FAILED", e)
PY
