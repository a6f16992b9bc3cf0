mkdir -p gpurun_out/agg
for mode in p2p nvls; do
  if [ $mode = nvls ]; then export FEDB200_AGG=nvls; else unset FEDB200_AGG; fi
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29768 bench.py --config wide --clients-per-gpu 32 --gpus 8 --steps 5 --warmup 2 2> gpurun_out/agg/wide_${mode}_8.err | grep "^{" > gpurun_out/agg/wide_${mode}_8.json
done
python - <<'PY'
import json, glob, os
for f in sorted(glob.glob("gpurun_out/agg/wide_*_8.json")):
    try:
        r = json.load(open(f)); print(os.path.basename(f), "ms/step", round(r["ms_per_step"], 4), "warm", round(r["warm_no_flush"]["ms_per_step"], 4), "acc", r["final_global_accuracy_train_shards"])
    except Exception as e:
        print(f, "FAILED", e)
PY
