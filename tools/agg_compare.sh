# stand-alone cross-GPU aggregation: peer loads (k_fedavg_p2p) vs NVLS multimem, headline config and config 5, on $1 GPUs
N=${1:-2}
mkdir -p gpurun_out/agg
for mode in p2p nvls; do
  if [ $mode = nvls ]; then export FEDB200_AGG=nvls; else unset FEDB200_AGG; fi
  FEDB200_NO_FUSED_P2P=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2975$N bench.py --gpus $N --steps 1000 --warmup 20 2> gpurun_out/agg/${mode}_$N.err | grep "^{" > gpurun_out/agg/${mode}_$N.json
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2976$N bench.py --config wide --clients-per-gpu 32 --gpus $N --steps 6 --warmup 2 2> gpurun_out/agg/wide_${mode}_$N.err | grep "^{" > gpurun_out/agg/wide_${mode}_$N.json
done
python - <<'PY'
import json, glob, os
for f in sorted(glob.glob("gpurun_out/agg/*.json")):
    try:
        r = json.load(open(f))
        print(os.path.basename(f), "ms/step", round(r["ms_per_step"], 4), "warm", round(r["warm_no_flush"]["ms_per_step"], 4), r["kernels"][-2:], "acc", r["final_global_accuracy_train_shards"])
    except Exception as e:
        print(f, "FAILED", e)
PY
