# ours only: headline config at N = 1, 2, 4, 8 and BASELINE config 5 (wide) at N = 1 and 8
mkdir -p gpurun_out/scale2
for N in 1 2 4 8; do
  if [ $N -eq 1 ]; then
    python bench.py --gpus 1 --steps 2000 --warmup 20 > gpurun_out/scale2/ours_$N.json 2> gpurun_out/scale2/ours_$N.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2970$N bench.py --gpus $N --steps 2000 --warmup 20 2> gpurun_out/scale2/ours_$N.err | grep "^{" > gpurun_out/scale2/ours_$N.json
  fi
done
python bench.py --config wide --gpus 1 --steps 10 --warmup 3 > gpurun_out/scale2/wide_1.json 2> gpurun_out/scale2/wide_1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29718 bench.py --config wide --gpus 8 --steps 10 --warmup 3 2> gpurun_out/scale2/wide_8.err | grep "^{" > gpurun_out/scale2/wide_8.json
python - <<'PY'
import json
for name in ("ours_1","ours_2","ours_4","ours_8","wide_1","wide_8"):
    try:
        r=json.load(open(f"gpurun_out/scale2/{name}.json"))
        print(name, "ms/step", round(r["ms_per_step"],4), "M samples/s", round(r["value"]/1e6,2), "warm", round(r["warm_no_flush"]["ms_per_step"],4), "e2e", round(r["e2e"]["ms_per_step"],4), "acc", r["final_global_accuracy_train_shards"], r["clocks"]["reasons"])
    except Exception as e:
        print(name, "FAILED", e)
PY
tail -3 gpurun_out/scale2/wide_8.err
