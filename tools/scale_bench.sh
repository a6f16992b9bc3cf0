mkdir -p gpurun_out/scale
for N in 1 2 4 8; do
  if [ $N -eq 1 ]; then
    python bench.py --gpus 1 --steps 1000 --warmup 20 > gpurun_out/scale/ours_$N.json 2> gpurun_out/scale/ours_$N.err
    python bench.py --impl nccl --gpus 1 --steps 300 --warmup 10 > gpurun_out/scale/nccl_$N.json 2> gpurun_out/scale/nccl_$N.err
    python bench.py --impl reference --gpus 1 --steps 200 --warmup 5 > gpurun_out/scale/ref_$N.json 2> gpurun_out/scale/ref_$N.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2960$N bench.py --gpus $N --steps 1000 --warmup 20 2> gpurun_out/scale/ours_$N.err | grep "^{" > gpurun_out/scale/ours_$N.json
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2961$N bench.py --impl nccl --gpus $N --steps 300 --warmup 10 2> gpurun_out/scale/nccl_$N.err | grep "^{" > gpurun_out/scale/nccl_$N.json
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2962$N bench.py --impl reference --gpus $N --steps 200 --warmup 5 2> gpurun_out/scale/ref_$N.err | grep "^{" > gpurun_out/scale/ref_$N.json
  fi
done
python - <<'PY'
import json,glob
for impl in ("ours","nccl","ref"):
    for N in (1,2,4,8):
        try:
            r=json.load(open(f"gpurun_out/scale/{impl}_{N}.json"))
            print(impl, N, round(r["ms_per_step"]*1e3,1), "us/round", round(r["value"]/1e6,2), "M samples/s", "acc", r.get("final_global_accuracy_train_shards"), (r.get("warm_no_flush") or {}).get("ms_per_step"), (r.get("e2e") or {}).get("ms_per_step"))
        except Exception as e:
            print(impl, N, "FAILED", e)
PY
