"""Worker of tests/test_multigpu.py, launched under torchrun: every rank hosts its share of the clients of ONE custom
federation in a DistFedEngine, runs a few rounds and rank 0 writes the global model, every client's slab as this rank sees
it, and the metric history to an .npz file."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    out, n_clients, rows, rounds = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
    precision = sys.argv[5] if len(sys.argv) > 5 else None
    from fedb200 import FedProblem, prepare
    from fedb200.parallel.dist_engine import DistFedEngine, init_distributed
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    init_distributed()
    data = prepare(with_mean=True)
    shards = [np.arange(c * rows, (c + 1) * rows, dtype=np.int32) for c in range(n_clients)]
    prob = FedProblem.custom(data.n_features, data.n_classes, shards, patience=0)
    eng = DistFedEngine([prob], data.X_train, data.y_train, data.X_test, data.y_test, rank=rank, world=world, max_rounds=rounds + 4,
                        precision=precision)
    h = eng.run(rounds)
    torch.cuda.synchronize()
    glob = [torch.empty_like(eng.global_) for _ in range(world)]
    torch.distributed.all_gather(glob, eng.global_)
    if rank == 0:
        np.savez(out, global_=eng.global_.cpu().numpy(), globals_all=torch.stack(glob).cpu().numpy(), hist=h["global"], local=h["local"],
                 kernels=np.asarray(eng.kernel_names), launches=eng.launches_per_round, precision=eng.precision_name)
    torch.distributed.barrier()
    eng.close()
    torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
