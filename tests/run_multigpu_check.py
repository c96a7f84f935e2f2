"""Multi-GPU parity check of the ray-sharded training step (run under torchrun on G GPUs; not collected by pytest):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 tests/run_multigpu_check.py [cfg2]

Every rank takes its slice of ONE cfg2 batch of rays and runs the CUDA-graph step (FusedTrainStep) with the shipped colour
loss (ray_masked / ray_miss / plain items: the masked-ray normaliser is a count summed over ranks).  Checked on every rank
(pointnerf_b200.train_step.sharded_gradient_check): the exchanged gradients are identical bits on all replicas; the
touched-row exchange equals the dense NCCL all-reduce (exactly at G = 2); both equal the single-process gradient of the
whole batch within 1e-4 of the gradient's max.  bench.py --gpus N prints the same check in its JSON line."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench  # noqa: E402  (workload builders only)
from pointnerf_b200.train_step import sharded_gradient_check  # noqa: E402


def main():
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
    name = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
    R = 4096
    opt, tables, batches, bg = bench.build_workload(name, R, 1, seed_base=77)       # the same batch on every rank
    model = bench.make_model(opt, tables, dev)
    bgp = torch.nn.Parameter(bg.to(dev))
    full = bench.to_dev(batches[0], dev)
    ok = True
    for items, wts in ((("ray_masked_coarse_raycolor", "ray_miss_coarse_raycolor", "coarse_raycolor"), (1.0, 0.0, 0.0)),
                       (("ray_masked_coarse_raycolor", "ray_miss_coarse_raycolor", "coarse_raycolor"), (1.0, 0.5, 0.25)),
                       (("coarse_raycolor",), (1.0,))):
        res = sharded_gradient_check(model, bgp, full, items, wts)
        res["loss_weights"] = list(wts)
        ok = ok and res["ok"]
        if rank == 0:
            print(json.dumps(res))
    torch.cuda.synchronize()
    dist.barrier()
    sys.stdout.flush()
    os._exit(0 if ok else 1)       # (no destroy_process_group after collectives captured in CUDA graphs: see bench.py)


if __name__ == "__main__":
    sys.exit(main())
