"""Multi-GPU parity check of the ray-sharded step (run under torchrun on G GPUs; not collected by pytest):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 tests/run_multigpu_check.py

Every rank takes its slice of ONE batch of rays.  Checked on every rank: (1) the sparse exchange (touched point rows,
rank-ordered merge) gives bit-identical gradients on all replicas, (2) it equals the dense NCCL all-reduce to fp32
rounding (bit-identical at G = 2, where a sum of two terms has one order), (3) both equal the single-process gradient of
the whole batch within 1e-4 of the gradient's max (the sums run in a different order), (4) two consecutive pipelined
steps (next batch's query prefetched on the side stream, row count piggybacked on the all-reduce) reproduce the same bits
as the unpipelined steps."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench  # noqa: E402  (workload builders only)
from pointnerf_b200.ray_shard import RayShardedStep, shard_bounds  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
    R = 4096
    opt, tables, batches, bg = bench.build_workload("cfg2", R, 2, seed_base=77)       # same batches on every rank
    model = bench.make_model(opt, tables, dev)
    bgp = torch.nn.Parameter(bg.to(dev))
    full = [bench.to_dev(b, dev) for b in batches]
    lo, hi = shard_bounds(R, world, rank)

    def mine(b):
        return dict(cam=b["cam"], raydir=b["raydir"][:, lo:hi].contiguous(), gt=b["gt"][:, lo:hi].contiguous(),
                    uniform=b["uniform"][:, lo:hi].contiguous())
    shards = [mine(b) for b in full]
    res = {}
    for name, kw in (("sparse", dict(sparse=True)), ("dense", dict(sparse=False))):
        st = RayShardedStep(model, bgp, **kw)
        b = shards[0]
        st(b["cam"], b["raydir"], b["gt"], uniform=b["uniform"], global_rays=R)
        res[name] = st.flat_grads().clone()
    # single-process gradient of the whole batch (no collective: reduce=False)
    st1 = RayShardedStep(model, bgp, sparse=False)
    b = full[0]
    st1(b["cam"], b["raydir"], b["gt"], uniform=b["uniform"], global_rays=R, reduce=False)
    whole = st1.flat_grads().clone()
    scale = float(whole.abs().max())
    e_sd = float((res["sparse"] - res["dense"]).abs().max()) / scale
    e_sw = float((res["sparse"] - whole).abs().max()) / scale
    gathered = [torch.empty_like(res["sparse"]) for _ in range(world)]
    dist.all_gather(gathered, res["sparse"])
    same = all(torch.equal(g, gathered[0]) for g in gathered)
    # pipelined: step 0 with next_batch = 1, then step 1; against unpipelined step 1
    stp = RayShardedStep(model, bgp, sparse=True)
    stp(shards[0]["cam"], shards[0]["raydir"], shards[0]["gt"], uniform=shards[0]["uniform"], global_rays=R, next_batch=shards[1])
    stp(shards[1]["cam"], shards[1]["raydir"], shards[1]["gt"], uniform=shards[1]["uniform"], global_rays=R)
    piped = stp.flat_grads().clone()
    stu = RayShardedStep(model, bgp, sparse=True, side_prefetch=False)
    stu(shards[1]["cam"], shards[1]["raydir"], shards[1]["gt"], uniform=shards[1]["uniform"], global_rays=R)
    pipe_ok = torch.equal(piped, stu.flat_grads())
    ok = same and e_sd <= (0.0 if world == 2 else 1e-6) and e_sw <= 1e-4 and pipe_ok and stp.caps_piggybacked == 1
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("world %d: replicas identical %s | sparse vs dense %.2e | sparse vs single-process %.2e | pipelined == unpipelined %s "
              "(capacity piggybacked %d) -> %s" % (world, same, e_sd, e_sw, pipe_ok, stp.caps_piggybacked, "OK" if int(flag) else "FAIL"))
    dist.destroy_process_group()
    return 0 if int(flag) else 1


if __name__ == "__main__":
    sys.exit(main())
