"""torchrun script: cost of the pieces of the touched-row exchange at the bench weak-scaling sizes (pack, all-gather, merge, MLP all-reduce):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/exchange_pieces.py"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch, torch.distributed as dist
from pointnerf_b200 import _lib
rank, world, lrank = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lrank)
dist.init_process_group("nccl", device_id=torch.device("cuda", lrank))
dev = torch.device("cuda", lrank)
T, F, cap = 1_000_000, 32, 54528
W = F + 8
lib = _lib.lib()
g = torch.Generator(device="cpu").manual_seed(rank)
ids = torch.randperm(T, generator=g)[:cap - 500].to(torch.int32).to(dev)
emb = torch.randn(T, F, device=dev); conf = torch.randn(T, device=dev); dirs = torch.randn(T, 3, device=dev); col = torch.randn(T, 3, device=dev)
packed = torch.empty(cap, W, device=dev); gathered = torch.empty(world * cap, W, device=dev)
slots = torch.full((T * world,), -1, dtype=torch.int32, device=dev)
flat = torch.randn(400_000, device=dev)
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
def pack(): _lib.check(lib.pn_rows_pack(_lib.ptr(ids), ids.numel(), cap, F, _lib.ptr(emb), _lib.ptr(conf), _lib.ptr(dirs), _lib.ptr(col), _lib.ptr(packed), _lib.stream()), "pack")
def gather(): dist.all_gather_into_tensor(gathered, packed)
def merge(): _lib.check(lib.pn_rows_merge(_lib.ptr(gathered), world, cap, F, T, _lib.ptr(slots), _lib.ptr(emb), _lib.ptr(conf), _lib.ptr(dirs), _lib.ptr(col), _lib.stream()), "merge")
def allred(): dist.all_reduce(flat)
r = dict(pack=t(pack), all_gather=t(gather), merge=t(merge), all_reduce_flat=t(allred))
if rank == 0:
    print("world", world, "cap", cap, "row bytes", W * 4, "MB per rank %.1f" % (cap * W * 4 / 1e6), {k: round(v, 1) for k, v in r.items()}, "us", flush=True)
torch.cuda.synchronize(); dist.barrier(); os._exit(0)
