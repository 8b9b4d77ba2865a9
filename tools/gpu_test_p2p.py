"""Multi-GPU check of the fused peer-memory collectives (run under torchrun, one rank per GPU).

Correctness (every result against NCCL / a single-GPU restatement):
  F1 fedavg_            local delta reduce + cross-GPU reduce + update + broadcast
  allreduce_axpby_      fused all-reduce of a staged vector (also behind all_reduce_sum_)
  F3b weighted_reduce_  FoolsGold's weighted gradient reduce + server step
  F2 weiszfeld          fused per-iteration RFA loop vs the single-GPU device loop (flat.weiszfeld) on the union of rows
  gather_rows           evaluation exchange over peer pointers
Stress (``--stress N``, default 300): N back-to-back fused calls alternating fedavg / axpby / barrier with inputs that
change every call and random per-rank host delays, each result compared with NCCL - exercises the ticket barrier
protocol (the sanitizers are closed on this pool).
Timing: fused F1 / F2 vs their NCCL baselines, device-timed, max over ranks.  Prints 'P2P_OK' on rank 0.
"""
import os
import random
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dba_mod_b200.ops import flat  # noqa: E402
from dba_mod_b200.parallel.launcher import make_comm  # noqa: E402
from dba_mod_b200.parallel.symm import PeerComm  # noqa: E402

n_stress = 300
if "--stress" in sys.argv:
    n_stress = int(sys.argv[sys.argv.index("--stress") + 1])
comm = make_comm()
rank, world = comm.rank, comm.world
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
torch.cuda.set_device(dev)
D = int(os.environ.get("DBA_P2P_D", "2802410"))
pc = PeerComm(comm, D, dev, use_nvls={"1": True, "0": False}.get(os.environ.get("DBA_NVLS", "auto"), "auto"))
torch.manual_seed(1234)                       # identical streams on every rank
G0 = torch.randn(D, device=dev)
Kr = [3, 2, 1, 4, 2, 1, 3, 2][:world]
W_all = [G0[None] + 0.01 * torch.randn(k, D, device=dev) for k in Kr]
total = sum(Kr)
G = pc.set_global(G0.clone())
ok = True
log = []


def say(msg):
    if rank == 0:
        print(msg, flush=True)
    log.append(msg)


# ---------------------------------------------------------------------------------------------- F1
for it in range(3):
    ref = G.clone()
    acc = sum((w - ref[None]).sum(0) for w in W_all)
    ref = ref + 0.01 * acc
    pc.fedavg_(W_all[rank], 0.01)
    torch.cuda.synchronize()
    err = (G - ref).abs().max().item()
    ok &= err < 1e-5
    say(f"fedavg_ iter {it}: max err {err:.3e}")
# ------------------------------------------------------------------------------------- fused all-reduce
part = torch.full((D,), float(rank + 1), device=dev)
out = pc.allreduce_axpby_(part, 2.0, 0.0, target="aux")
torch.cuda.synchronize()
ok &= bool((out == 2.0 * sum(range(1, world + 1))).all())
t = torch.arange(10000, dtype=torch.float64, device=dev) * (rank + 1)
pc.all_reduce_sum_(t)
ok &= bool(torch.allclose(t, torch.arange(10000, dtype=torch.float64, device=dev) * sum(range(1, world + 1))))
say(f"allreduce_axpby_ / all_reduce_sum_: {'ok' if ok else 'FAIL'}")
# ---------------------------------------------------------------------------------------------- F3b
Dp = D - 10010
S_all = [torch.randn(k, Dp, device=dev) for k in Kr]
wv_all = [torch.rand(k, device=dev, dtype=torch.float64) for k in Kr]
G_before = G.clone()
ref = 0.95 * G_before[:Dp] - 0.01 * sum((w.float()[:, None] * s).sum(0) for w, s in zip(wv_all, S_all))
pc.weighted_reduce_(S_all[rank], wv_all[rank], Dp, -0.01, 0.95)
torch.cuda.synchronize()
err = (G[:Dp] - ref).abs().max().item()
ok &= err < 1e-4 and bool(torch.equal(G[Dp:], G_before[Dp:]))
say(f"weighted_reduce_ (FoolsGold reduce + server step): max err {err:.3e}")
# ---------------------------------------------------------------------------------------------- F2
G.copy_(G0)
torch.cuda.synchronize(); dist.barrier()
Wz = [w.clone() for w in W_all]
Wz[0][0] = G0 + 60 * (Wz[0][0] - G0)              # one scaled attacker
ns = [500.0 + 13 * i for i in range(total)]
tot = sum(ns)
slot0 = sum(Kr[:rank])
a_loc = torch.tensor(ns[slot0:slot0 + Kr[rank]], device=dev) / tot
med, table, state, obj = pc.weiszfeld(Wz[rank], G, a_loc, slot0, total, 10)
torch.cuda.synchronize()
pc.check()
Wcat = torch.cat(Wz, 0)
alphas = torch.tensor(ns, device=dev) / tot
m1, d1, w1, o1, s1 = flat.weiszfeld(Wcat, G0, alphas.float(), 10)
torch.cuda.synchronize()
e_med = (med - m1).abs().max().item() / m1.abs().max().item()
e_w = (table[0] - w1).abs().max().item()
e_d = ((table[1] - d1.float()).abs() / d1.float()).max().item()
same_iters = int(state[1]) == int(s1[1]) and int(state[2]) == int(s1[2])
ok &= e_med < 1e-4 and e_w < 1e-5 and e_d < 1e-4 and same_iters
say(f"weiszfeld fused vs single-GPU loop: median rel err {e_med:.2e}, wv err {e_w:.2e}, dist rel err {e_d:.2e}, "
    f"iters {int(state[1])}/{int(s1[1])}, calls {int(state[2])}/{int(s1[2])}, attacker weight {float(table[0, 0]):.2e}")
# --------------------------------------------------------------------------------------- gather_rows
rows = torch.full((Kr[rank], 1000), float(rank), device=dev) + torch.arange(Kr[rank], device=dev)[:, None] * 0.01
got = pc.gather_rows(rows, Kr)
exp = torch.cat([torch.full((k, 1000), float(r), device=dev) + torch.arange(k, device=dev)[:, None] * 0.01
                 for r, k in enumerate(Kr)])
torch.cuda.synchronize()
ok &= bool(torch.equal(got, exp))
say(f"gather_rows over peer pointers: {'ok' if torch.equal(got, exp) else 'FAIL'}")
# -------------------------------------------------------------------------------------------- stress
rng = random.Random(100 + rank)
G.copy_(G0)
Gn = G0.clone()                                   # NCCL-side replica of the same sequence
torch.cuda.synchronize(); dist.barrier()
bad = 0
for i in range(n_stress):
    if rng.random() < 0.3:
        time.sleep(rng.random() * 0.002)          # ranks drift apart on the host
    kind = i % 3
    if kind == 0:
        Wi = Gn[None] + (0.001 * (i + 1)) * torch.randn(Kr[rank], D, device=dev)
        pc.fedavg_(Wi, 0.01)
        acc = (Wi - Gn[None]).sum(0)
        dist.all_reduce(acc)
        Gn.add_(acc, alpha=0.01)
        if not torch.allclose(G, Gn, rtol=1e-5, atol=1e-5):
            bad += 1
    elif kind == 1:
        v = torch.randn(50000, device=dev) * (i + 1)
        exp = v.clone()
        dist.all_reduce(exp)
        got = pc.allreduce_axpby_(v, 1.0, 0.0, target="aux").clone()
        if not torch.allclose(got, exp, rtol=1e-5, atol=1e-4 * (i + 1)):
            bad += 1
    else:
        pc.device_barrier()
torch.cuda.synchronize()
pc.check()
ok &= bad == 0
say(f"stress: {n_stress} alternating fused calls with host jitter, {bad} mismatches vs NCCL")


# -------------------------------------------------------------------------------------------- timing
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
Gb = G0.clone()


def fused():
    pc.fedavg_(Wr, 0.0)


def baseline():
    acc = (Wr - Gb[None]).sum(0)
    dist.all_reduce(acc)
    Gb.add_(acc, alpha=0.0)


def wz_fused():
    pc.weiszfeld(Wz[rank], G, a_loc, slot0, total, 10)


tf, tb = timeit(fused), timeit(baseline)
tw = timeit(wz_fused, n=5)
nvl_bytes = 2.0 * D * 4 * (world - 1) / world          # two-shot all-reduce: bytes in + out per GPU over NVLink
hbm_bytes = (Kr[rank] + 1) * D * 4 + D * 4
roof_us = max(nvl_bytes / 770e9, hbm_bytes / 6553e9) * 1e6 if world > 1 else hbm_bytes / 6553e9 * 1e6
flag = torch.tensor([1.0 if ok else 0.0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
say(f"path: {'NVLS multimem' if pc.use_nvls else 'P2P loads/stores'} (multicast available: {pc.has_multicast})")
say(f"F1 fedavg fused: {tf * 1e3:.1f} us   nccl baseline (sum + all_reduce + add): {tb * 1e3:.1f} us   roofline "
    f"max(NVLink {nvl_bytes / 1e6:.1f} MB @770 GB/s, HBM {hbm_bytes / 1e6:.1f} MB @6553 GB/s) = {roof_us:.1f} us -> "
    f"{roof_us / (tf * 1e3):.2f} of roofline   world {world} K_local {Kr[rank]} D {D}")
say(f"F2 weiszfeld fused: {tw * 1e3:.0f} us for the whole loop (12 launches, {int(state[1])} iterations before "
    f"convergence)   world {world} n {total}")
say("P2P_OK" if flag.item() == 1.0 else "P2P_FAIL")
if rank == 0 and os.path.isdir("gpurun_out"):
    with open(f"gpurun_out/p2p_{world}gpu_{'nvls' if pc.use_nvls else 'p2p'}.txt", "w") as f:
        f.write("\n".join(log) + "\n")
dist.barrier()
dist.destroy_process_group()
