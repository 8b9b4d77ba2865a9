"""GPU probe: correctness of conv_tc for both descriptor conventions + timing per layer shape.  Writes
gpurun_out/probe_conv.json.  Usage: timeout 300 python tools/gpu_probe_conv.py"""
import json
import os
import sys
import time

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dba_mod_b200.ops import conv, layout, native  # noqa: E402

torch.backends.cudnn.allow_tf32 = False
d = torch.device("cuda", 0)
out = {"correctness": [], "timing": []}
os.makedirs("gpurun_out", exist_ok=True)


def check(K, B, H, W, Cin, Cout, ks, sub, swap):
    torch.manual_seed(0)
    x = torch.randn(K, B, Cin, H, W, device=d)
    w = torch.randn(K, Cout, Cin, ks, ks, device=d) / (Cin * ks * ks) ** 0.5
    y = conv.conv_tc(layout.to_planar(x), layout.pack_conv_weight(w), K, B, H, W, ks, sub, debug_swap=swap)
    torch.cuda.synchronize()
    got = layout.from_planar(y, B, H // sub, W // sub)
    xb, wb = x.bfloat16().float(), w.bfloat16().float()
    ref = torch.stack([F.conv2d(xb[k], wb[k], None, sub, ks // 2) for k in range(K)])
    err = (got - ref).abs().max().item()
    return err, ref.abs().max().item()


for shp in [(1, 1, 8, 8, 32, 32, 1, 1), (1, 1, 8, 8, 32, 32, 3, 1), (2, 3, 8, 8, 64, 64, 3, 1),
            (1, 2, 32, 32, 32, 64, 3, 2), (2, 5, 4, 4, 256, 256, 3, 1), (1, 2, 2, 2, 512, 512, 3, 1), (2, 64, 16, 16, 64, 128, 1, 2)]:
    for swap in (0,):
        try:
            err, mag = check(*shp, swap)
            rec = {"shape": shp, "swap": swap, "max_err": err, "ref_max": mag}
        except Exception as e:  # noqa: BLE001
            rec = {"shape": shp, "swap": swap, "error": str(e)[:200]}
        print(rec, flush=True)
        out["correctness"].append(rec)
    json.dump(out, open("gpurun_out/probe_conv.json", "w"), indent=1)

good = [r for r in out["correctness"] if r.get("max_err", 1e9) < 0.05]
swap = good[0]["swap"] if good else 0
out["chosen_swap"] = swap

# timing: CIFAR layer shapes at K=10 clients x B=64 (training step) and eval (10 models x 1000 images, shared input)
layers = [(32, 32, 32, 32, 3, 1), (32, 32, 32, 64, 3, 2), (16, 16, 64, 64, 3, 1), (16, 16, 64, 128, 3, 2),
          (8, 8, 128, 128, 3, 1), (8, 8, 128, 256, 3, 2), (4, 4, 256, 256, 3, 1), (32, 32, 32, 64, 1, 2)]
for (K, B, shared) in [(10, 64, False), (10, 1000, True), (1, 64, False)]:
    for (H, W, Cin, Cout, ks, sub) in layers:
        x = layout.to_planar(torch.randn(1 if shared else K, B, Cin, H, W, device=d))
        wt = layout.pack_conv_weight(torch.randn(K, Cout, Cin, ks, ks, device=d) * 0.05)
        g = layout.Geom(B, H // sub, W // sub)
        y = torch.zeros(K, Cout // 8, g.rows, 8, dtype=torch.bfloat16, device=d)
        for _ in range(3):
            conv.conv_tc(x, wt, K, B, H, W, ks, sub, out=y, x_shared=shared, debug_swap=swap)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 10
        e0.record()
        for _ in range(n):
            conv.conv_tc(x, wt, K, B, H, W, ks, sub, out=y, x_shared=shared, debug_swap=swap)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        flops = 2.0 * K * B * H * W * Cin * Cout * ks * ks   # full-resolution MACs actually executed
        useful = flops / (sub * sub)
        # cuDNN reference timing (grouped conv, bf16, channels_last)
        xr = torch.randn(B, K * Cin, H, W, device=d, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
        wr = torch.randn(K * Cout, Cin, ks, ks, device=d, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
        try:
            for _ in range(3):
                F.conv2d(xr, wr, None, sub, ks // 2, groups=K)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(n):
                F.conv2d(xr, wr, None, sub, ks // 2, groups=K)
            e1.record()
            torch.cuda.synchronize()
            ms_ref = e0.elapsed_time(e1) / n
        except Exception as e:  # noqa: BLE001
            ms_ref = None
        rec = {"K": K, "B": B, "shared": shared, "layer": (H, W, Cin, Cout, ks, sub), "ms": ms,
               "tflops_executed": flops / ms / 1e9, "tflops_useful": useful / ms / 1e9, "cudnn_grouped_ms": ms_ref}
        print(rec, flush=True)
        out["timing"].append(rec)
        json.dump(out, open("gpurun_out/probe_conv.json", "w"), indent=1)
print("done")
