"""Per-role cycle breakdown of conv_tc for the CIFAR layer shapes (profiling aid, uses dba_conv_set_debug)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dba_mod_b200.ops import conv, layout, native
lib = native.load()
d = torch.device("cuda", 0)
dbg = torch.zeros(148 * 12, dtype=torch.int64, device=d)
layers = [(32, 32, 32, 32, 3, 1), (32, 32, 32, 64, 3, 2), (16, 16, 64, 64, 3, 1), (8, 8, 128, 128, 3, 1),
          (4, 4, 256, 256, 3, 1), (32, 32, 32, 64, 1, 2)]
if len(sys.argv) > 1 and sys.argv[1] == "nprobe":      # per-MMA cost vs N
    layers = [(32, 32, 32, 32, 3, 1), (32, 32, 32, 64, 3, 1), (32, 32, 32, 96, 3, 1), (32, 32, 32, 128, 3, 1),
              (16, 16, 64, 64, 3, 1), (16, 16, 64, 192, 3, 1)]
cases = [(10, 64, False, True), (10, 500, True, False)]
if len(sys.argv) > 1 and sys.argv[1] == "k1":
    cases = [(1, 64, False, True), (2, 64, False, True)]
for (K, B, shared, stats) in cases:
    for (H, W, Cin, Cout, ks, sub) in layers:
        x = layout.to_planar(torch.randn(1 if shared else K, B, Cin, H, W, device=d))
        wt = layout.pack_conv_weight(torch.randn(K, Cout, Cin, ks, ks, device=d) * 0.05)
        g = layout.Geom(B, H // sub, W // sub)
        y = torch.zeros(K, Cout // 8, g.rows, 8, dtype=torch.bfloat16, device=d)
        st = torch.zeros(K, Cout, 2, device=d) if stats else None
        sc = None if stats else torch.ones(K, Cout, device=d)
        for _ in range(2):
            conv.conv_tc(x, wt, K, B, H, W, ks, sub, out=y, x_shared=shared, stats=st, scale=sc, shift=sc, relu=not stats)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        conv.conv_tc(x, wt, K, B, H, W, ks, sub, out=y, x_shared=shared, stats=st, scale=sc, shift=sc, relu=not stats)
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3
        dbg.zero_()
        lib.dba_conv_set_debug(native.ptr(dbg))
        conv.conv_tc(x, wt, K, B, H, W, ks, sub, out=y, x_shared=shared, stats=st, scale=sc, shift=sc, relu=not stats)
        torch.cuda.synchronize()
        lib.dba_conv_set_debug(None)
        v = dbg.reshape(148, 3, 4).double()
        used = (v.sum((1, 2)) > 0).sum().item()
        m = v[:used].mean(0) / 1e3
        print(f"K={K} B={B} {'stats' if stats else 'eval '} layer {(H, W, Cin, Cout, ks, sub)}: {us:7.1f} us  ctas {used} | kcyc/CTA  "
              f"prod wait_a_empty {m[0,0]:.1f} | mma wait_acc_empty {m[1,0]:.1f} wait_a_full {m[1,1]:.1f} issue {m[1,2]:.1f} | "
              f"epi wait_acc_full {m[2,0]:.1f} work {m[2,1]:.1f} | lifetime prod {m[0,3]:.1f} mma {m[1,3]:.1f} epi {m[2,3]:.1f} "
              f"max {v[:used, :, 3].max().item() / 1e3:.1f}", flush=True)
