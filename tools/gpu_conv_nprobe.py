"""How does the per-MMA cost grow with N?  Times 3x3 convolutions Cin -> Cout for several Cout (profiling aid)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dba_mod_b200.ops import conv, layout, native
lib = native.load()
d = torch.device("cuda", 0)
K, B = 10, 500
for (H, Cin, couts) in [(32, 32, (32, 64, 96, 128, 192)), (16, 64, (64, 128, 192, 256))]:
    x = layout.to_planar(torch.randn(1, B, Cin, H, H, device=d))
    for Cout in couts:
        wt = layout.pack_conv_weight(torch.randn(K, Cout, Cin, 3, 3, device=d) * 0.05)
        g = layout.Geom(B, H, H)
        y = torch.zeros(K, Cout // 8, g.rows, 8, dtype=torch.bfloat16, device=d)
        sc = torch.ones(K, Cout, device=d)
        for _ in range(2):
            conv.conv_tc(x, wt, K, B, H, H, 3, 1, out=y, x_shared=True, scale=sc, shift=sc, relu=True)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            conv.conv_tc(x, wt, K, B, H, H, 3, 1, out=y, x_shared=True, scale=sc, shift=sc, relu=True)
        e1.record(); torch.cuda.synchronize()
        print(f"H={H} Cin={Cin} Cout={Cout}: {e0.elapsed_time(e1) / 5 * 1e3:.1f} us", flush=True)
