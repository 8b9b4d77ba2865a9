"""Does the +-1 row misalignment of the 3x3 tap views cost tensor-core operand bandwidth?  (profiling aid)"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dba_mod_b200.ops import conv, layout, native
lib = native.load()
d = torch.device("cuda", 0)
for (H, W, Cin, Cout) in [(32, 32, 32, 32), (16, 16, 64, 64), (8, 8, 128, 128), (4, 4, 256, 256)]:
    K, B = 10, 500
    x = layout.to_planar(torch.randn(1, B, Cin, H, W, device=d))
    wt = layout.pack_conv_weight(torch.randn(K, Cout, Cin, 3, 3, device=d) * 0.05)
    g = layout.Geom(B, H, W)
    y = torch.zeros(K, Cout // 8, g.rows, 8, dtype=torch.bfloat16, device=d)
    sc = torch.ones(K, Cout, device=d)
    for mode in (0, 1, 2):
        lib.dba_conv_set_debug_shift(mode)
        for _ in range(2):
            conv.conv_tc(x, wt, K, B, H, W, 3, 1, out=y, x_shared=True, scale=sc, shift=sc, relu=True)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            conv.conv_tc(x, wt, K, B, H, W, 3, 1, out=y, x_shared=True, scale=sc, shift=sc, relu=True)
        e1.record(); torch.cuda.synchronize()
        print(f"{(H, W, Cin, Cout)} shift mode {mode}: {e0.elapsed_time(e1) / 5 * 1e3:.1f} us", flush=True)
    lib.dba_conv_set_debug_shift(0)
