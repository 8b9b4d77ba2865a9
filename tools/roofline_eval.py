"""Per-layer roofline of one evaluation chunk (CIFAR ResNet-18, M models x Bt images) from an ncu launch list.

    python tools/roofline_eval.py gpurun_out/launches_eval2.csv [M] [Bt] [out.md]

Useful FLOPs = 2 * MACs of the real convolution (no halo rows, no padding); bytes = activations read + residual read +
activations written + weights (bf16).  Peaks: MEASURED_PEAKS.json (driver-written copy bandwidth and cuBLAS bf16)."""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
path = sys.argv[1]
M = int(sys.argv[2]) if len(sys.argv) > 2 else 10
Bt = int(sys.argv[3]) if len(sys.argv) > 3 else 500
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
PB = peaks["hbm_gbs"]

lines = [ln for ln in open(path) if ln.startswith('"')]
rows = [r for r in csv.DictReader(lines) if r["Metric Name"] == "gpu__time_duration.sum"]
names = [(r["Kernel Name"].split("(")[0], float(r["Metric Value"].replace(",", "")) / 1e3) for r in rows]
starts = [i for i, n in enumerate(names) if "input_stage" in n[0]]
# the first evaluate() call's chunks are warm-up; take the last shared-input chunk of the second call
chunk = None
for a, b in zip(starts, starts[1:]):
    seg = names[a:b]
    if sum(1 for n in seg if "conv_tc" in n[0]) == 20 and chunk is None or (chunk is not None and False):
        chunk = seg
# prefer a later (warm) chunk
for a, b in zip(starts, starts[1:]):
    seg = names[a:b]
    if sum(1 for n in seg if "conv_tc" in n[0]) == 20:
        chunk = seg
        if a >= starts[len(starts) // 4]:
            break
convs = [n for n in chunk if "conv_tc" in n[0]]
# compute mode from the kernel's template arguments: conv_tc_kernel<KC2, TF32>
TF32 = any(", 1>" in n[0] or "true>" in n[0] for n in convs)
ELEM = 4.0 if TF32 else 2.0                      # bytes per activation element
PF = peaks["bf16_tflops"] / (2.0 if TF32 else 1.0)   # dense tf32 rate = half the measured bf16 rate
MODE = "tf32 (fp32 activations, kind::tf32)" if TF32 else "bf16"
layers = [("stem (im2col 27->32)", 32, 32, 32, 1, 1, False)]
cin, res = 32, 32
for w in (32, 64, 128, 256):
    for bi in range(2):
        sub = 2 if (w > 32 and bi == 0) else 1
        ro = res // sub
        layers.append((f"{w}@{ro} c1 3x3" + (" s2 (space-to-depth)" if sub == 2 else ""), cin, w, ro, 9, sub, False))
        if sub == 2:
            layers.append((f"{w}@{ro} shortcut 1x1 s2", cin, w, ro, 1, sub, False))
        layers.append((f"{w}@{ro} c2 3x3 +res", w, w, ro, 9, 1, True))
        cin, res = w, ro
assert len(layers) == len(convs), (len(layers), len(convs))
n_img = M * Bt
out = [f"Evaluation chunk, {MODE}: {M} models x {Bt} images = {n_img} image-forwards; peaks {PF:.0f} TFLOP/s "
       f"({'half of ' if TF32 else ''}the measured cuBLAS bf16 rate), {PB:.0f} GB/s (copy, measured)", "",
       "| layer | us | useful TFLOP/s | % of peak | GB/s moved | % of copy peak | bound |", "|---|---|---|---|---|---|---|"]
tot_us = tot_fl = 0.0
for (name, ci, co, ro, taps, sub, has_res), (_, us) in zip(layers, convs):
    real_ci = 27 if name.startswith("stem") else ci
    macs = n_img * ro * ro * taps * real_ci * co
    fl = 2.0 * macs
    in_elems = n_img * ro * ro * ci * (4 if (sub == 2 and taps == 9) else 1)     # s2d input: four parity planes
    if name.startswith("stem"):
        in_elems = Bt * ro * ro * 32                                              # shared input
    byts = ELEM * (in_elems + n_img * ro * ro * co * (2 if has_res else 1))
    tf, gb = fl / us / 1e6, byts / us / 1e3
    bound = "HBM" if gb / PB > tf / PF else "tensor"
    out.append(f"| {name} | {us:.0f} | {tf:.0f} | {100 * tf / PF:.1f} | {gb:.0f} | {100 * gb / PB:.1f} | {bound} |")
    tot_us += us
    tot_fl += fl
oth = sum(us for n, us in chunk if "conv_tc" not in n)
out += ["", f"convolutions {tot_us:.0f} us, other kernels {oth:.0f} us; whole chunk {tot_fl / (tot_us + oth) / 1e6:.0f} "
        f"useful TFLOP/s = {100 * tot_fl / (tot_us + oth) / 1e6 / PF:.1f} % of that peak; "
        f"{(tot_us + oth) / n_img:.3f} us per image-forward"]
txt = "\n".join(out) + "\n"
if len(sys.argv) > 4:
    open(sys.argv[4], "w").write(txt)
print(txt)
