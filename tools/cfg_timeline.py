"""Kernel-time breakdown of one eager train step of BASELINE config 1 or 4 (torch.profiler / CUPTI)."""
import sys, collections, re, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/perceiver-io-1_b200')
import perceiver
from perceiver import (ClassificationOutputAdapter, ImageInputAdapter, PerceiverDecoder, PerceiverEncoder, PerceiverIO, TextInputAdapter)
from torch.profiler import profile, ProfilerActivity
which = sys.argv[1] if len(sys.argv) > 1 else "cfg4"
perceiver.set_compute_mode("bf16")
dev = torch.device("cuda"); torch.manual_seed(0)
if which == "cfg1":
    enc = PerceiverEncoder(ImageInputAdapter((28, 28, 1), 32), (32, 128), 3, 4, 4, 3)
    dec = PerceiverDecoder(ClassificationOutputAdapter(10, num_output_channels=128), (32, 128), 1)
    args = (torch.rand(128, 28, 28, 1, device=dev) * 2 - 1,); y = torch.randint(0, 10, (128,), device=dev)
else:
    enc = PerceiverEncoder(TextInputAdapter(10003, 16384, 512), (256, 512), 3, 8, 4, 6)
    dec = PerceiverDecoder(ClassificationOutputAdapter(2, num_output_channels=512), (256, 512), 1)
    args = (torch.randint(3, 10003, (32, 16384), device=dev),); y = torch.randint(0, 2, (32,), device=dev)
model = PerceiverIO(enc, dec).to(dev)
opt = torch.optim.AdamW(model.parameters(), lr=1e-3, fused=True)
def step():
    loss = perceiver.cross_entropy(model(*args), y)
    opt.zero_grad(set_to_none=True); loss.backward(); opt.step()
for _ in range(2): step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    step(); torch.cuda.synchronize()
agg = collections.defaultdict(lambda: [0, 0.0]); tot = 0.0
for ev in prof.events():
    if ev.device_type == torch.autograd.DeviceType.CUDA:
        n = re.sub(r'\(.*', '', ev.name).replace('void ', '')
        agg[n][0] += 1; agg[n][1] += ev.device_time; tot += ev.device_time
print(f"{which}: sum of kernel time {tot/1e3:.3f} ms/step")
for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:22]:
    print(f"{t:9.1f} us {100*t/tot:5.1f}% x{c:4d} avg {t/c:8.1f} us  {n[:90]}")
