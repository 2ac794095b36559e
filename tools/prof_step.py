"""Kernel timeline of the bench step under torch.profiler (CUPTI): real, overlapped execution."""
import sys, collections, re, torch
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/perceiver-io-1_b200')
import bench, perceiver
from torch.profiler import profile, ProfilerActivity
mode = sys.argv[1] if len(sys.argv) > 1 else "bf16"
perceiver.set_compute_mode(mode)
dev = torch.device("cuda")
torch.manual_seed(0)
model = bench.build_model().to(dev)
opt = torch.optim.AdamW(model.parameters(), lr=3e-3, fused=True)
ids, pad = bench.synthetic_batch(0, 64, 512, 10003)
ids, pad = ids.to(dev), pad.to(dev)
def step():
    if mode == "bf16":
        loss, _ = model(ids.clone(), pad, return_loss=True)
    else:
        logits, labels = model(ids.clone(), pad)
        loss = perceiver.cross_entropy(logits, labels)
    opt.zero_grad(set_to_none=True); loss.backward(); opt.step()
for _ in range(3): step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3): step()
    torch.cuda.synchronize()
agg = collections.defaultdict(lambda: [0, 0.0])
tot = 0.0
for ev in prof.events():
    if ev.device_type == torch.autograd.DeviceType.CUDA:
        n = re.sub(r'\(.*', '', ev.name); n = re.sub(r'^void ', '', n)
        agg[n][0] += 1; agg[n][1] += ev.device_time if hasattr(ev, "device_time") else ev.cuda_time
        tot += ev.device_time if hasattr(ev, "device_time") else ev.cuda_time
print(f"3 steps: sum of kernel time {tot/3e3:.3f} ms/step")
for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:24]:
    print(f"{t/3:9.1f} us/step {100*t/tot:5.1f}% x{c//3:4d} avg {t/c:7.1f} us  {n[:90]}")
