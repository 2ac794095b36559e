import sys, torch
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/perceiver-io-1_b200')
import perceiver
from perceiver import functional as F, _native as N
from perceiver import PerceiverDecoder, PerceiverEncoder, PerceiverMLM, TextInputAdapter, TextMasking, TextOutputAdapter
perceiver.set_compute_mode("bf16")
torch.manual_seed(0)
enc = PerceiverEncoder(TextInputAdapter(503, 64, 64), (16, 64), 2, 4, 4, 2)
dec = PerceiverDecoder(TextOutputAdapter(503, 64, num_output_channels=64), (16, 64), 4)
model = PerceiverMLM(enc, dec, TextMasking(503, 1, 2, 3)).cuda()
opt = torch.optim.AdamW(model.parameters(), lr=1e-2, fused=True)
ids = torch.randint(3, 503, (8, 64), device="cuda"); pad = torch.zeros(8, 64, dtype=torch.bool, device="cuda")
names = {id(p): n for n, p in model.named_parameters()}
for it in range(3):
    c0 = N.lib().pio_launch_count()
    loss, _ = model(ids.clone(), pad, return_loss=True)
    print("step", it, "launches", N.lib().pio_launch_count() - c0, "epoch", F._SH, "table", model in F._TABLES and len(F._TABLES[model][2]))
    for p in model.parameters():
        h = getattr(p, "_pio_w16", None)
        if h is None: continue
        ok = torch.equal(h[2][:, :p.shape[1]], p.detach().to(torch.bfloat16))
        if not ok:
            d = (h[2][:, :p.shape[1]].float() - p.detach()).abs().max().item()
            print("  MISMATCH", names[id(p)], tuple(p.shape), tuple(h[2].shape), "ver", p._version, h[0], "epoch", h[3], "maxdiff", d)
    opt.zero_grad(set_to_none=True); loss.backward(); opt.step()
