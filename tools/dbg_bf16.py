import sys, torch
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/perceiver-io-1_b200'); sys.path.insert(0,'/root/repo/tests')
import perceiver
from perceiver import CrossAttention, SelfAttention, mlp, self_attention_block, PerceiverEncoder, TextInputAdapter
from oracle import perceiver_oracle as O
from gpu_util import rel_err
perceiver.set_compute_mode("bf16")
DEV="cuda"
torch.manual_seed(0)
def pert(m):
    g=torch.Generator().manual_seed(1)
    with torch.no_grad():
        for k,p in m.named_parameters():
            if k.endswith("bias") or "norm" in k: p.add_(torch.randn(p.shape,generator=g)*0.1)
for (B,Nn,C,H) in [(2,4,8,1),(2,8,8,2),(3,8,16,2),(4,8,16,2),(1,16,16,2),(2,8,16,1),(2,8,32,2),(4,40,16,2),(2,128,16,2),(2,8,24,2)]:
    x=torch.randn(B,Nn,C)
    ff=mlp(C); pert(ff)
    sd={"module."+k:v for k,v in ff.state_dict().items()}
    e_mlp=rel_err(ff.to(DEV)(x.to(DEV)), O.mlp_residual(sd,"",x)-x)
    sa=SelfAttention(C,H,0.0); pert(sa)
    sd={"module."+k:v for k,v in sa.state_dict().items()}
    e_sa=rel_err(sa.to(DEV)(x.to(DEV)), O.self_attention_residual(sd,"",x,H)-x)
    errs=[]
    for S in (1,2,3):
        blk=self_attention_block(S,C,H,0.0); pert(blk)
        sd=blk.state_dict()
        ref=O.self_attention_block(sd,"",x,H,S)
        errs.append(rel_err(blk.to(DEV)(x.to(DEV)), ref))
    print(f"B{B} N{Nn} C{C} H{H}: mlp {e_mlp:.2e} sa {e_sa:.2e} block S=1,2,3: "+" ".join(f"{e:.2e}" for e in errs))
