"""Error of the tensor-core attention forward against fp64 at a long-key shape (B2 N256 M16384 H2 dh64)."""
import sys, ctypes, torch
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/perceiver-io-1_b200')
from perceiver import _native as N
L=N.lib(); DEV="cuda"
torch.manual_seed(0)
for scale_in in (1.0, 3.0):
    b,n,m,heads,dh=2,256,16384,2,64; e=heads*dh
    q=(torch.randn(b,n,e)*scale_in).to(DEV).to(torch.bfloat16); kv=torch.randn(b,m,2*e).to(DEV).to(torch.bfloat16)
    o=torch.empty(b,n,e,device=DEV); lse=torch.empty(b,heads,n,device=DEV)
    d=N.AttnDesc()
    d.q,d.k,d.v,d.pad,d.o,d.lse=q.data_ptr(),kv.data_ptr(),kv.data_ptr()+2*e,None,o.data_ptr(),lse.data_ptr()
    d.q_bs,d.q_rs,d.k_bs,d.k_rs,d.v_bs,d.v_rs,d.o_bs,d.o_rs=n*e,e,m*2*e,2*e,m*2*e,2*e,n*e,e
    d.batch,d.heads,d.n,d.m,d.dh,d.scale=b,heads,n,m,dh,dh**-0.5
    N.check(L.pio_attn_fwd(ctypes.byref(d),1,N.stream_ptr()))
    torch.cuda.synchronize()
    qd=q.double().view(b,n,heads,dh).transpose(1,2); kd=kv[...,:e].double().view(b,m,heads,dh).transpose(1,2); vd=kv[...,e:].double().view(b,m,heads,dh).transpose(1,2)
    s=(qd@kd.transpose(-1,-2))*dh**-0.5
    ref=(torch.softmax(s,-1)@vd).transpose(1,2).reshape(b,n,e)
    lref=torch.logsumexp(s,-1)
    err=(o.double()-ref).abs().max()/ref.abs().max()
    print(f"q scale {scale_in}: o max-abs err / max |o| = {err.item():.2e}; lse max abs err = {(lse.double()-lref).abs().max().item():.2e}")
