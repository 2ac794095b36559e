import sys, ctypes, torch
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/perceiver-io-1_b200')
from perceiver import _native as N
L=N.lib(); DEV="cuda"
which = sys.argv[1] if len(sys.argv)>1 else "all"
def run(b,n,m,heads,reps=5,bwd=True,dh=64):
    e=heads*dh
    q=torch.randn(b,n,e,device=DEV).to(torch.bfloat16); kv=torch.randn(b,m,2*e,device=DEV).to(torch.bfloat16)
    do=torch.randn(b,n,e,device=DEV).to(torch.bfloat16)
    o=torch.empty(b,n,e,device=DEV); lse=torch.empty(b,heads,n,device=DEV)
    d=N.AttnDesc()
    d.q,d.k,d.v,d.pad,d.o,d.lse=q.data_ptr(),kv.data_ptr(),kv.data_ptr()+2*e,None,o.data_ptr(),lse.data_ptr()
    d.q_bs,d.q_rs,d.k_bs,d.k_rs,d.v_bs,d.v_rs,d.o_bs,d.o_rs=n*e,e,m*2*e,2*e,m*2*e,2*e,n*e,e
    d.batch,d.heads,d.n,d.m,d.dh,d.scale=b,heads,n,m,dh,dh**-0.5
    st=N.stream_ptr()
    def t(fn):
        ts=[]
        for i in range(reps+2):
            e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); e1.synchronize()
            if i>=2: ts.append(e0.elapsed_time(e1))
        return sum(ts)/len(ts)
    f=t(lambda: N.check(L.pio_attn_fwd(ctypes.byref(d),1,st)))
    fl=4.0*b*n*m*e
    msg=f"B{b} N{n} M{m} H{heads}: fwd {f*1e3:8.1f} us {fl/f/1e9:7.1f} TFLOP/s"
    if bwd:
        dq=torch.empty_like(q); dkv=torch.empty_like(kv); delta=torch.empty(b,heads,n,device=DEV)
        g=t(lambda: N.check(L.pio_attn_bwd_bf16(ctypes.byref(d),do.data_ptr(),dq.data_ptr(),dkv.data_ptr(),dkv.data_ptr()+2*e,delta.data_ptr(),st)))
        msg+=f" | bwd {g*1e3:8.1f} us {2.5*fl/g/1e9:7.1f} TFLOP/s (10NME convention)"
    print(msg)
if which=="fwdonly":
    run(32,256,16384,8,reps=1,bwd=False); sys.exit(0)
run(32,256,16384,8)
run(32,256,256,4,dh=128)
run(128,32,784,4,dh=32)
run(128,32,32,4,dh=32)
run(32,256,8192,8)
run(8,256,65536,8)
run(64,64,512,4)
run(64,512,64,4)
run(64,64,64,4)
