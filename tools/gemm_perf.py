import sys, ctypes, torch
ONLY = sys.argv[1] if len(sys.argv)>1 else None
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/perceiver-io-1_b200')
from perceiver import _native as N
L=N.lib(); DEV="cuda"
def run(m,n,k,a_k,b_k,out32=True,reps=20):
    p8=lambda x:(x+7)//8*8
    a=torch.randn((m,p8(k)) if a_k else (k,p8(m)),device=DEV).to(torch.bfloat16)
    b=torch.randn((n,p8(k)) if b_k else (k,p8(n)),device=DEV).to(torch.bfloat16)
    d=N.GemmBf16Desc(); d.a,d.b=a.data_ptr(),b.data_ptr(); d.m,d.n,d.k=m,n,k; d.lda,d.ldb=a.shape[1],b.shape[1]
    d.a_kmajor,d.b_kmajor=int(a_k),int(b_k)
    c32=torch.empty(m,n,device=DEV) if out32 else None
    c16=None if out32 else torch.empty(m,n,device=DEV,dtype=torch.bfloat16)
    d.c32,d.ldc32=N.ptr(c32),n; d.c16,d.ldc16=N.ptr(c16),n
    ws=torch.empty(L.pio_gemm_bf16_ws_bytes(m,n,k),dtype=torch.uint8,device=DEV)
    flush=torch.empty(256<<20,dtype=torch.uint8,device=DEV)
    st=N.stream_ptr(); ts=[]
    for i in range(reps+3):
        flush.zero_()
        e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        e0.record(); N.check(L.pio_gemm_bf16(ctypes.byref(d),ws.data_ptr(),ws.numel(),st)); e1.record(); e1.synchronize()
        if i>=3: ts.append(e0.elapsed_time(e1))
    ms=sum(ts)/len(ts); fl=2.0*m*n*k
    byts=(m*k+n*k)*2+m*n*(4 if out32 else 2)
    print(f"m{m} n{n} k{k} {'K' if a_k else 'M'}{'K' if b_k else 'M'} out{'32' if out32 else '16'}: {ms*1e3:8.1f} us  {fl/ms/1e9:7.1f} TFLOP/s  {byts/ms/1e6:7.1f} GB/s")
if ONLY:
    run(32768,10003,256,True,True,reps=2); sys.exit(0)
run(32768,10003,256,True,True)
run(32768,10003,256,True,True,out32=False)
run(32768,256,10003,True,False)
run(10003,256,32768,False,False)
run(32768,512,256,True,True)
run(32768,512,256,True,True,out32=False)
run(4096,256,256,True,True)
run(4096,768,256,True,True)
run(256,256,4096,False,False)
run(768,256,4096,False,False)
run(8192,8192,8192,True,True,out32=False,reps=5)
run(8192,8192,8192,True,False,out32=False,reps=5)
