import sys; sys.path.insert(0,".")
import torch, ctypes as C
from rnnwavefunctions_b200 import _lib
L=_lib.lib
L.rnnwf_debug_gemm_tc.argtypes=[C.c_void_p,C.c_longlong,C.c_int,C.c_void_p,C.c_longlong,C.c_int,C.c_void_p,C.c_longlong,C.c_longlong,C.c_int,C.c_longlong,C.c_int,C.c_void_p]
L.rnnwf_debug_gemm_tc.restype=C.c_int
dev="cuda"
torch.manual_seed(0)
def run(Aop, a_kc, Bop, b_kc, M,N,K, mode, Cinit=None):
    Cm = torch.zeros(M,N,device=dev) if Cinit is None else Cinit.clone()
    rc=L.rnnwf_debug_gemm_tc(Aop.data_ptr(), Aop.stride(0), a_kc, Bop.data_ptr(), Bop.stride(0), b_kc, Cm.data_ptr(), Cm.stride(0), M,N,K,mode,None)
    torch.cuda.synchronize()
    assert rc==0, L.rnnwf_last_error()
    return Cm
def tf32(x): return (x.view(torch.int32) & ~0x1FFF).view(torch.float32)
for (M,N,K) in [(128,128,32),(300,100,104),(1000,800,100),(100,800,5000),(200,100,3332),(37,5,77*4)]:
    A=torch.randn(M,K,device=dev); B=torch.randn(N,K,device=dev)
    ref=(A.double()@B.double().T)
    # nt: A K-contig, B K-contig
    c=run(A,1,B,1,M,N,K,0); e1=((c.double()-ref).abs().max()/ref.abs().max()).item()
    # nn: B stored [K,N] N-contig
    Bt=B.T.contiguous(); c=run(A,1,Bt,0,M,N,K,0); e2=((c.double()-ref).abs().max()/ref.abs().max()).item()
    # tn: A stored [K,M] M-contig, B stored [K,N]; atomic split
    At=A.T.contiguous(); c=run(At,0,Bt,0,M,N,K,2); e3=((c.double()-ref).abs().max()/ref.abs().max()).item()
    c=run(At,0,Bt,0,M,N,K,1,Cinit=torch.ones(M,N,device=dev)); e4=((c.double()-ref-1).abs().max()/ref.abs().max()).item() if True else 0
    print((M,N,K), "nt %.2e nn %.2e tn(atomic) %.2e tn(+=) %.2e"%(e1,e2,e3,e4))
# timing of the backward-sized shapes
import time
for (M,N,K,a_kc,b_kc,mode,name) in [(10000,800,100,1,0,0,"dT=ds*W2 (nn)"),(100,800,10000,0,0,2,"dW2+=ds^T T (tn)"),(10000,200,100,1,1,0,"dhc=dm Wm^T (nt)")]:
    if a_kc: A=torch.randn(M,K,device=dev)
    else: A=torch.randn(K,M,device=dev)
    B=torch.randn(N,K,device=dev) if b_kc else torch.randn(K,N,device=dev)
    Cm=torch.zeros(M,N,device=dev)
    for _ in range(3): L.rnnwf_debug_gemm_tc(A.data_ptr(),A.stride(0),a_kc,B.data_ptr(),B.stride(0),b_kc,Cm.data_ptr(),Cm.stride(0),M,N,K,mode,None)
    torch.cuda.synchronize(); e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True); e0.record()
    for _ in range(20): L.rnnwf_debug_gemm_tc(A.data_ptr(),A.stride(0),a_kc,B.data_ptr(),B.stride(0),b_kc,Cm.data_ptr(),Cm.stride(0),M,N,K,mode,None)
    e1.record(); torch.cuda.synchronize(); ms=e0.elapsed_time(e1)/20
    print(name, "%.1f us  %.1f TFLOP/s"%(ms*1e3, 2*M*N*K/ms/1e9))
