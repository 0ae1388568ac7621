import torch, sys
from singa_b200.ops import kernels as K
dev=torch.device("cuda:0")
def t(fn,n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)/n
for (M,N,Kd) in [(8192,8192,8192),(4096,16384,16384),(16384,4096,16384)]:
    a=torch.randn(M,Kd,device=dev).bfloat16(); b=torch.randn(N,Kd,device=dev).bfloat16()
    o=torch.empty(M,N,device=dev,dtype=torch.bfloat16); o32=torch.empty(M,N,device=dev)
    bias=torch.randn(N,device=dev)
    for rep in range(2):
        ms=t(lambda: K.gemm(a,b,M,N,Kd,out16=o,bias=bias,act=1))
        ms32=t(lambda: K.gemm(a,b,M,N,Kd,out32=o32))
        print(f"{M}x{N}x{Kd}: bf16+bias+relu {ms:.3f} ms {2*M*N*Kd/ms/1e9:.0f} TF | fp32 out {ms32:.3f} ms {2*M*N*Kd/ms32/1e9:.0f} TF")
