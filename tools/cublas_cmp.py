import torch,sys
for (M,N,K) in [(25616,2304,768),(25616,3072,768),(25616,768,3072),(25616,768,768)]:
    A=torch.randn(M,K,device='cuda').half(); B=torch.randn(N,K,device='cuda').half()
    for _ in range(3): C=A@B.t()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): C=A@B.t()
    e1.record(); torch.cuda.synchronize()
    ms=e0.elapsed_time(e1)/20
    print(f"cublas {M},{N},{K}: {ms:.3f} ms {2*M*N*K/ms/1e9:.1f} TF/s")
