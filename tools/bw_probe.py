import torch
def t(fn, n=5):
    fn(); torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(True),torch.cuda.Event(True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)/n
out=torch.empty(32*50*480*640,dtype=torch.float64,device='cuda')
src=torch.empty(out.numel()//2,dtype=torch.float32,device='cuda')
ms=t(lambda: out.zero_()); print("write-only 4.9 GB f64 zero_: %.3f ms %.0f GB/s"%(ms,out.numel()*8/ms/1e6))
ms=t(lambda: out.fill_(1.5)); print("fill_: %.3f ms %.0f GB/s"%(ms,out.numel()*8/ms/1e6))
a=torch.empty(out.numel(),dtype=torch.float32,device='cuda')
ms=t(lambda: out.copy_(a)); print("f32->f64 copy (read 1.97 GB, write 3.9 GB): %.3f ms %.0f GB/s"%(ms,out.numel()*12/ms/1e6))
b=torch.empty_like(out)
ms=t(lambda: b.copy_(out)); print("f64 copy: %.3f ms %.0f GB/s"%(ms,out.numel()*16/ms/1e6))
