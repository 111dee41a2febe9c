import sys, statistics, torch
sys.path.insert(0, '.')
import veles.simd_b200 as v
v.use_torch_stream()
def t(fn, reps=9):
    for _ in range(3): fn()
    torch.cuda.synchronize(); ts=[]
    for _ in range(reps):
        a,b=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return statistics.median(ts)
n=2**29; hl=8192
buf=torch.rand(n+hl+8,device='cuda'); h=torch.rand(hl,device='cuda')
y=torch.empty(n+hl+8,device='cuda')
hd=v.Handle(n,hl)
print('handle full conv  (pad=h-1)      ', t(lambda: hd(buf[:n],h,out=y[:n+hl-1])))
print('shard pad=0 aligned in/out       ', t(lambda: v.filter_shard(buf[:n+hl-1],h,0,0,n,out=y[:n])))
print('shard pad=0 x+1 (in misaligned)  ', t(lambda: v.filter_shard(buf[1:n+hl],h,0,0,n,out=y[:n])))
print('shard pad=0 out+8191 (odd offset)', t(lambda: v.filter_shard(buf[:n+hl-1],h,0,0,n-8191,out=y[8191:n])))
print('shard small job h-1 outputs      ', t(lambda: v.filter_shard(buf[:2*(hl-1)],h,0,0,hl-1,out=y[:hl-1])))
