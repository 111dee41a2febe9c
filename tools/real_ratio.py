import os, sys, statistics, torch
sys.path.insert(0, '.')
import veles.simd_b200 as v
v.use_torch_stream()
def t(fn, reps=7):
    for _ in range(2): fn()
    torch.cuda.synchronize(); ts=[]
    for _ in range(reps):
        a,b=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b)*1e3)
    return statistics.median(ts)
for n in (1024, 2048, 4096, 8192, 16384):
    hl = int(n*0.4); xl=262144; B=256
    x=torch.rand((B,xl),device='cuda'); h=torch.rand((hl,),device='cuda'); y=torch.empty((B,xl+hl-1),device='cuda')
    res=[]
    for real in ('0','1'):
        os.environ['VELES_B200_REAL']=real; os.environ['VELES_B200_FORCE_N']=str(n)
        hd=v.Handle(xl,hl,'_overlap_save')
        inner=v.ConvolutionHandle(); inner.algorithm, inner.x_length, inner.h_length = 2, xl, hl; inner.handle.os=hd.handle
        fn=lambda: v.lib.convolve_batch(inner,x.data_ptr(),xl,h.data_ptr(),y.data_ptr(),xl+hl-1,B)
        us=t(fn); hd.close()
        step = (n-hl+1) if real=='0' else ((2*n-(hl-1+((hl-1)&1)))&~1)
        nseg=-(-(xl+hl-1)//step); units = (-(-nseg//2)) if real=='0' else nseg
        res.append((us, us/units))
    print(f"N={n} h={hl}: paired {res[0][0]:.0f} us, real {res[1][0]:.0f} us, per-transform cost ratio real/paired = {res[1][1]/res[0][1]:.2f}")
