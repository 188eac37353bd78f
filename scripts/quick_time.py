import torch, numpy as np, sys, time
sys.path.insert(0,'.')
from detectron2_centernet_b200 import target as T, losses as L, decode as D, synthetic as S
dev='cuda:0'
B,C,H,W,K=32,80,128,128,100
bl,cl=S.make_batch_objects(1236,B,512)
boxes,classes,offsets=S.pack_objects(bl,cl)
tb,tc,to=[torch.from_numpy(x).to(dev) for x in (boxes,classes,offsets)]
g=torch.Generator(device=dev).manual_seed(1)
hm=(torch.randn((B,C,H,W),device=dev,generator=g)*1.5-2.19).requires_grad_(True)
wh=(torch.rand((B,2,H,W),device=dev,generator=g)*64).requires_grad_(True)
reg=torch.rand((B,2,H,W),device=dev,generator=g).requires_grad_(True)
flush=torch.empty(512*1024*1024//4,device=dev)
def timeit(fn,n=int(sys.argv[1]) if len(sys.argv)>1 else 20):
    for _ in range(3): fn()
    ts=[]
    for _ in range(n):
        flush.zero_()
        s=torch.cuda.Event(enable_timing=True); e=torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize(); ts.append(s.elapsed_time(e))
    return np.median(ts)*1e3, np.min(ts)*1e3
gt=T.target_gen(tb,tc,to,B,C,H,W)
el=B*C*H*W
t=timeit(lambda: T.target_gen(tb,tc,to,B,C,H,W,out=gt)); print('target_gen us',t,'GB/s',el*4/t[0]/1e3)
def loss():
    o=L.ctdet_losses({'hm':hm,'wh':wh,'reg':reg},gt,alpha=[1.0]); return o
t=timeit(loss); print('loss fwd(+grad) us',t,'GB/s',el*12/t[0]/1e3)
def fused():
    return L.ctdet_losses_from_objects({'hm':hm,'wh':wh,'reg':reg},tb,tc,to,alpha=[1.0])
t=timeit(fused); print('fused target+loss fwd(+grad) us',t,'GB/s (8 B/cell)',el*8/t[0]/1e3)
def lossb():
    o=L.ctdet_losses({'hm':hm,'wh':wh,'reg':reg},gt,alpha=[1.0]); (o['hm_loss']+o['wh_loss']+o['off_loss']).backward(); hm.grad=None;wh.grad=None;reg.grad=None
t=timeit(lossb); print('loss fwd+autograd bwd us',t)
hd=hm.detach();whd=wh.detach();rd=reg.detach()
t=timeit(lambda: D.ctdet_decode_batch(hd,whd,rd,4,False,K,logits=True)); print('decode(logits) B=32 us',t,'GB/s',el*4/t[0]/1e3)
B2=64
hm2=torch.randn((B2,C,H,W),device=dev,generator=g)*1.5-2.19
heat2=torch.clamp(torch.sigmoid(hm2),1e-4,1-1e-4)
wh2=torch.rand((B2,2,H,W),device=dev,generator=g)*64; reg2=torch.rand((B2,2,H,W),device=dev,generator=g)
t=timeit(lambda: D.ctdet_decode_batch(heat2,wh2,reg2,4,False,K)); print('decode(probs) B=64 us',t,'GB/s',B2*C*H*W*4/t[0]/1e3)
