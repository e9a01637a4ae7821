import torch, numpy as np
from torchbooster_b200.transforms import BatchTransform, TransformSpec, CropFlipParams
IM=((0.485, 0.456, 0.406), (0.229, 0.224, 0.225))
B=512
bt=BatchTransform(TransformSpec(crop=224,flip_p=.5,mean=IM[0],std=IM[1],dtype=torch.bfloat16),device="cuda")
xs=[torch.randint(0,256,(B,256,256,3),dtype=torch.uint8,device="cuda") for _ in range(4)]
t,l,f=CropFlipParams(42).draw(B,256,256,(224,224),(0,0),.5)
prm=tuple(torch.from_numpy(a).cuda() for a in (t,l,f))
out=torch.empty((B,3,224,224),dtype=torch.bfloat16,device="cuda")
for i in range(6): bt(xs[i%4],params=prm,out=out)
torch.cuda.synchronize()
for rep in range(4):
    for i in range(3): bt(xs[i],params=prm,out=out)      # back to back; the last one's stamps survive
    torch.cuda.synchronize()
    t=out.view(-1)[:32].view(torch.int64).cpu().numpy()
    print("init %d  gridwait %d  params+issue %d  table %d  first tile ready +%d  second +%d  rest %d  total %d ns" % (t[1]-t[0],t[2]-t[1],t[3]-t[2],t[4]-t[3],t[5]-t[4],t[6]-t[5],t[7]-t[6],t[7]-t[0]))
