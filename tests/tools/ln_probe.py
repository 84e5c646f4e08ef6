import sys, torch
sys.path.insert(0,'/root/repo/self-attention-experiments-vision_b200')
from sav import ops
torch.manual_seed(0)
dev='cuda'
for C in (192, 384, 768):
    x=torch.randn(1576,C,device=dev); sc=torch.randn(C,device=dev); bi=torch.randn(C,device=dev)
    y,m,r=ops.layernorm_fwd(x,sc,bi); yh,mh,rh=ops.layernorm_fwd(x[:788].contiguous(),sc,bi)
    print(C,'stats equal', torch.equal(m[:788],mh), torch.equal(r[:788],rh))
    dy=torch.randn(1576,C,device=dev).bfloat16(); g=torch.randn(1576,C,device=dev)
    ds=torch.zeros(C,device=dev); db=torch.zeros(C,device=dev)
    dx,dx16=ops.layernorm_bwd(dy,x,sc,m,r,ds,db,dx_add=g)
    dx2,_=ops.layernorm_bwd(dy,x,sc,m,r,ds,db,dx_add=g)
    dxh,dx16h=ops.layernorm_bwd(dy[:788].contiguous(),x[:788].contiguous(),sc,mh,rh,ds,db,dx_add=g[:788].contiguous())
    torch.cuda.synchronize()
    d=(dx[:788]-dxh).abs()
    print(C,'repeat equal', torch.equal(dx,dx2),'full-vs-half max', d.max().item(), 'rows differing', int((d.amax(1)>0).sum()), 'first rows', (d.amax(1)>0).nonzero().flatten()[:8].tolist(), 'cols', (d.amax(0)>0).nonzero().flatten()[:8].tolist())
    # no dx_add
    dxa,_=ops.layernorm_bwd(dy,x,sc,m,r,ds,db); dxb,_=ops.layernorm_bwd(dy[:788].contiguous(),x[:788].contiguous(),sc,mh,rh,ds,db)
    torch.cuda.synchronize()
    print(C,'no dx_add: equal', torch.equal(dxa[:788],dxb))
