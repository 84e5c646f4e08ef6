import sys, torch
sys.path.insert(0,'/root/repo/self-attention-experiments-vision_b200')
from sav import ops
torch.manual_seed(0)
dev='cuda'
C=192
for (N,K,epi) in ((576,192,'plain'),(768,192,'gelu2'),(192,768,'f32res'),(192,192,'f32res'),(768,192,'dgelu')):
    M=1576
    A=torch.randn(M,K,device=dev).bfloat16(); W=(torch.randn(K,N,device=dev)/K**0.5).bfloat16()
    kw1={}; kw2={}
    if epi=='gelu2':
        b=torch.randn(N,device=dev)
        p1=torch.empty(M,N,device=dev,dtype=torch.bfloat16); p2=torch.empty(M//2,N,device=dev,dtype=torch.bfloat16)
        kw1=dict(bias=b,act=1,aux_out=p1); kw2=dict(bias=b,act=1,aux_out=p2)
    elif epi=='f32res':
        b=torch.randn(N,device=dev); r=torch.randn(M,N,device=dev)
        kw1=dict(bias=b,residual=r,out_fp32=True); kw2=dict(bias=b,residual=r[:M//2],out_fp32=True)
    elif epi=='dgelu':
        h=torch.randn(M,N,device=dev).bfloat16()
        kw1=dict(aux_in=h); kw2=dict(aux_in=h[:M//2])
    full=ops.gemm(A,W,b_mn=True,**kw1)
    half=ops.gemm(A[:M//2],W,b_mn=True,**kw2)
    torch.cuda.synchronize()
    d=(full[:M//2].float()-half.float()).abs().max().item()
    print(epi,N,K,'max abs diff full-vs-half', d, 'equal', torch.equal(full[:M//2],half))
# attention fwd/bwd batch invariance
B,N,H,D=8,197,3,64
qkv=torch.randn(B,N,3,H,D,device=dev).bfloat16()
q,k,v=qkv[:,:,0],qkv[:,:,1],qkv[:,:,2]
o2,lse=ops.attn_fwd(q,k,v)
qh=qkv[:4].contiguous(); o2h,lseh=ops.attn_fwd(qh[:,:,0],qh[:,:,1],qh[:,:,2])
print('attn fwd equal', torch.equal(o2[0][:4],o2h[0]), torch.equal(lse[:4],lseh))
do=torch.randn(B,N,H,D,device=dev).bfloat16()
dqkv=torch.empty_like(qkv); ops.attn_bwd(q,k,v,o2,do,lse,dqkv[:,:,0],dqkv[:,:,1],dqkv[:,:,2])
dqh=torch.empty_like(qh); ops.attn_bwd(qh[:,:,0],qh[:,:,1],qh[:,:,2],o2h,do[:4].contiguous(),lseh,dqh[:,:,0],dqh[:,:,1],dqh[:,:,2])
torch.cuda.synchronize()
print('attn bwd equal', torch.equal(dqkv[:4],dqh), (dqkv[:4].float()-dqh.float()).abs().max().item())
# LN
x=torch.randn(1576,C,device=dev); sc=torch.randn(C,device=dev); bi=torch.randn(C,device=dev)
y,m,r=ops.layernorm_fwd(x,sc,bi); yh,mh,rh=ops.layernorm_fwd(x[:788].contiguous(),sc,bi)
print('ln fwd equal', torch.equal(y[:788],yh))
dy=torch.randn(1576,C,device=dev).bfloat16(); g=torch.randn(1576,C,device=dev)
ds=torch.zeros(C,device=dev); db=torch.zeros(C,device=dev)
dx,dx16=ops.layernorm_bwd(dy,x,sc,m,r,ds,db,dx_add=g)
dxh,dx16h=ops.layernorm_bwd(dy[:788].contiguous(),x[:788].contiguous(),sc,mh,rh,ds,db,dx_add=g[:788].contiguous())
torch.cuda.synchronize()
print('ln bwd equal', torch.equal(dx[:788],dxh), torch.equal(dx16[:788],dx16h))
