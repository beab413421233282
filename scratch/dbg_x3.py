import sys, torch
sys.path.insert(0, '/root/repo')
from distributed_compute_pytorch_b200.ops import cuda, tc
cuda.lib()
def rel(a,b): return ((a.double()-b.double()).abs().max()/(b.double().abs().max()+1e-12)).item()
torch.manual_seed(1)
for (M,N,K,ks) in [(128,128,256,1),(128,128,9216,36),(200,72,384,1)]:
  for a_mn,b_mn in [(0,0),(0,1),(1,1),(1,0)]:
    A=torch.randn(M,K,device='cuda'); B=torch.randn(N,K,device='cuda')
    ref=A.double()@B.double().t()
    Ast=(A.t().contiguous() if a_mn else A).contiguous(); Bst=(B.t().contiguous() if b_mn else B).contiguous()
    o3=torch.zeros(M,N,device='cuda'); tc.gemm(Ast,Bst,o3,None,M,N,K,Ast.shape[1],Bst.shape[1],N,a_mn,b_mn,False,ks,x3=True)
    o1=torch.zeros(M,N,device='cuda'); tc.gemm(Ast,Bst,o1,None,M,N,K,Ast.shape[1],Bst.shape[1],N,a_mn,b_mn,False,ks,x3=False)
    torch.cuda.synchronize()
    print(M,N,K,ks,a_mn,b_mn,'x3',rel(o3,ref),'x1',rel(o1,ref),'fp32',rel(A@B.t(),ref))
