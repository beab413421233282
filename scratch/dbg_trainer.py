import sys, copy, torch, torch.nn.functional as F
sys.path.insert(0, '/root/repo')
import distributed_compute_pytorch_b200 as dcp
from distributed_compute_pytorch_b200 import ops
from distributed_compute_pytorch_b200.parallel.convnet_trainer import FusedConvNetTrainer
torch.manual_seed(0)
N=128
net = dcp.models.ConvNet(); net.dropout1.p=0.0; net.dropout2.p=0.0
net = net.cuda().train()
opt = ops.Adadelta(net.parameters(), lr=0.0)
tr = FusedConvNetTrainer(net, opt, N, use_graph=False)
x = torch.randn(N,1,28,28, device='cuda'); y = torch.randint(0,10,(N,), device='cuda')
loss = tr.step(x, y); torch.cuda.synchronize()
def rel(a,b): return ((a.double()-b.double()).norm()/(b.double().norm()+1e-30)).item()
torch.backends.cudnn.allow_tf32=False; torch.backends.cuda.matmul.allow_tf32=False
w1,b1,w2,b2 = net.conv1.weight, net.conv1.bias, net.conv2.weight, net.conv2.bias
a1 = F.relu(F.conv2d(x, w1, b1))                      # [N,32,26,26]
print('a1', rel(tr.a1.view(N,26,26,32).permute(0,3,1,2), a1))
a2 = F.relu(F.conv2d(a1, w2, b2))                     # [N,64,24,24]
a2p = torch.zeros(N,64,26,26, device='cuda'); a2p[:,:,:24,:24]=a2
print('a2', rel(tr.a2.view(N,26,26,64).permute(0,3,1,2), a2p))
t = tr.a2.view(N,26,26,64).permute(0,3,1,2)
print('a2 valid-region', rel(t[:,:,:24,:24], a2), 'invalid abs max', t[:,:,24:,:].abs().max().item(), t[:,:,:,24:].abs().max().item())
pooled = F.max_pool2d(a2,2).flatten(1)
print('pooled', rel(tr.pooled, pooled))
h1 = F.linear(pooled, net.fc1.weight, net.fc1.bias)
print('h1', rel(tr.h1, h1))
mean, var = h1.mean(0), h1.var(0, unbiased=False)
h2 = F.relu((h1-mean)/torch.sqrt(var+1e-5)*net.batchnorm.weight+net.batchnorm.bias)
print('h2', rel(tr.h2, h2))
logits = F.linear(h2, net.fc2.weight, net.fc2.bias)
print('logits', rel(tr.logits, logits))
print('loss', loss.item(), F.cross_entropy(logits, y).item())
# backward reference via autograd on torch ops
params = [p.detach().clone().requires_grad_(True) for p in (w1,b1,w2,b2,net.fc1.weight,net.fc1.bias,net.fc2.weight,net.fc2.bias,net.batchnorm.weight,net.batchnorm.bias)]
W1,B1,W2,B2,FW1,FB1,FW2,FB2,G,Bt = params
A1 = F.relu(F.conv2d(x,W1,B1)); A1.retain_grad()
A2 = F.relu(F.conv2d(A1,W2,B2)); A2.retain_grad()
PL = F.max_pool2d(A2,2).flatten(1); PL.retain_grad()
H1 = F.linear(PL,FW1,FB1); H1.retain_grad()
H2 = F.relu(F.batch_norm(H1,None,None,G,Bt,True)); H2.retain_grad()
LG = F.linear(H2,FW2,FB2); LG.retain_grad()
L = F.cross_entropy(LG,y); L.backward()
print('dlogits', rel(tr.dlogits, LG.grad)); print('dh2', rel(tr.dh2, H2.grad)); print('dh1', rel(tr.dh1, H1.grad)); print('dpooled', rel(tr.dpooled, PL.grad))
dy2 = torch.zeros(N,64,26,26,device='cuda'); 
# grad wrt pre-relu conv2 output = A2.grad * (A2>0)
dy2[:,:,:24,:24] = A2.grad*(A2>0)
print('dy2', rel(tr.dy2.view(N,26,26,64).permute(0,3,1,2), dy2))
dz1 = A1.grad*(A1>0)
print('dz1', rel(tr.dz1.view(N,26,26,32).permute(0,3,1,2), dz1))
names=['conv1.weight','conv1.bias','conv2.weight','conv2.bias','fc1.weight','fc1.bias','fc2.weight','fc2.bias','batchnorm.weight','batchnorm.bias']
pd = dict(net.named_parameters())
for n,pr in zip(names,params): print(n, rel(pd[n].grad, pr.grad))
