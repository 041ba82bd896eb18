import importlib, sys, numpy as np, torch
sys.path.insert(0,'/root/repo')
nv = importlib.import_module("nano-vectordb_b200")
dev=torch.device('cuda',0)
N,D,Q=100000,384,200
g=torch.Generator(device=dev); g.manual_seed(1)
x=torch.nn.functional.normalize(torch.randn(N,D,device=dev,generator=g),dim=1).half()
idx=nv.FlatIndexCuda.from_device(x.data_ptr(),None,nv.F16,N,D,keepalive=x)
ivf=nv.IvfPqIndex(idx,nlist=1024,pq_m=64,ntrain=50000,niter=5,seed=1)
q=torch.nn.functional.normalize(x[:Q].float()+0.05*torch.randn(Q,D,device=dev,generator=g),dim=1).cpu().numpy()
for R in (50,100,128,200):
    b,_=ivf.search(q,8,R)
    s=np.stack([ivf.search(q[i:i+1],8,R)[0][0] for i in range(Q)])
    print('ivf search R',R,'batch==single',np.array_equal(b,s))
    rb,db=idx.refine_l2(q,b,10)
    rs=np.stack([idx.refine_l2(q[i:i+1],b[i:i+1],10)[0][0] for i in range(Q)])
    print('  refine batch==single',np.array_equal(rb,rs), 'bad rows', np.nonzero((rb!=rs).any(1))[0][:5])
    sb,_=ivf.search_refine(q,8,R,10)
    print('  search_refine == search+refine', np.array_equal(sb,rb))
    # oracle check of refine: brute force
    xb=x.float()
    for i in range(3):
        c=b[i]; c=c[c<N]
        d=((torch.from_numpy(q[i]).to(dev)[None,:]-xb[c.astype(np.int64)])**2).sum(1)
        o=c[torch.argsort(d,stable=True)[:10].cpu().numpy()]
        if not np.array_equal(o,rb[i]): print('   refine vs torch mismatch q',i,o,rb[i])
