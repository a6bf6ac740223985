import sys, os, time, ctypes as C
sys.path.insert(0,'/root/repo/risk-aware-hybrid-lqr-mpc-navigation-for-autonomous-systems_b200')
import numpy as np, torch
from hybrid_controller import _native as nat, workloads as wl
B,N,M=65536,20,5
dev=torch.device('cuda',0)
x0,xr,ur,ob=wl.config4_instances(B,N=N,seed=100)
host=[torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in (x0,xr.reshape(B,-1),ur.reshape(B,-1),ob.reshape(B,-1))]
big=torch.cat([h.reshape(-1) for h in host]).pin_memory(); d=torch.empty_like(big,device=dev)
torch.cuda.synchronize()
for _ in range(3):
    t0=time.perf_counter(); d.copy_(big,non_blocking=True); torch.cuda.synchronize(); t=time.perf_counter()-t0
print("raw H2D %.1f MB: %.3f ms = %.1f GB/s"%(big.numel()*8/1e6,t*1e3,big.numel()*8/t/1e9))
h_u0=torch.empty((B,2),dtype=torch.float64).pin_memory(); h_cost=torch.empty(B,dtype=torch.float64).pin_memory()
h_st,h_it,h_sl=(torch.empty(B,dtype=torch.int32).pin_memory() for _ in range(3))
prm=nat.mpc_params(nat.MPC_LTV,N,M,1,[15,15,50],[.1,.1],[30,30,40],.3,5000.,2.,3.,.02)
for ch in (1,2,3,4,6,8):
    os.environ["B2C_HOST_CHUNKS"]=str(ch)
    ctx=nat.Context(0)
    def step():
        ctx.call("b2c_mpc_host",B,C.byref(prm),host[0].data_ptr(),host[1].data_ptr(),host[2].data_ptr(),host[3].data_ptr(),None,h_u0.data_ptr(),h_cost.data_ptr(),h_st.data_ptr(),h_it.data_ptr(),h_sl.data_ptr(),None,None)
    for _ in range(3): step()
    t0=time.perf_counter()
    for _ in range(20): step()
    t=(time.perf_counter()-t0)/20
    print("chunks",ch,"ms %.3f"%(t*1e3),"Msolves/s %.1f"%(B/t/1e6))
    ctx.close()
