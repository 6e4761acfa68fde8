import sys,time; sys.path.insert(0,'.')
import numpy as np, torch
import awesomeslam_b200 as A
for (mp,msd,B) in ((A.map_cfg1(),30,16384),(A.map_cfg1(),20,16384),(A.map_cfg3(),36,16384)):
    lmx,lmy=mp; L=len(lmx)
    rp=A.Replay(lmx,lmy,B,24+8*8,seed0=0x0CF20000)
    dev=torch.device("cuda",0)
    dt=[torch.from_numpy(a).to(dev) for a in (rp.x,rp.y,rp.yaw,rp.vx,rp.az,rp.dt,rp.range,rp.bearing)]
    f=A.BatchedFilter("ukf",B,0,max_state_dim=msd)
    def run(m0,n): 
        p=[t.data_ptr()+m0*t[0].numel()*t.element_size() for t in dt]; f.run_device(n,*p,L)
    run(0,24); f.sync()
    run(24,8); f.sync()
    t0=time.perf_counter()
    for k in range(6): run(32+8*k,8)
    f.sync(); t=time.perf_counter()-t0
    print("n=%d cap %d: %.2f M updates/s"%(3+2*L, msd, B*48/t/1e6))
    f.close()
