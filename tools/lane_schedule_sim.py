#!/usr/bin/env python
"""What could a different lane schedule buy k_raycast32?  (CPU only, numpy.)  Replays trace_ray (laser_models.py:107-146) for 400
raceline poses of the bench workload, keeps the number of marching steps of every beam, and counts warp iterations / issue slots
of: the lockstep rounds of 32 consecutive beams the kernel uses; rounds capped at T iterations whose unfinished rays are regrouped
densely (one or several levels); lanes refilled from the scan's beam queue every R iterations.  Reproduces the lane efficiency ncu
measures (0.64) and shows why neither scheme was built: regrouping long rays barely helps (0.64 -> 0.70 at best: the long rays are
heavy-tailed among themselves), and refilling costs more issue slots than it saves while the loop body is 10 instructions
(profiles/r2c_lane_schedule_sim.txt).      python tools/lane_schedule_sim.py"""
import os, sys, numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from red_gym_b200 import host      # map loading only (numpy / scipy on the host); no GPU needed, but the library must be built
mm = host.MapData(os.path.join(host.MAPS_DIR, 'example_map.yaml'), '.png', edt='scipy')
dt=mm.edt; res=mm.resolution; ox=mm.origin[0]; oy=mm.origin[1]; H,W=dt.shape
wp=np.loadtxt(os.path.join(host.MAPS_DIR, 'example_waypoints.csv'),delimiter=';',skiprows=3)
rng=np.random.default_rng(0)
n=400
idx=rng.integers(0,wp.shape[0],n)
px=wp[idx,1]+rng.uniform(-0.5,0.5,n); py=wp[idx,2]+rng.uniform(-0.5,0.5,n); yaw=wp[idx,3]+np.pi/2+rng.uniform(-0.5,0.5,n)
B=1080; fov=2*np.pi
ang=yaw[:,None]-fov/2+np.arange(B)[None,:]*(fov/(B-1))
cs=np.cos(ang); sn=np.sin(ang)
x=np.repeat(px[:,None],B,1); y=np.repeat(py[:,None],B,1)
def look(x,y):
    c=np.floor((x-ox)/res).astype(int); r=np.floor((y-oy)/res).astype(int)
    oob=(c<0)|(c>=W)|(r<0)|(r>=H)
    v=dt[np.clip(r,0,H-1),np.clip(c,0,W-1)]
    return np.where(oob,dt[-1,-1],v)
d=look(x,y); tot=d.copy(); steps=np.zeros((n,B),int)
alive=(d>1e-4)&(tot<=30)
while alive.any():
    x=np.where(alive,x+d*cs,x); y=np.where(alive,y+d*sn,y)
    dn=look(x,y); d=np.where(alive,dn,d); tot=np.where(alive,tot+dn,tot); steps+=alive
    alive=alive&(d>1e-4)&(tot<=30)
print('mean steps/beam',steps.mean(),'median',np.median(steps),'p95',np.percentile(steps,95),'max',steps.max(), 'lookups/scan', steps.sum(1).mean())
# baseline lockstep rounds of 32 consecutive beams
pad=(-B)%32
s=np.pad(steps,((0,0),(0,pad)))
r=s.reshape(n,-1,32)
it=r.max(2).sum(1)
print('baseline iterations/scan %.1f, efficiency %.3f'%(it.mean(), steps.sum(1).mean()/32/it.mean()))
# capped rounds with dense continuation (multi-level caps)
def capped(T_list):
    tot_it=0; tot_over=0.0
    for sc in steps:
        rem=sc.copy().astype(int)
        # level 0: rounds of consecutive beams, cap T0
        cur_groups=[np.arange(i,min(i+32,B)) for i in range(0,B,32)]
        its=0; batches=0
        for lvl,T in enumerate(T_list+[10**9]):
            nxt=[]
            for g in cur_groups:
                m_=rem[g].max()
                its+=min(m_,T)
                rem[g]=np.maximum(rem[g]-T,0)
                nxt.extend(g[rem[g]>0].tolist())
            if lvl>0: batches+=len(cur_groups)
            nxt=np.array(nxt,int)
            if nxt.size==0: break
            # dense regroup: sort by remaining? keep beam order
            cur_groups=[nxt[i:i+32] for i in range(0,nxt.size,32)]
        tot_it+=its; tot_over+=batches
    return tot_it/len(steps), tot_over/len(steps)
for T in ([6],[8],[10],[12],[6,12],[8,16],[8,16,32],[6,12,24,48],[4,8,16,32,64]):
    it_,b_=capped(T)
    print('caps',T,'iterations/scan %.1f (eff %.3f) continuation batches/scan %.1f'%(it_, steps.sum(1).mean()/32/it_, b_))

# refill schemes: 32 lanes, a queue of the scan's beams in order; every R iterations idle lanes take the next beam
def refill(R, dynamic=True):
    tot=0
    for sc in steps[:200]:
        q=list(sc); qi=0
        lane=np.zeros(32,int)
        it=0
        while True:
            if it % R == 0:
                for l in range(32):
                    if lane[l]==0 and qi<len(q):
                        lane[l]=q[qi]; qi+=1
                        # zero-step beams finish immediately; take another
                        while lane[l]==0 and qi<len(q):
                            lane[l]=q[qi]; qi+=1
            if qi>=len(q) and not lane.any(): break
            lane=np.maximum(lane-1,0); it+=1
        tot+=it
    return tot/200
base=steps[:200].sum(1).mean()/32
for R in (1,2,3,4,6,8):
    it_=refill(R)
    print('refill every %d iterations: iterations/scan %.1f efficiency %.3f ; issue slots/scan at 10/iter + 13 per refill point: %.0f (baseline %.0f)'%(R,it_,base/it_, it_*10 + (it_/R)*13, 352.8*10))
