import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/visual-slam_b200')
import numpy as np, time
from orbslam2 import utils
K_G=np.array([796.7,0,239,0,787.8,425,0,0,1.]).reshape(3,3); K_S=np.array([[320.,0,320],[0,320,240],[0,0,1]])
def jac(a,b): a=a.ravel()!=0; b=b.ravel()!=0; return (a&b).sum()/max((a|b).sum(),1)
for name,tag,K in (('synth640','s640',K_S),('pair01','pair01',K_G),('pair12','pair12',K_G)):
    g=np.load('tests/golden/%s.npz'%name); p1,p2=g[tag+'_p1'],g[tag+'_p2']
    for seed in (0x5EED,1,2,3):
        t0=time.time(); F,m=utils.find_fundamental_ransac(p1,p2,3.0,0.99,seed=seed); tf=time.time()-t0
        out=[tag,'seed',seed,'F',int(m.sum()),'ref',int(g[tag+'_Fmask'].sum()),'jac %.3f'%jac(m,g[tag+'_Fmask']),'%.1fms'%(tf*1e3)]
        for thr in (1.0,3.0):
            t0=time.time(); E,m=utils.calculate_essential_matrix(p1,p2,K,prob=0.999,threshold=thr,seed=seed); te=time.time()-t0
            ref=g['%s_Emask_%d'%(tag,int(thr))]
            out+=['E%d'%thr,int(m.sum()),'ref',int(ref.sum()),'jac %.3f'%jac(m,ref),'%.1fms'%(te*1e3)]
        print(*out)
