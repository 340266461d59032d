import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/visual-slam_b200')
import numpy as np
from oracle import oracle as O
from orbslam2.synth import SyntheticVideo
from orbslam2._native import OrbHandle
v=SyntheticVideo(640,480,0); f=v.frame(3)
h=OrbHandle(640,480,max_batch=1)
h.debug_run_levels_host(f)
x,y,s=h.debug_candidates(0,0)
m=O.fast_score_map(f)
rx,ry,rs=O.level_fast_candidates(f,0)
ref={(int(a),int(b)):int(c) for a,b,c in zip(rx,ry,rs)}
got={(int(a),int(b)):int(c) for a,b,c in zip(x,y,s)}
print(len(got),len(ref), len(x))
ex=[k for k in got if k not in ref]; mi=[k for k in ref if k not in got]; df=[k for k in got if k in ref and got[k]!=ref[k]]
print('extra',len(ex),'missing',len(mi),'diffscore',len(df))
for k in ex[:30]:
    print(k, 'got',got[k],'true',m[k[1],k[0]],'tile',k[0]%64,k[1]%32, 'nbmax', m[k[1]-1:k[1]+2,k[0]-1:k[0]+2].max())
for k in df[:10]: print('df',k,got[k],ref[k])
for k in mi[:10]: print('mi',k,ref[k])
