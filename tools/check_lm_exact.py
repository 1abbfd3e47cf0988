import numpy as np, sys
sys.path.insert(0,'.')
import __graft_entry__
from oracle import pyorc
from tools.synth import synth
from lidarslam_b200 import capi
from lidarslam_b200.slam import Slam
g = Slam(device=0); o = pyorc.Oracle(nb_threads=pyorc.max_threads())
seq = synth.Sequence(); sens = synth.sensor("vlp16")
p0,T0,_ = seq.frame(sens,0); p1,T1,_ = seq.frame(sens,1)
_,_,k0 = o.ExtractKeypoints(p0); _,_,k1 = o.ExtractKeypoints(p1)
prior = np.linalg.inv(T0)@T1; prior[:3,3]+= [0.05,-0.03,0.01]
for single in (0,1):
  for it in (1,3,15,40):
    mp = capi.MatchParams(single,5.0,8 if single else 10,3 if single else 4,0.2,5,0.04,0.2,10,2.0)
    rg = g.MatchAndSolve(k1,k0,mp,it,prior); ro = o.MatchAndSolve(k1,k0,mp,it,prior)
    print(single, it, rg['iterations'], ro['iterations'], rg['accepted'], ro['accepted'], 'pose maxdiff', np.abs(rg['pose']-ro['pose']).max(), 'cost diff', rg['cost']-ro['cost'],
          'bit-equal', np.array_equal(rg['pose'], ro['pose']) and rg['cost']==ro['cost'])
