"""CPU (tests/host_emul + oracle/_ref): where the closest hit of a triangle mesh depends on the TREE.
The reference does not range-check triangle hits; what bounds them is box culling, so with a different tree
(our BVH8 instead of the reference's own) results can differ when the eye is inside the mesh or the depth buffer is
pre-filled.  Counts depth mismatches of our BVH8 (kernel source on the CPU) against the compiled reference per situation.
usage: PYTHONPATH=. python tools/tree_dependence.py   (needs oracle/_ref)"""
import numpy as np

import triangles_sdf_cpu_raytracing_b200 as rt
from oracle import ref_oracle as R
from tests import util
from tests import host_emul as E
G=util.Golden()
W,H=64,40
for name in ("stanford-bunny","spot","as1-oc-214","cube"):
    a=G.scene_arrays(name)
    ro=R.RefOracle(); ref=ro.mesh(a["vpos4"],a["indices"])
    s=E.EmulScene.mesh(a["vpos4"],a["indices"],brute=False)
    pinv=ro.projection(W,H)[1]
    rng=np.random.default_rng(11)
    for radii in ([1.8,2.5,4.0],[1.0,1.3],[0.2,0.5]):
        for prefill in (False,True):
            for away in (False,True):
                bad=0; n=0
                for k in range(10):
                    v=rng.normal(size=3); v/=np.linalg.norm(v); r=float(rng.choice(radii))
                    eye=tuple(float(x) for x in r*v)
                    target=tuple(float(x) for x in (2.5*r*v+rng.uniform(-0.5,0.5,3))) if away else tuple(float(x) for x in rng.uniform(-0.3,0.3,3))
                    cam=R.CamSpec(eye=eye,target=target,lock_up=bool(k%2))
                    pos,_,vinv=ro.camera(cam)
                    t0=(rng.choice(np.array([np.inf,0.5,1.2,2.0,2.6,3.5],np.float32),size=(H,W)).astype(np.float32) if prefill else np.full((H,W),np.inf,np.float32))
                    c0=rng.integers(0,2**32,size=(H,W),dtype=np.uint32)
                    wc,wt,_=ref.render(cam,W,H,proj_inv=pinv,color=c0,t=t0)
                    c,t,_,_=s.render(rt.make_frame(pos,vinv,pinv,W,H),color=c0,t=t0)
                    bad+=int((util.bits(t)!=util.bits(wt)).sum()); n+=W*H
                print(f"{name:15s} radii {str(radii):16s} prefill {prefill!s:5s} away {away!s:5s} mismatches {bad} of {n}")
