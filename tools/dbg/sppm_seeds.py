import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np
import rust_raytracer_b200 as rt
from scenes_util import scene_builder
from oracle_ffi import OracleScene
W=H=40; SPP=4; ITER=6; PH=20000
sb=scene_builder("cornell_ref")
ctx=rt.Context(0); scene=rt.Scene(ctx,sb); orc=OracleScene(sb)
G=[];O=[]
for seed in range(1,9):
    p=rt.copy_params(sb.params,width=W,height=H,spp_count=SPP,seed=seed)
    acc,px,st=scene.render_sppm(sb.camera,p,iterations=ITER,photons_per_iter=PH,want_pixel_stats=True)
    G.append((st.global_photons,st.caustic_photons,float(np.nan_to_num(acc[...,:3]/SPP).mean()),float(px[...,4].mean()),float(px[...,3].mean()), st.photon_rays))
    a,s,(ng,nc)=orc.sppm_render(sb.camera,rt.copy_params(p,seed=seed+100),iterations=ITER,photons_per_iter=PH)
    O.append((ng,nc,float(np.nan_to_num(a/SPP,posinf=0).mean()),float(s[...,4].mean()),float(s[...,3].mean()),0))
G=np.array(G,float);O=np.array(O,float)
print("gpu    mean",G.mean(0),"\n       std ",G.std(0,ddof=1))
print("oracle mean",O.mean(0),"\n       std ",O.std(0,ddof=1))
print("rel diff",(G.mean(0)-O.mean(0))/np.maximum(O.mean(0),1e-9),"\nsigma of diff (rel)",np.sqrt(G.var(0,ddof=1)/8+O.var(0,ddof=1)/8)/np.maximum(O.mean(0),1e-9))
