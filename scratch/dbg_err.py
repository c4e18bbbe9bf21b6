import sys, numpy as np, torch
sys.path.insert(0,'.'); sys.path.insert(0,'tests')
from helpers import reference_camera
from oracle import pyoracle as O
from brush_b200 import scene, make_uniforms, render_forward
O.set_threads(8)
for (n,w,h,deg,seed,sig) in [(30000,400,300,0,21,(0.5,4.0)),(16384,256,256,0,1001,(0.5,4.0)),(20000,333,200,3,7,(0.5,4.0))]:
    sc=scene.make_scene(n,w,h,deg,seed,sig)
    vm,focal,center=reference_camera(w,h)
    ref,raux=O.render_forward(O.make_uniforms(vm,focal,(w,h),center),sc.means,sc.log_scales,sc.quats,sc.sh_coeffs,sc.raw_opacity)
    T=lambda a: torch.as_tensor(a,device='cuda')
    img,aux=render_forward(make_uniforms(vm,focal,(w,h),center),T(sc.means),T(sc.log_scales),T(sc.quats),T(sc.sh_coeffs),T(sc.raw_opacity),False)
    err=np.abs(img.cpu().numpy()-ref)
    e=err.max(axis=2)
    print(n,w,h,'max',e.max(),'n>1e-5',(e>1e-5).sum(),'n>2e-6',(e>2e-6).sum(),'mean',e.mean(), 'fi mismatch',(aux.final_index.cpu().numpy().view(np.uint32)!=raux.final_index).sum())
    ys,xs=np.where(e>1e-5)
    for y,x in list(zip(ys,xs))[:5]: print('  px',x,y,err[y,x], img.cpu().numpy()[y,x], ref[y,x])
