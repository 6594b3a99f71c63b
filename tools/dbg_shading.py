import sys, os
ROOT='/root/repo'
sys.path[:0]=[ROOT+'/tests', ROOT+'/vk-renderer_b200']
import numpy as np, torch, chain, vkr
w,h=1280,720
scene=vkr.synth_scene(0,64)
oc=chain.Chain(w,h,chain.OracleBackend(),lut_size=256)
cc=chain.Chain(w,h,chain.CudaBackend(),luts=(oc.pdf.clone_to('cuda'), oc.brdf.clone_to('cuda')))
P=oc.begin_frame(scene); cc.begin_frame(scene, gbuffer_from=oc.gb)
for (name,ot,outs),(_,ct,_) in zip(oc.steps(P), cc.steps(P)):
    d,s=cc.images(),oc.images()
    for k in d: d[k].buf.copy_(s[k].buf)
    ot(); ct(); torch.cuda.synchronize()
    if name=='DeferedShading':
        g=cc.color_out.download().astype(int); r=oc.color_out.download().astype(int)
        diff=np.abs(g-r)[...,:3].max(axis=2)
        ys,xs=np.nonzero(diff>1)
        print('bad pixels',len(ys))
        dep=oc.gb.depth.download(0); nrm=oc.gb.normal.download(); mat=oc.gb.material.download(); alb=oc.gb.albedo.download()
        d1=oc.gb.depth.download(1)
        for y,x in list(zip(ys,xs))[:12]:
            print((x,y),'got',g[y,x,:3],'ref',r[y,x,:3],'depth',dep[y,x]&0xFFFFFF, 'nbr depth', [int(dep[min(y+dy,h-1),min(x+dx,w-1)]&0xFFFFFF) for dy in (-1,0,1) for dx in (-1,0,1)], 'mat',mat[y,x],'alb',alb[y,x],'nrm',nrm[y,x])
