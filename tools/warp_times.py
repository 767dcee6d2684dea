"""Developer probe: distribution of per-warp finish times of the ray-march kernel (needs VR_WARP_TIMES)."""
import ctypes as C, os, sys, numpy as np
os.environ["VR_WARP_TIMES"]="1"
sys.path.insert(0,"/root/repo"); sys.path.insert(0,"/root/repo/tests")
import torch
from volume_renderer_b200 import _abi, multigpu
import bench
args = bench.parse()
lib=_abi.lib(); ctx=C.c_void_p(); _abi.check(lib.vr_create(C.byref(ctx)))
scene=bench.Scene(args); n=args.size
dims=(C.c_uint64*3)(n,n,n)
E=torch.empty(n**3,dtype=torch.float32,device="cuda"); _abi.check(lib.vr_synth_volume(ctx,0,bench.SEED,dims,0,C.c_void_p(E.data_ptr())),ctx)
def vol(t,st):
    v=_abi.VrVolume(); v.data=t.data_ptr(); v.dims[0]=v.dims[1]=v.dims[2]=n; v.last_update=st; v.dtype=0; v.memspace=1; return v
vE=vol(E,1); R=E*E; vR=vol(R,2)
G=[torch.empty(n**3,dtype=torch.float32,device="cuda") for _ in range(3)]
_abi.check(lib.vr_volume_gradient(ctx,C.byref(vE),1,*[C.c_void_p(t.data_ptr()) for t in G]),ctx)
g=[vol(G[i],4+i) for i in range(3)]
_abi.check(lib.vr_sync_volumes(ctx,0,C.byref(vE),C.byref(vR),C.byref(vE),*[C.byref(x) for x in g]),ctx)
lut=np.zeros(64**3,dtype=np.float32); _abi.check(lib.vr_henyey_greenstein(64,0.8,lut.ctypes.data))
lutv=_abi.VrVolume(); lutv.data=lut.ctypes.data; lutv.dims[0]=lutv.dims[1]=lutv.dims[2]=64; lutv.last_update=9
lights=(_abi.VrLight*2)()
for i,(p,c) in enumerate(scene.lights): lights[i].position[:]=p; lights[i].color[:]=c
img=torch.zeros(scene.W*scene.H*3,device="cuda")
resp=scene.res.ctypes.data_as(C.POINTER(C.c_uint64)); f=bench.f32p
for nr in (1,2,8):
  multigpu.configure_partition(ctx,0,nr,args.tile,args.tile)
  for it in range(3):
    rot=scene.rot(0)
    _abi.check(lib.vr_render_device(ctx,lights,2,C.byref(lutv),f(scene.factors),f(scene.es),resp,f(rot),f(scene.props),C.c_float(scene.thr),f(scene.color),C.c_void_p(img.data_ptr()),None),ctx)
    s=_abi.VrStats(); _abi.check(lib.vr_get_stats(ctx,C.byref(s)),ctx)
  NW=148*32; wt=np.zeros(2*NW,dtype=np.uint64); _abi.check(lib.vr_debug_warp_times(ctx,wt.ctypes.data,NW),ctx)
  st=wt[0::2].astype(np.int64); en=wt[1::2].astype(np.int64); t0=st.min()
  end=(en-t0)/1e6
  print("nranks",nr,"kernel_ms %.2f"%s.kernel_ms,"warp end ms: min %.2f p10 %.2f p50 %.2f p90 %.2f p99 %.2f max %.2f mean %.2f"%(end.min(),*np.percentile(end,[10,50,90,99]),end.max(),end.mean()), "still running at T-4,-3,-2,-1,-0.5 ms:", [int((end>end.max()-d).sum()) for d in (4,3,2,1,0.5)])
