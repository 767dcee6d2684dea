import sys,os,numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import vrt
from volume_renderer_b200 import _abi, api
from test_golden_refcuda import envelope
for name in sorted(vrt.golden_scenes()):
    g=vrt.golden_scenes()[name](api.VolumeRender); g.setOption(_abi.OPT_TEXTURE_UNIT,1)
    o=vrt.golden_scenes()[name](vrt.OracleVolumeRender); o._backend.lib.orc_set_option(o.objectHandle, vrt.ORC_OPT_FILTER_FRAC_BITS, 9)
    img,ref=g.render(),o.render()
    m=vrt.compare(img,ref); e=envelope(img,ref)
    print('%-28s max_abs_norm %.2e psnr %.1f within1e-3 %.5f'%(name,m['max_abs_norm'],m['psnr_db'],e['frac_within_1e-3']), flush=True)
brain=vrt.synth_brain((64,64,64)); lut=api.HenyeyGreenstein(64)
for lights in (0,1,2):
  for res in ((160,128),(257,131)):
    g=vrt.example1_scene(api.VolumeRender, brain, res, lights=lights, lut=lut); g.setOption(_abi.OPT_TEXTURE_UNIT,1)
    o=vrt.example1_scene(vrt.OracleVolumeRender, brain, res, lights=lights, lut=lut); o._backend.lib.orc_set_option(o.objectHandle, vrt.ORC_OPT_FILTER_FRAC_BITS, 9)
    if lights==0: g.LightSources=False; o.LightSources=False
    img,ref=g.render(),o.render(); m=vrt.compare(img,ref)
    print('example1 lights %d res %s max_abs_norm %.2e psnr %.1f'%(lights,res,m['max_abs_norm'],m['psnr_db']), flush=True)
