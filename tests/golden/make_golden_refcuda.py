"""Generates tests/golden/refcuda_<scene>.npy: images rendered ON A B200 by the reference's own CUDA
kernel and upload code (oracle/_ref/libref_cuda.so = src/C/vr/volumeRender_kernel.cu compiled through
oracle/ref_cuda/texref_shim.h, + src/C/vr/volumeRender.cpp), one fresh process per scene because the
reference keeps its textures and role indices in process globals.

    make -C oracle ref refcuda            # here, where /root/reference exists
    gpurun -- python tests/golden/make_golden_refcuda.py gpurun_out/golden    # on the GPU box
    cp gpurun_out/golden/refcuda_*.npy tests/golden/

The scenes are vrt.golden_scenes(); inputs are regenerated from their seeds by the tests."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))


def one(name, out_dir):
    import numpy as np
    import vrt

    r = vrt.golden_scenes()[name](vrt.RefCudaVolumeRender)
    img = np.asarray(r.render(), dtype=np.float32)
    np.save(os.path.join(out_dir, "refcuda_%s.npy" % name), img)
    print(name, img.shape, float(img.max()), int(np.isnan(img).sum()))


if __name__ == "__main__":
    out_dir = sys.argv[1] if len(sys.argv) > 1 else HERE
    os.makedirs(out_dir, exist_ok=True)
    if len(sys.argv) > 2:
        one(sys.argv[2], out_dir)
    else:
        import vrt

        for name in vrt.golden_scenes():
            subprocess.run([sys.executable, os.path.abspath(__file__), out_dir, name], check=True)
