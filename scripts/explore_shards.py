"""How noisy is relMSE(film, converged)? Sharded vs unsharded GPU films vs the live reference."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import common, pine_b200 as pb
from common import relmse
from oracle import ref
for name in ("c1_64", "c3_48", "mesh_48", "cones_48"):
    g = dict(np.load(os.path.join(common.GOLDEN, name + ".npz")))
    R, S = ref.RefScene(), pb.Scene()
    common.build(R, name); common.build(S, name)
    depth = common.SMALL_SCENES[name][3]
    B = pb.Backend(0); B.upload_scene(S)
    for spp in (16, 64, 256):
        fr, _ = R.render(spp, depth)
        e_ref = relmse(fr, g["film_converged"])
        e_gpu = relmse(B.render(spp, depth), g["film_converged"])
        out = [f"{name} spp={spp} e_ref={e_ref:.5f} e_gpu={e_gpu:.5f}"]
        for n in (2, 4, 8):
            acc = sum(B.render(spp, depth, samples=(s, s + spp // n)).astype(np.float64) for s in range(0, spp, spp // n))
            out.append(f"sh{n}={relmse(acc, g['film_converged']):.5f}")
        print(" ".join(out), flush=True)
    B.close()
