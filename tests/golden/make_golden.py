"""Freezes outputs of the CPU oracle (faithful linear scan) for small cases into tests/golden/*.npz.

    python tests/golden/make_golden.py

The reference (C# on UnityEngine) cannot run in this container or on the GPU box, so these are NOT outputs of the reference
itself; they freeze the oracle — whose primitives are pinned by the KATs in tests/test_oracle.py — so both the oracle
and the CUDA path are checked against committed numbers.  Arrays: hit_id (n,3) int32, hit_t (n,) f32, rgba8 (n,4) u8,
stats (3,) = primary/shadow/mirror ray counts.
"""
import importlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))

CASES = {
    "cfg1_rt_unity_50x50": lambda sc: (sc.rt_unity(), sc.rt_unity_params()),
    "cfg2_mirror_96x54_d4": lambda sc: (sc.rt_unity(mirror_sphere=True), sc.rt_unity_params(96, 54, 4)),
    "soup7_64x48_d3": lambda sc: (sc.random_soup(seed=7), sc.soup_params(64, 48, 3, 1)),
    "soup3_48x32_d2_spp2": lambda sc: (sc.random_soup(seed=3), sc.soup_params(48, 32, 2, 2)),
    # edge cases (builders in tests/conftest.py): NaN / infinite / zero-area / repeated primitives; an oblique and a sheared
    # camera; 40 lights (beyond the fused kernel's 32); reflection chains that run to the bounce limit
    "degenerate_nan_64x40_d2": lambda sc: (_conftest().degenerate_scene(sc, True), sc.soup_params(64, 40, 2, 1)),
    "soup4_oblique_camera_56x40_d2": lambda sc: (sc.random_soup(seed=4, n_tris=500, n_spheres=60),
                                                 sc.make_params(56, 40, 10.0, 8.0, 6.0, (0.1, 0.2, 0.3), 2, 1, cam=_conftest().camera_set()[5])),
    "soup4_sheared_camera_56x40_d2": lambda sc: (sc.random_soup(seed=4, n_tris=500, n_spheres=60),
                                                 sc.make_params(56, 40, 10.0, 8.0, 6.0, (0.1, 0.2, 0.3), 2, 1, cam=_conftest().camera_set()[6])),
    "soup13_40_lights_48x32_d2": lambda sc: (_conftest().many_lights_scene(sc), sc.soup_params(48, 32, 2, 1)),
    "mirror_walls_48x32_d64": lambda sc: (_conftest().mirror_walls_scene(sc), sc.make_params(48, 32, 10.0, 8.0, 5.0, (0.05, 0.05, 0.1), 64, 1)),
}


def _conftest():
    sys.path.insert(0, os.path.dirname(HERE))
    import conftest
    return conftest


def main():
    sys.path.insert(0, ROOT)
    sc = importlib.import_module("unity-raytracer_b200.scenes")
    from oracle import oracle as O
    O.build()
    for name, mk in CASES.items():
        s, p = mk(sc)
        r = O.OracleScene(s).render(p, O.FAITHFUL)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), hit_id=r["hit_id"], hit_t=r["hit_t"], rgba8=r["rgba8"],
                            stats=np.array(r["stats"], np.uint64))
        print(name, r["stats"])


if __name__ == "__main__":
    main()
