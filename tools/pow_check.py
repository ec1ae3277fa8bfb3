"""Development aid (GPU box): float-colour mismatches against the oracle on the reduced cfg3 scene (Blinn-Phong pow path) and on a
scene with non-integer / large Phong exponents (library pow path).  usage: python tools/pow_check.py [lib.so]"""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
rtb = importlib.import_module("unity-raytracer_b200")
if len(sys.argv) > 1: rtb.host.LIB_PATH = sys.argv[1]
from oracle import oracle
from conftest import compare_renders
sc = rtb.scenes
rt = rtb.RayTracer((0,))
for phong in (20.0, 1.0, 64.0, 3.0, 20.5, 100.0, 0.0):
    s, p = sc.cfg3_scene(n=40, phong=phong), sc.cfg3_params(640, 360)
    ref = oracle.OracleScene(s).render(p, oracle.ACCELERATED)
    rt.set_scene(s); got = rt.render(p)
    print("phong", phong, compare_renders(ref, got), flush=True)
