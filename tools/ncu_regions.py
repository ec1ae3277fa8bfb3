"""Aggregate `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass --kernel-name regex:K` by code region (node loop,
triangle test, leaf loop, gates, shading ...): share of executed warp instructions.  usage: python tools/ncu_regions.py src.csv"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
H = None; fname = '?'; agg = {}
import re
src = {}
def region(fname, ln):
    if fname == 'rt_packet.cuh':
        t = open('/root/repo/unity-raytracer_b200/csrc/rt_packet.cuh').read().split('\n')
        # find markers
        global _m
        try: _m
        except NameError:
            _m = {}
            for i, l in enumerate(t, 1):
                if 'rt_packet_pop(STK& stk' in l: _m['pop'] = i
                if 'rt_packet_descend(const float4*' in l: _m['desc'] = i
                if '__device__ __forceinline__ bool rt_traverse_packet' in l: _m['trav'] = i
                if 'int32_t ref = ~node;' in l: _m['leaf'] = i
        if ln >= _m['leaf']: return 'packet leaf loop (loads, accept, gate call)'
        if ln >= _m['trav']: return 'packet setup/dispatch'
        if ln >= _m['desc']: return 'packet node loop'
        if ln >= _m['pop']: return 'packet pop (distance cull)'
        return 'packet other'
    if fname == 'rt_core.h':
        if 100 <= ln <= 127: return 'ray_tri'
        if 129 <= ln <= 141: return 'ray_sphere'
        if 175 <= ln <= 255: return 'per-lane traverse'
        if 70 <= ln <= 86: return 'exact aabb (scene/mesh gates)'
        if 256 <= ln <= 400: return 'shading'
        return 'ray gen / finalize / mapping'
    if fname == 'rt_hd.h':
        t = open('/root/repo/unity-raytracer_b200/csrc/rt_hd.h').read().split('\n')[ln - 1]
        for k in ('rt_fma', '__fmaf_rn', 'um_dot', 'um_cross', '__fdiv_rn', '__fsqrt_rn', 'um_min', 'um_max', 'um_pow', 'pow('):
            if k in t: return 'hd: ' + k
        if 'float r' in t and '* b.' in t: return 'hd: um_cross'
        return 'hd: vector ops'
    if fname == 'rt_render.cu': return 'kernel body (rt_render.cu)'
    return fname
for r in rows:
    if len(r) == 2 and r[0] == "File Path": fname = r[1].split('/')[-1]; continue
    if len(r) > 5 and r[0] == "Line No": H = r; iI = H.index("Instructions Executed"); continue
    if H is None or len(r) < len(H) or r[0] == "": continue
    try: ie = int(r[iI]); ln = int(r[0])
    except ValueError: continue
    if ie <= 0: continue
    k = region(fname, ln)
    agg[k] = agg.get(k, 0) + ie
tot = sum(agg.values())
print("total", tot)
for k, v in sorted(agg.items(), key=lambda x: -x[1]): print(f"{v / tot * 100:5.1f}% {k}")
