"""Runs the two roofline microbenchmarks (unity-raytracer_b200/csrc/rt_microbench.cu) and prints JSON."""
import ctypes as C, json, os, sys
lib = C.CDLL(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "unity-raytracer_b200", "librtb200_microbench.so"))
lib.rtmb_l2_read_gbs.restype = C.c_double; lib.rtmb_l2_read_gbs.argtypes = [C.c_int, C.c_int, C.c_int]
lib.rtmb_fp32_tflops.restype = C.c_double; lib.rtmb_fp32_tflops.argtypes = [C.c_int, C.c_int]
out = {"l2_read_gbs": {f"{mb}MiB": round(lib.rtmb_l2_read_gbs(0, mb, 200 if mb <= 64 else 20), 1) for mb in (16, 32, 48, 64, 96, 512)},
       "fp32_tflops": round(lib.rtmb_fp32_tflops(0, 1 << 16), 2)}
print(json.dumps(out))
