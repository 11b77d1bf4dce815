import ctypes as C, json, sys
sys.path.insert(0, ".")
from mim_solvers_b200 import _capi
lib = _capi.load_library()
a, b, c = C.c_double(), C.c_double(), C.c_double()
rc = lib.mimgpu_measure_peaks(0, C.byref(a), C.byref(b), C.byref(c))
print(json.dumps(dict(rc=rc, dfma_tflops=a.value, dmma_tflops=b.value, hbm_read_gbs=c.value)))
