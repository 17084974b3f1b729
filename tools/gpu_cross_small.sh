timeout 300 python -m pytest tests/test_gpu_step_kernels.py -x -q 2>&1 | tail -2
python - <<'PY'
import ctypes as C, json, sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from whisper_rust_b200 import api
import torch
lib = api.load_library()
f = lib.whisper_b200_bench_kernel
f.restype = C.c_int
f.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
for n in (1, 2, 4, 7, 8, 16):
    us, by = C.c_double(), C.c_double()
    r = f(b"cross_attn_step", n, 20, 1500, 32, 5, C.byref(us), C.byref(by))
    print(n, r, round(us.value, 2), "us", round(by.value / us.value / 1e3, 1), "GB/s")
PY
