#!/usr/bin/env python
"""Where does a multi-token prompt pass (second window of a long clip, prompt_past carried inside the call: wc:7073-7085) spend its
time?  One large-v3 state, one 65 s clip; prints the scheduler log lines of the general decode path and the profiler categories."""
import ctypes as C, os, sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
os.environ["WB200_SCHED_LOG"] = "/tmp/diag_sched.log"
import conftest as cf
import bench
from whisper_rust_b200 import FullParams, SamplingStrategy, WhisperContext, WhisperContextParameters, api, synth
lib = cf.declare(api.load_library(), cf._B200_SIGS)
lib.whisper_b200_profile.argtypes = [C.c_int]
lib.whisper_b200_profile_read.restype = C.c_int
lib.whisper_b200_profile_read.argtypes = [C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_double), C.POINTER(C.c_double)]
api.install_quiet_logging(lib)
arch = sys.argv[1] if len(sys.argv) > 1 else "large-v3"
if len(sys.argv) > 2 and sys.argv[2] == "rich":   # a recipe that emits timestamps: later windows carry prompt_past (wc:7077-7085)
    ctx = WhisperContext.new_from_buffer_with_params(synth.model_bytes(arch, synth.InitRecipe(te_std=0.7, pe_std=0.3, gain=2.0, ts_scale=1.3)),
                                                     WhisperContextParameters(use_gpu=True), lib=lib)
else:
    ctx = WhisperContext.new_with_params(bench.model_path(arch), WhisperContextParameters(use_gpu=True), lib=lib)
st = ctx.create_state()
p = bench.make_params(lib)
if len(sys.argv) > 3:    # n prompt tokens through FullParams::set_tokens: the first window's prompt pass has n + 4 rows
    p.set_tokens([300 + 7 * i for i in range(int(sys.argv[3]))])
clip = synth.synth_audio(65.0, seed=4242)
st.full(p, clip)                      # warm-up (graphs, workspaces)
open("/tmp/diag_sched.log", "w").close()
lib.whisper_b200_profile(100000)
t0 = time.perf_counter(); st.full(p, clip); dt = time.perf_counter() - t0
print(f"65 s clip alone: {dt*1e3:.1f} ms, segments {st.full_n_segments()}, tokens {sum(st.full_n_tokens(i) for i in range(st.full_n_segments()))}")
n = lib.whisper_b200_profile_read(0, None, 0, None, None, None, None, None)
for c in range(n):
    name = C.create_string_buffer(64); ms, ns, nt, by, fl = C.c_double(), C.c_uint64(), C.c_uint64(), C.c_double(), C.c_double()
    lib.whisper_b200_profile_read(c, name, 64, C.byref(ms), C.byref(ns), C.byref(nt), C.byref(by), C.byref(fl))
    if nt.value: print(f"  {name.value.decode():28s} sampled {ms.value:9.2f} ms over {ns.value:6d} launches (total launches {nt.value})")
print(open("/tmp/diag_sched.log").read())
