import ctypes as C, sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_pipeline_b200.ops import lib
libs = {"new": lib.load(), "old": C.CDLL(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_tmp", "libdlp_old.so"))}
vp, i32 = C.c_void_p, C.c_int
for L in libs.values():
    L.dlp_gemm_f16.restype = i32
    L.dlp_gemm_f16.argtypes = [vp, vp, i32, vp, i32, vp, i32, i32, i32, i32, i32, i32, i32, vp]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def run(L, w, x, y, epi, up=None):
    T, K = x.shape; N = w.shape[0]
    rc = L.dlp_gemm_f16(w.data_ptr(), 0 if up is None else up.data_ptr(), K, x.data_ptr(), K, y.data_ptr(), y.stride(0), T, N, K, epi, 0, 0, torch.cuda.current_stream().cuda_stream)
    assert rc == 0
def timeit(fn, reps=15):
    for _ in range(3): fn()
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort(); return ts[len(ts) // 2]
for T in (512, 2048):
    for name, N, K, epi, dual in (("qkv", 6144, 4096, 0, False), ("o(add)", 4096, 4096, 2, False), ("gate_up", 14336, 4096, 3, True), ("down(add)", 4096, 14336, 2, False)):
        w = (torch.randn(N, K, device="cuda") * 0.02).half(); up = (torch.randn(N, K, device="cuda") * 0.02).half() if dual else None
        x = torch.randn(T, K, device="cuda").half()
        y = torch.zeros(T, N, device="cuda", dtype=torch.float32 if epi == 2 else torch.float16)
        res = {k: timeit(lambda L=L: run(L, w, x, y, epi, up)) for k, L in libs.items()}
        fl = 2 * T * N * K * (2 if dual else 1)
        print(f"T={T} {name:10s} old {res['old']:8.1f} us ({fl/res['old']/1e6:6.0f} TF)   new {res['new']:8.1f} us ({fl/res['new']/1e6:6.0f} TF)", flush=True)
