import ctypes, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
L = ctypes.CDLL(os.path.join(ROOT, "dbg", "libbist_trace.so"))
c_p, c_i, c_i64, c_f, c_sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float, ctypes.c_size_t
L.bist_vidproj_fwd.argtypes = [c_p]*6 + [c_i64]*3 + [c_f, c_p, c_sz, c_i, c_i, c_p]
L.bist_vidproj_workspace_bytes.restype = c_sz
L.bist_vidproj_workspace_bytes.argtypes = [c_i64]*3 + [c_i]
L.bist_last_error.restype = ctypes.c_char_p
dev = "cuda:0"
N, D, d = 40960, 2048, 512
x = torch.rand(N, D, device=dev); w = torch.randn(d, D, device=dev) * 0.02
b = torch.randn(d, device=dev) * 0.02; a2 = torch.ones(d, device=dev); b2 = torch.zeros(d, device=dev)
out = torch.empty(N, d, device=dev, dtype=torch.bfloat16)
ws = torch.empty(L.bist_vidproj_workspace_bytes(N, D, d, 1), dtype=torch.uint8, device=dev)
for _ in range(3):
    rc = L.bist_vidproj_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), a2.data_ptr(), b2.data_ptr(), out.data_ptr(), N, D, d, 1e-6, ws.data_ptr(), ws.numel(), 1, 1, None)
    assert rc == 0, L.bist_last_error()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    L.bist_vidproj_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), a2.data_ptr(), b2.data_ptr(), out.data_ptr(), N, D, d, 1e-6, ws.data_ptr(), ws.numel(), 1, 1, None)
e1.record(); torch.cuda.synchronize()
print("vidproj us", e0.elapsed_time(e1) * 100)
buf = (ctypes.c_longlong * (4 * 64))()
assert L.bist_debug_vp_trace(buf) == 0
t = [list(buf[i*64:(i+1)*64]) for i in range(4)]
t0 = min(v for v in t[0] if v > 0)
for role, name in enumerate(["tma_issue", "mma_commit", "conv_start", "conv_done"]):
    print(" ", name, [v - t0 if v > 0 else None for v in t[role][:48]])
