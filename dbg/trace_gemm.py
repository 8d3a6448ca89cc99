import ctypes, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
L = ctypes.CDLL(os.path.join(ROOT, "dbg", "libbist_trace.so"))
c_p, c_i, c_i64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64
L.bist_linear_fwd.argtypes = [c_p]*5 + [c_i64]*3 + [c_i]*4 + [c_p]
L.bist_last_error.restype = ctypes.c_char_p
dev = "cuda:0"
for (M, N, K) in [(40960, 1024, 512), (2048, 512, 512)]:
    x = torch.randn(M, K, device=dev).to(torch.bfloat16)
    w = torch.randn(N, K, device=dev).to(torch.bfloat16)
    y = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    for _ in range(3):
        rc = L.bist_linear_fwd(x.data_ptr(), w.data_ptr(), None, None, y.data_ptr(), M, N, K, 0, 1, 1, 1, None)
        assert rc == 0, L.bist_last_error()
    torch.cuda.synchronize()
    buf = (ctypes.c_longlong * (3 * 64))()
    assert L.bist_debug_gemm_trace(buf) == 0
    t = [list(buf[i*64:(i+1)*64]) for i in range(3)]
    t0 = min(v for v in t[0] if v > 0)
    print("shape", M, N, K)
    for role, name in enumerate(["tma_issue", "mma_commit", "epi(start,end)"]):
        print(" ", name, [v - t0 if v > 0 else None for v in t[role][:40]])
