"""ctypes binding of libbist_b200.so (C ABI: include/bist_b200.h) and its in-tree build.

The library is compiled from bist_b200/csrc/*.cu with nvcc for sm_100a only and kept
in-tree at bist_b200/lib/libbist_b200.so (git-ignored, shipped to the GPU box by gpurun).
There is no fallback: if the library is missing or a call fails, a RuntimeError is raised.
"""
import ctypes
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libbist_b200.so")
OBJ_DIR = os.path.join(LIB_DIR, "obj")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC"]

F32, BF16 = 0, 1
MASK_NONE, MASK_KEY, MASK_FULL = 0, 1, 2

c_p, c_i, c_i64, c_f, c_sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float, ctypes.c_size_t


class PackItem(ctypes.Structure):
    _fields_ = [("src", ctypes.c_void_p), ("dst_off", ctypes.c_int64), ("rows", ctypes.c_int32), ("cols", ctypes.c_int32),
                ("transpose", ctypes.c_int32), ("reserved", ctypes.c_int32)]


class MhaParams(ctypes.Structure):
    _fields_ = [("w", c_p * 4), ("b", c_p * 4)]


class Dropout(ctypes.Structure):
    "bist_dropout_t"
    _fields_ = [("p_attn", ctypes.c_float), ("p_sublayer", ctypes.c_float), ("p_ffn", ctypes.c_float),
                ("seed", ctypes.c_uint64)]


class StLayerParams(ctypes.Structure):
    _fields_ = [("attn_w", (c_p * 4) * 6), ("attn_b", (c_p * 4) * 6),
                ("ff_w1", c_p * 2), ("ff_b1", c_p * 2), ("ff_w2", c_p * 2), ("ff_b2", c_p * 2),
                ("ln_a", c_p * 8), ("ln_b", c_p * 8)]


# name -> (restype, argtypes); every symbol declared in include/bist_b200.h
SIGNATURES = {
    "bist_version": (c_i, []),
    "bist_last_error": (ctypes.c_char_p, []),
    "bist_device_check": (c_i, []),
    "bist_launch_count": (c_i64, []),
    "bist_prof_enable": (c_i, [c_i]),
    "bist_prof_collect": (c_i, [ctypes.c_char_p, c_sz]),
    "bist_ln_fwd": (c_i, [c_p, c_p, c_p, c_p, c_i64, c_i64, c_f, c_i, c_i, c_p]),
    "bist_linear_fwd": (c_i, [c_p, c_p, c_p, c_p, c_p, c_i64, c_i64, c_i64, c_i, c_i, c_i, c_i, c_p]),
    "bist_mha_workspace_bytes": (c_sz, [c_i64] * 6),
    "bist_mha_fwd": (c_i, [c_p, c_p, c_p, c_p, c_i, ctypes.POINTER(MhaParams), c_p, c_p,
                           c_i64, c_i64, c_i64, c_i64, c_i64, c_i64, c_p, c_sz, c_i, c_p]),
    "bist_ffn_fwd": (c_i, [c_p, c_p, c_p, c_p, c_p, c_p, c_i64, c_i64, c_i64, c_i64, c_p, c_sz, c_i, c_p]),
    "bist_vidproj_workspace_bytes": (c_sz, [c_i64, c_i64, c_i64, c_i]),
    "bist_vidproj_fwd": (c_i, [c_p, c_p, c_p, c_p, c_p, c_p, c_i64, c_i64, c_i64, c_f, c_p, c_sz, c_i, c_i, c_p]),
    "bist_st_layer_packed_bytes": (c_sz, [c_i64, c_i64, c_i, c_i]),
    "bist_st_layer_pack": (c_i, [ctypes.POINTER(StLayerParams), c_p, c_sz, c_i64, c_i64, c_i64, c_i, c_i, c_p]),
    "bist_st_layer_workspace_bytes": (c_sz, [c_i64] * 7 + [c_i, c_i]),
    "bist_st_layer_fwd": (c_i, [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p,
                                c_i64, c_i64, c_i64, c_i64, c_i64, c_i64, c_i64,
                                c_i, c_i, c_f, c_p, c_sz, c_i, c_p]),
    # training step
    "bist_vidproj_tape_bytes": (c_sz, [c_i64, c_i64, c_i64, c_i]),
    "bist_vidproj_train_fwd": (c_i, [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_sz, c_i64, c_i64, c_i64, c_f, c_i, c_i, c_p]),
    "bist_vidproj_bwd_workspace_bytes": (c_sz, [c_i64, c_i64, c_i64, c_i]),
    "bist_vidproj_bwd": (c_i, [c_p, c_p, c_p, c_sz] + [c_p] * 5 + [c_i64, c_i64, c_i64, c_f, c_p, c_sz, c_i, c_p]),
    "bist_st_layer_train_packed_bytes": (c_sz, [c_i64, c_i64, c_i, c_i]),
    "bist_st_layer_train_pack": (c_i, [ctypes.POINTER(StLayerParams), c_p, c_sz, c_i64, c_i64, c_i64, c_i, c_i, c_p]),
    "bist_st_layer_grad_elems": (c_sz, [c_i64, c_i64, c_i]),
    "bist_st_layer_tape_bytes": (c_sz, [c_i64] * 7 + [c_i, c_i]),
    "bist_st_layer_bwd_workspace_bytes": (c_sz, [c_i64] * 7 + [c_i, c_i]),
    "bist_st_layer_train_fwd": (c_i, [c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p,
                                      c_i64, c_i64, c_i64, c_i64, c_i64, c_i64, c_i64,
                                      c_i, c_i, c_f, c_p, c_sz, c_i, ctypes.POINTER(Dropout), c_p]),
    "bist_st_layer_bwd": (c_i, [c_p] * 9 + [c_sz] + [c_p] * 4 + [c_i64] * 7 +
                          [c_i, c_i, c_f, c_p, c_sz, c_i, ctypes.POINTER(Dropout), c_p]),
    "bist_dropout_mask": (c_i, [ctypes.c_uint64, ctypes.c_uint32, c_f, c_i64, c_p, c_p]),
    "bist_ln_bwd": (c_i, [c_p, c_p, c_p, c_p, c_p, c_p, c_i64, c_i64, c_f, c_i, c_i, c_i, c_p]),
    "bist_linear_bwd": (c_i, [c_p, c_p, c_p, c_p, c_p, c_p, c_i64, c_i64, c_i64, c_i, c_i, c_i, c_i, c_p]),
    "bist_dropout_fwd": (c_i, [c_p, c_p, c_i64, c_i64, c_f, ctypes.c_uint64, ctypes.c_uint32, c_i, c_p]),
    "bist_pack_params": (c_i, [ctypes.POINTER(PackItem), c_i64, c_p, c_i, c_p]),
    "bist_dropout_add_fwd": (c_i, [c_p, c_p, c_p, c_i64, c_i64, c_f, ctypes.c_uint64, ctypes.c_uint32, c_i, c_p]),
    "bist_mha_train_fwd": (c_i, [c_p, c_p, c_p, c_p, c_i, ctypes.POINTER(MhaParams), c_p, c_p,
                                 c_i64, c_i64, c_i64, c_i64, c_i64, c_i64, c_p, c_sz, c_i,
                                 c_f, ctypes.c_uint64, ctypes.c_uint32, c_p]),
    "bist_mha_bwd_workspace_bytes": (c_sz, [c_i64, c_i64, c_i64, c_i64, c_i]),
    "bist_mha_bwd": (c_i, [c_p, c_p, c_p, c_p, c_p, c_i, ctypes.POINTER(MhaParams), c_p, c_sz, c_p, c_p, c_p,
                           ctypes.POINTER(c_p), ctypes.POINTER(c_p), c_i64, c_i64, c_i64, c_i64, c_i64, c_i64,
                           c_p, c_sz, c_i, c_f, ctypes.c_uint64, ctypes.c_uint32, c_p]),
    "bist_ffn_train_fwd": (c_i, [c_p] * 7 + [c_i64] * 4 + [c_i, c_f, ctypes.c_uint64, ctypes.c_uint32, c_p]),
    "bist_ffn_bwd": (c_i, [c_p] * 10 + [c_i64] * 4 + [c_p, c_sz, c_i, c_f, c_p]),
    "bist_set_dropout_step": (c_i, [c_p]),
    "bist_ingest_gather": (c_i, [ctypes.POINTER(c_p), c_p, ctypes.POINTER(c_i64), ctypes.POINTER(c_i64),
                                 ctypes.POINTER(c_i64), ctypes.POINTER(c_i64), c_i64, c_p]),
    "bist_vidproj_fwd_bf16in": (c_i, [c_p] * 6 + [c_i64] * 3 + [c_f, c_p, c_sz, c_i, c_p]),
    "bist_debug_qside": (c_i, [c_i]),
    "bist_adam_chunk_elems": (c_i, []),
    "bist_adam_step": (c_i, [c_p, c_p, c_i, c_f, c_f, c_f, c_f, c_i64, c_p]),
    "bist_vidproj_fwd_noalloc": (c_i, [c_p] * 6 + [c_i64] * 3 + [c_f, c_i, c_p]),
    "bist_embed_fwd": (c_i, [c_p, c_p, c_p, c_p, c_i64, c_i64, c_i64, c_i64, c_f, ctypes.c_uint64, ctypes.c_uint32, c_i, c_p]),
    "bist_embed_bwd": (c_i, [c_p, c_p, c_p, c_i64, c_i64, c_f, ctypes.c_uint64, ctypes.c_uint32, c_i, c_p]),
    "bist_ptr_attn_fwd": (c_i, [c_p] * 6 + [c_i64] * 4 + [c_i, c_p]),
    "bist_ptr_attn_bwd": (c_i, [c_p] * 9 + [c_i64] * 4 + [c_i, c_p]),
    "bist_genloss_fwd_bwd": (c_i, [c_p, c_p, c_i, ctypes.POINTER(c_p), ctypes.POINTER(c_p), ctypes.POINTER(c_p),
                                   ctypes.POINTER(c_i64), c_p, c_p, c_p, c_p, c_i64, c_i64, c_i64, c_i64, c_f, c_i, c_i, c_p]),
    "bist_gate_combine_fwd": (c_i, [ctypes.POINTER(c_p), c_p, c_p, c_i64, c_i64, c_i, c_i, c_p]),
    "bist_stage1_fused_supported": (c_i, [c_i64] * 6),
    "bist_stage1_fused": (c_i, [c_p, c_p, c_i, ctypes.POINTER(c_i), ctypes.POINTER(c_p), ctypes.POINTER(c_p),
                                ctypes.POINTER(c_p), ctypes.POINTER(c_p)] + [c_i64] * 6 + [c_p, c_p]),
    "bist_gate_combine_bwd": (c_i, [c_p, ctypes.POINTER(c_p), c_p, ctypes.POINTER(c_p), c_p, c_i64, c_i64, c_i, c_i, c_p]),
}


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _stale():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = sources() + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    deps.append(os.path.join(os.path.dirname(HERE), "include", "bist_b200.h"))
    return any(os.path.getmtime(p) > t for p in deps)


def build(force=False, verbose=False):
    """Compile csrc/*.cu for sm_100a into lib/libbist_b200.so (parallel per-file, then link)."""
    if not force and not _stale():
        return LIB_PATH
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    os.makedirs(OBJ_DIR, exist_ok=True)

    def compile_one(src):
        obj = os.path.join(OBJ_DIR, os.path.basename(src)[:-3] + ".o")
        cmd = [nvcc] + NVCC_FLAGS + ["-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed: %s\n%s%s" % (" ".join(cmd), r.stdout, r.stderr))
        if verbose and r.stderr:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, sources()))
    cmd = [nvcc, "-shared", "-o", LIB_PATH] + objs + ["-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed: %s\n%s%s" % (" ".join(cmd), r.stdout, r.stderr))
    return LIB_PATH


_LIB = None


def lib():
    """Load (once) and return the ctypes handle; raises if the library has not been built."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "libbist_b200.so is not built (%s). Run `python -c 'import __graft_entry__ as g; g.build()'`. "
                "bist_b200 has no CPU or eager fallback." % LIB_PATH)
        h = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(h, name)          # AttributeError here = header and library disagree
            fn.restype, fn.argtypes = res, args
        _LIB = h
    return _LIB


def check(rc, what):
    if rc != 0:
        raise RuntimeError("%s failed (status %d): %s" % (what, rc, lib().bist_last_error().decode()))


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def dtype_code(torch_dtype):
    import torch
    if torch_dtype == torch.float32:
        return F32
    if torch_dtype == torch.bfloat16:
        return BF16
    raise TypeError("bist_b200 supports float32 and bfloat16, got %s" % torch_dtype)


def stream_ptr():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("bist_b200 runs on CUDA tensors only (no CPU fallback); got a %s tensor" % t.device)
