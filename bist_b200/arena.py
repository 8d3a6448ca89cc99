"""One zero fill per training step instead of one per module.

Every hand-written backward (modules.py, encoder.py, generator.py) accumulates its parameter gradients into a zero-initialised fp32
block.  Allocated with ``torch.zeros`` that is one fill kernel per block — about 400 two-microsecond launches in a full BiST step,
most of them on the critical path of a captured graph.  Inside ``with arena.step(device):`` the blocks are carved instead from ONE
buffer that is zeroed with ONE memset when the step begins; the sequence of requests is the same every step, so every block gets
the same address every step (a captured graph keeps them, and ``p.grad`` of the previous step is simply overwritten).  Outside the
context ``zeros`` is ``torch.zeros``.  GraphedTrainStep (graph.py) and bench.py's eager loop use it.

Contract: the gradients of step k are dead once step k + 1 begins (the optimizer has consumed them: stream order guarantees it).
"""
import torch

_ALIGN = 64                      # floats (256 bytes)
_arenas = {}
_active = None


class _Arena:
    def __init__(self, device):
        self.device = device
        self.buf = None
        self.off = 0
        self.high = 0            # elements used by the largest step so far
        self.spill = 0           # elements that did not fit this step (served by torch.zeros)

    def begin(self):
        need = max(self.high, self.off + self.spill)
        if self.buf is None or need > self.buf.numel():
            if torch.cuda.is_current_stream_capturing():
                raise RuntimeError("gradient arena too small inside a graph capture (run warm-up steps first)")
            self.buf = torch.empty(int(need * 1.05) + 4096, dtype=torch.float32, device=self.device)
        self.high = max(self.high, need)
        self.off, self.spill = 0, 0
        if self.high:
            self.buf[:self.high].zero_()

    def take(self, n):
        n_al = (n + _ALIGN - 1) // _ALIGN * _ALIGN
        if self.buf is None or self.off + n_al > self.buf.numel() or self.off + n_al > self.high:
            # first step(s): not sized yet.  Serve from torch and remember the demand; the next begin() grows the buffer.
            self.spill += n_al
            return torch.zeros(n, dtype=torch.float32, device=self.device)
        out = self.buf[self.off:self.off + n]
        self.off += n_al
        return out


class step:
    """``with arena.step(device): loss = ...; loss.backward()`` — gradient blocks of this step come from the arena."""

    def __init__(self, device):
        self.device = torch.device(device)

    def __enter__(self):
        global _active
        key = (self.device.type, self.device.index)
        a = _arenas.get(key)
        if a is None:
            a = _arenas[key] = _Arena(self.device)
        a.begin()
        self.prev, _active = _active, a
        return a

    def __exit__(self, *exc):
        global _active
        _active = self.prev
        return False


def zeros(shape, device, dtype=torch.float32):
    """Zero-initialised fp32 gradient block: a view of the step's arena when one is active on `device`, else torch.zeros."""
    if isinstance(shape, int):
        shape = (shape,)
    a = _active
    if a is None or dtype != torch.float32 or torch.device(device) != a.device:
        return torch.zeros(shape, dtype=dtype, device=device)
    n = 1
    for s in shape:
        n *= int(s)
    return a.take(n).view(shape)


def owns(t):
    "True when the tensor lives in a gradient arena (dist.allreduce_grads groups such gradients into contiguous runs)."
    for a in _arenas.values():
        if a.buf is not None and t.untyped_storage().data_ptr() == a.buf.untyped_storage().data_ptr():
            return True
    return False
