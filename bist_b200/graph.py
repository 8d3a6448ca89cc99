"""CUDA-graph capture of a training step's forward + loss + backward.

Every kernel of libbist_b200.so is stream-ordered, never synchronises and takes its buffers from torch's caching
allocator, so the forward with its tape, the loss and the whole hand-written backward record into one graph; a replay
costs one launch on the host instead of several hundred.  Three things make a replay equal to an eager step:

  * packed / bf16 / transposed weight copies are re-derived INSIDE the graph (the caches are invalidated right before
    the capture), so a replay after an optimizer step sees the new weights;
  * dropout masks are counter based; a device counter registered with ``bist_set_dropout_step`` is incremented by the
    graph's first node and added to every site's seed at run time, so each replay draws fresh masks;
  * parameter gradients come back in static buffers (``p.grad`` keeps pointing at them): run the optimizer (and
    ``dist.allreduce_grads``) eagerly after the replay and do NOT call ``zero_grad(set_to_none=True)`` in between —
    the graph re-zeroes its gradient blocks itself.
"""
import torch

from . import _lib
from . import arena as _arena
from .modules import weights_updated


class GraphedTrainStep:
    """``step = GraphedTrainStep(loss_fn, params)``; then per iteration: ``loss = step(); allreduce; opt.step()``.

    ``loss_fn()`` must build the loss from STATIC input tensors (copy new batches into them in place) and contain no
    host synchronisation.  ``params``: the parameters whose ``.grad`` the step produces.  No autograd graph over
    these parameters from an earlier eager step may still be alive (drop old loss tensors first): torch keeps their
    AccumulateGrad nodes on the stream they were created on, which invalidates a capture.
    """

    def __init__(self, loss_fn, params, warmup=3, allreduce=False):
        """``allreduce=True`` (N > 1): the gradient all-reduce is captured too — each video layer's block starts its
        NCCL all-reduce as soon as its backward is enqueued (``dist.overlap_grad_allreduce``), the rest follows the
        backward as one grouped launch, all inside the graph; the caller then only runs the optimizer.
        ``self.collectives`` is the number of NCCL launches per replay."""
        from . import dist as _dist
        self.collectives = 0

        def backward(loss):
            if not allreduce:
                loss.backward()
                return
            with _dist.overlap_grad_allreduce():
                loss.backward()
            self.collectives = _dist.allreduce_grads(self.params)

        def step_once():
            # every zero-initialised gradient block of the step is carved from one arena: ONE memset per step (arena.py)
            with _arena.step(dev):
                loss = loss_fn()
                backward(loss)
            return loss

        self.params = [p for p in params if p.requires_grad]
        dev = self.params[0].device
        _lib.require_cuda(self.params[0])
        L = _lib.lib()
        self.counter = torch.zeros(1, dtype=torch.int64, device=dev)
        _lib.check(L.bist_set_dropout_step(self.counter.data_ptr()), "bist_set_dropout_step")
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):            # sets kernel attributes, warms the allocator
                for p in self.params:
                    p.grad = None
                step_once()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        for p in self.params:
            p.grad = None
        weights_updated()                              # every cached weight copy is rebuilt inside the capture
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.counter.add_(1)
            self.loss = step_once()
        weights_updated()                              # the copies made during capture are not valid eager caches

    def __call__(self):
        self.graph.replay()
        weights_updated()
        return self.loss

    def close(self):
        """Unregister the dropout counter (eager dropout goes back to its per-call seeds only)."""
        _lib.check(_lib.lib().bist_set_dropout_step(None), "bist_set_dropout_step")
