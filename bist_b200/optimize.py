"""Drop-in mirror of /root/reference/model/optimize.py: ``NoamOpt`` (optimize.py:9-34) and ``SimpleLossCompute``
(optimize.py:36-94).  The loss of every head — the (pointer) generator on ``decoded_text`` against ``trg_y`` and the
auto-encoder heads on ``cap_ft`` / ``temporal_ft`` / ``spatial_ft`` (/ ``audio_ft``) against the query — goes through
``generator.loss_rows``: one fused kernel pass per vocabulary row that yields the label-smoothed KL and its gradient
(bist_genloss_fwd_bwd), normalised by ``ntokens`` / ``qntokens`` on the device (no host synchronisation).
 """
import torch

from . import _lib


class NoamOpt:
    "optimize.py:9-34 — lr = factor * d_model^-0.5 * min(step^-0.5, step * warmup^-1.5)."

    def __init__(self, model_size, factor, warmup, optimizer):
        self.optimizer = optimizer
        self._step = 0
        self.warmup = warmup
        self.factor = factor
        self.model_size = model_size
        self._rate = 0

    def step(self):
        self._step += 1
        self._rate = self.rate()
        for group in self.optimizer.param_groups:
            group['lr'] = self._rate
        self.optimizer.step()

    def rate(self, step=None):
        step = self._step if step is None else step
        return self.factor * (self.model_size ** (-0.5) * min(step ** (-0.5), step * self.warmup ** (-1.5)))


class SimpleLossCompute:
    "optimize.py:36-94.  ``criterion``: an object with ``padding_idx`` and ``smoothing`` (generator.LabelSmoothing)."

    def __init__(self, generator, ae_generator, criterion, opt=None, l=1.0, args=None):
        self.generator = generator
        self.ae_generator = ae_generator
        self.criterion = criterion
        self.opt = opt
        self.l = l
        self.args = args

    def ae_keys(self):
        a = self.args
        if not a.auto_encoder:
            return []
        if getattr(a, "enc_st_combine", "none") != "none":
            raise NotImplementedError("auto-encoder head on spatiotemporal_ft (enc_st_combine != 'none') is not built")
        keys = (['cap_ft'] if a.nb_cenc_blocks > 0 else []) + (['audio_ft'] if getattr(a, "nb_aenc_blocks", 0) > 0 else [])
        if a.nb_venc_blocks > 0:
            keys += (['temporal_ft'] if a.s2t else []) + (['spatial_ft'] if a.t2s else [])
        return keys

    def loss(self, ft, batch):
        """(total normalised loss, {'out': ..., '<key>_ae': ...}) as device tensors (optimize.py:47-82).

        The auto-encoder heads and the generator loss are independent chains (vocabulary GEMM + fused loss kernel each,
        pointer attentions for the generator): on a CUDA device each head is enqueued on its own branch stream beside
        the generator's chain and joined before the sum (``BIST_BRANCH_STREAMS=0``: one stream)."""
        from .decoder import _branch_streams_enabled, _loss_streams
        from .modules import refresh_weight_copies
        pad, eps = self.criterion.padding_idx, self.criterion.smoothing
        # row-sharded training: dist.sync_token_counts(batch) leaves (global count / world size) in batch.loss_norm, which makes
        # the all-reduce-averaged gradient equal the reference's single-process one (optimize.py:46-50 on the whole batch)
        gn = getattr(batch, "loss_norm", None)
        norm, ae_norm = (gn[0], gn[1]) if gn is not None else (batch.ntokens.float(), batch.qntokens.float())
        keys = self.ae_keys()
        x = ft['decoded_text']
        side = bool(keys) and x.is_cuda and _branch_streams_enabled()
        heads = {}
        if side:
            main = torch.cuda.current_stream(x.device)
            refresh_weight_copies(x.device)
            streams = _loss_streams(x.device, len(keys))
            for key, s in zip(keys, streams):
                s.wait_stream(main)
                with torch.cuda.stream(s):
                    heads[key] = self.ae_generator.loss_rows(ft, batch, self.args, batch.query, pad, eps, key).sum() / ae_norm
        else:
            for key in keys:
                heads[key] = self.ae_generator.loss_rows(ft, batch, self.args, batch.query, pad, eps, key).sum() / ae_norm
        parts = {'out': self.generator.loss_rows(ft, batch, self.args, batch.trg_y, pad, eps).sum() / norm}
        if side:
            for key, s in zip(keys, streams):
                main.wait_stream(s)
                heads[key].record_stream(main)
        for key in keys:
            parts[key + '_ae'] = heads[key]
        total = parts['out']
        for key, v in parts.items():
            if key != 'out':
                total = total + v
        return total, parts

    def __call__(self, ft, batch):
        total, parts = self.loss(ft, batch)
        if self.opt is not None:
            total.backward()
            self.opt.step()
            self.opt.optimizer.zero_grad()
        zero = 0.0
        return {'out': parts['out'].item() * batch.ntokens.float(),
                'temporal_ae': (parts['temporal_ft_ae'].detach() * batch.qntokens.float()) if 'temporal_ft_ae' in parts else zero,
                'spatial_ae': (parts['spatial_ft_ae'].detach() * batch.qntokens.float()) if 'spatial_ft_ae' in parts else zero}


class FusedAdam:
    """torch.optim.Adam's update (no weight decay, no amsgrad — what the reference's NoamOpt wraps, optimize.py:9-34, train.py) for
    all parameters in ONE kernel launch (bist_adam_step): a device table of {p, g, m, v, n} rows, rebuilt only when a gradient
    tensor's address changes (inside a captured training step they never do).  Duck-types the parts of ``torch.optim.Optimizer``
    that NoamOpt and the training loops use: ``param_groups`` (lr, betas, eps), ``step()``, ``zero_grad()``, ``state_dict()``."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        import numpy as np
        self._np = np
        self.params = [p for p in params if p.requires_grad]
        assert self.params and all(p.is_cuda and p.dtype == torch.float32 and p.is_contiguous() for p in self.params), \
            "FusedAdam: contiguous fp32 CUDA parameters only"
        self.param_groups = [{"params": self.params, "lr": lr, "betas": tuple(betas), "eps": eps}]
        dev = self.params[0].device
        n = sum(p.numel() for p in self.params)
        self._m = torch.zeros(n, dtype=torch.float32, device=dev)      # one flat buffer each: views per parameter
        self._v = torch.zeros(n, dtype=torch.float32, device=dev)
        self.state, off = {}, 0
        for p in self.params:
            self.state[p] = {"exp_avg": self._m[off:off + p.numel()].view_as(p), "exp_avg_sq": self._v[off:off + p.numel()].view_as(p)}
            off += p.numel()
        self.step_count = 0
        self._key, self._rows, self._chunks, self._n_chunks, self._live = None, None, None, 0, None

    def _table(self, live):
        np = self._np
        key = tuple((p.data_ptr(), p.grad.data_ptr()) for p in live)
        if key == self._key:
            return
        L = _lib.lib()
        ce = L.bist_adam_chunk_elems()
        rows = np.zeros((len(live), 5), dtype=np.int64)
        chunks = []
        for i, p in enumerate(live):
            g = p.grad
            if g.dtype != torch.float32 or not g.is_contiguous():
                raise RuntimeError("FusedAdam: gradients must be contiguous fp32 tensors")
            st = self.state[p]
            rows[i] = (p.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(), p.numel())
            chunks += [(i, c) for c in range((p.numel() + ce - 1) // ce)]
        dev = live[0].device
        self._rows = torch.from_numpy(rows.view(np.uint8).reshape(-1).copy()).to(dev)
        self._chunks = torch.tensor(chunks, dtype=torch.int32).reshape(-1).to(dev)
        self._n_chunks, self._key, self._live = len(chunks), key, live

    @torch.no_grad()
    def step(self):
        live = [p for p in self.params if p.grad is not None]
        if not live:
            return
        self._table(live)
        self.step_count += 1
        g = self.param_groups[0]
        _lib.check(_lib.lib().bist_adam_step(_lib.ptr(self._rows), _lib.ptr(self._chunks), self._n_chunks, float(g["lr"]), float(g["betas"][0]),
                                             float(g["betas"][1]), float(g["eps"]), self.step_count, _lib.stream_ptr()), "bist_adam_step")
        from .modules import weights_updated
        weights_updated()                      # cached bf16 / packed / transposed weight copies are stale now

    def zero_grad(self, set_to_none=True):
        for p in self.params:
            if set_to_none:
                p.grad = None
            elif p.grad is not None:
                p.grad.zero_()

    def state_dict(self):
        return {"step": self.step_count, "exp_avg": self._m, "exp_avg_sq": self._v, "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups]}

    def load_state_dict(self, sd):
        self.step_count = int(sd["step"])
        self._m.copy_(sd["exp_avg"])
        self._v.copy_(sd["exp_avg_sq"])
