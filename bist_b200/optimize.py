"""Drop-in mirror of /root/reference/model/optimize.py: ``NoamOpt`` (optimize.py:9-34) and ``SimpleLossCompute``
(optimize.py:36-94).  The loss of every head — the (pointer) generator on ``decoded_text`` against ``trg_y`` and the
auto-encoder heads on ``cap_ft`` / ``temporal_ft`` / ``spatial_ft`` (/ ``audio_ft``) against the query — goes through
``generator.loss_rows``: one fused kernel pass per vocabulary row that yields the label-smoothed KL and its gradient
(bist_genloss_fwd_bwd), normalised by ``ntokens`` / ``qntokens`` on the device (no host synchronisation).
 """
import torch


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
