"""bist_b200 — B200-native implementation of the BiST s2t/t2s video hot path.

Drop-in ``nn.Module``s (same constructor/forward signatures and parameter names as
/root/reference/model/modules.py and model/encoder.py) over hand-written sm_100a CUDA
kernels reached through the C-ABI library ``libbist_b200.so`` (include/bist_b200.h).
There is no CPU or eager fallback: using a module without the built library raises.
"""
__version__ = "0.1.0"

_LAZY = {
    "LayerNorm": "modules", "SublayerConnection": "modules", "MultiHeadedAttention": "modules",
    "PositionwiseFeedForward": "modules", "clones": "modules",
    "VidEncoder8": "encoder", "VidEncoderLayer4": "encoder", "VideoPath": "encoder",
    "set_precision": "encoder", "GraphedVideoPath": "encoder", "PipelinedVideoPath": "encoder",
    "CapEncoderLayer": "encoder", "AudioEncoderLayer": "encoder",
    "Encoder": "encoder", "Embeddings": "modules", "PositionalEncoding": "modules", "EmbedSequence": "modules",
    "Generator": "generator", "PointerGenerator": "generator", "MultiPointerGenerator": "generator",
    "LabelSmoothing": "generator", "MTN": "mtn", "make_model": "mtn", "SimpleLossCompute": "optimize", "NoamOpt": "optimize",
    "MultimodalDecoderLayer12": "decoder", "MultimodalDecoder8": "decoder", "make_decoder": "decoder",
}


def __getattr__(name):
    if name in _LAZY:
        import importlib
        return getattr(importlib.import_module(f"{__name__}.{_LAZY[name]}"), name)
    raise AttributeError(name)
