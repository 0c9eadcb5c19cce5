"""Import shim that executes the reference's own hot-path files verbatim (TEST INFRASTRUCTURE).

Only usable where ``/root/reference`` is mounted (the build container); nothing on the GPU box may call
it.  ``import fairseq`` from the reference fails on Python 3.12 (omegaconf/hydra/bitarray missing and
``fairseq/dataclass/configs.py:1011`` raises on mutable dataclass defaults), so — following SURVEY.md
Appendix B — the package skeleton is faked in ``sys.modules`` and the leaf files that make up the hot
path are imported UNMODIFIED from the read-only mount.  Stand-ins exist only for glue the path does not
compute with (registries, checkpoint utils, matplotlib, editdistance) and for four small helpers whose
bodies are restated from ``fairseq/utils.py:470-481,293-298`` and ``fairseq/data/data_utils.py:518-522``.

Used by ``oracle/make_golden.py`` to (a) validate the restatement in ``oracle/encoder.py`` and (b) write
the golden fixtures under ``tests/golden/``.
"""
from __future__ import annotations

import importlib
import os
import sys
from argparse import Namespace
from types import ModuleType
from unittest.mock import MagicMock

import torch
import torch.nn as nn
import torch.nn.functional as F

REF = os.environ.get("S2T_REFERENCE", "/root/reference")
_loaded = None


def available() -> bool:
    return os.path.isdir(os.path.join(REF, "fairseq", "models", "speech_to_text"))


def _pkg(name: str, path: str) -> ModuleType:
    m = ModuleType(name)
    m.__path__ = [path]
    m.__package__ = name
    sys.modules[name] = m
    return m


def _fake_utils() -> ModuleType:
    u = ModuleType("fairseq.utils")

    def softmax(x, dim: int, onnx_trace: bool = False):       # fairseq/utils.py:470-474
        return F.softmax(x, dim=dim, dtype=torch.float32)

    def log_softmax(x, dim: int, onnx_trace: bool = False):   # fairseq/utils.py:477-481
        return F.log_softmax(x, dim=dim, dtype=torch.float32)

    def make_positions(tensor, padding_idx: int, onnx_trace: bool = False):  # fairseq/utils.py:286-298
        mask = tensor.ne(padding_idx).int()
        return (torch.cumsum(mask, dim=1).type_as(mask) * mask).long() + padding_idx

    def item(t):
        return t.item() if hasattr(t, "item") else t

    def get_activation_fn(activation: str):
        return {"relu": F.relu, "gelu": F.gelu, "tanh": torch.tanh, "linear": lambda x: x,
                "swish": lambda x: x * torch.sigmoid(x)}[activation]

    u.softmax, u.log_softmax, u.make_positions, u.item = softmax, log_softmax, make_positions, item
    u.get_activation_fn = get_activation_fn
    u.get_available_activation_fns = lambda: ["relu", "gelu", "gelu_fast", "gelu_accurate", "tanh", "linear"]
    u.safe_hasattr = lambda obj, k: getattr(obj, k, None) is not None
    u.safe_getattr = lambda obj, k, default=None: getattr(obj, k, default)
    u.eval_str_list = lambda x, type=float: ([type(v) for v in eval(x)] if isinstance(x, str) else x)
    u.has_parameters = lambda module: any(True for _ in module.parameters())
    u.set_incremental_state = lambda *a, **k: None
    u.get_incremental_state = lambda *a, **k: None
    return u


def _fake_data_utils() -> ModuleType:
    d = ModuleType("fairseq.data.data_utils")

    def lengths_to_padding_mask(lens):                        # fairseq/data/data_utils.py:518-522
        bsz, max_lens = lens.size(0), torch.max(lens).item()
        mask = torch.arange(max_lens).to(lens.device).view(1, max_lens)
        return mask.expand(bsz, -1) >= lens.view(bsz, 1).expand(-1, max_lens)

    def post_process(sentence: str, symbol: str):
        return sentence

    d.lengths_to_padding_mask, d.post_process = lengths_to_padding_mask, post_process
    return d


class FakeDictionary:
    """Minimal stand-in for fairseq.data.Dictionary: len() == V, blank/bos 0, pad 1, eos 2, unk 3."""

    def __init__(self, size: int):
        self.size = size

    def __len__(self):
        return self.size

    def pad(self):
        return 1

    def eos(self):
        return 2

    def bos(self):
        return 0

    def unk(self):
        return 3

    def index(self, sym):
        return 0

    def __eq__(self, other):
        return isinstance(other, FakeDictionary) and other.size == self.size

    def __hash__(self):
        return hash(self.size)


class FakeTask:
    def __init__(self, src_vocab: int, tgt_vocab: int):
        self.source_dictionary = FakeDictionary(src_vocab)
        self.target_dictionary = FakeDictionary(tgt_vocab)

    def get_source_dictionary(self, *a, **k):
        return self.source_dictionary


def load():
    """Install the shim and import the reference's hot-path modules.  Returns a namespace of them."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError(f"reference tree not found at {REF} (the shim only works in the build container)")
    os.environ["PYTHONDONTWRITEBYTECODE"] = "1"
    sys.dont_write_bytecode = True
    fs = _pkg("fairseq", REF + "/fairseq")
    mods = _pkg("fairseq.modules", REF + "/fairseq/modules")
    _pkg("fairseq.data", REF + "/fairseq/data")
    _pkg("fairseq.data.audio", REF + "/fairseq/data/audio")
    _pkg("fairseq.data.audio.feature_transforms", REF + "/fairseq/data/audio/feature_transforms")
    models = _pkg("fairseq.models", REF + "/fairseq/models")
    _pkg("fairseq.models.speech_to_text", REF + "/fairseq/models/speech_to_text")

    sys.modules["fairseq.utils"] = fs.utils = _fake_utils()
    sys.modules["fairseq.data.data_utils"] = _fake_data_utils()
    sys.modules["fairseq.checkpoint_utils"] = fs.checkpoint_utils = MagicMock()
    for name in ("matplotlib", "matplotlib.pyplot", "editdistance"):
        sys.modules.setdefault(name, MagicMock())

    # fairseq.models: the real FairseqEncoder, identity registries, plain nn.Module model bases
    fe = importlib.import_module("fairseq.models.fairseq_encoder")
    models.FairseqEncoder = fe.FairseqEncoder
    models.register_model = lambda *a, **k: (lambda cls: cls)
    models.register_model_architecture = lambda *a, **k: (lambda fn: fn)
    models.FairseqEncoderDecoderModel = type("FairseqEncoderDecoderModel", (nn.Module,), {})
    models.FairseqEncoderModel = type("FairseqEncoderModel", (nn.Module,), {})
    tr = ModuleType("fairseq.models.transformer")

    def Embedding(num_embeddings, embedding_dim, padding_idx):   # fairseq/models/transformer.py Embedding
        m = nn.Embedding(num_embeddings, embedding_dim, padding_idx=padding_idx)
        nn.init.normal_(m.weight, mean=0, std=embedding_dim ** -0.5)
        nn.init.constant_(m.weight[padding_idx], 0)
        return m

    tr.Embedding, tr.TransformerDecoder = Embedding, type("TransformerDecoder", (nn.Module,), {})
    sys.modules["fairseq.models.transformer"] = tr

    # feature transforms need a registry decorator only
    ft = sys.modules["fairseq.data.audio.feature_transforms"]
    ft.AudioFeatureTransform = type("AudioFeatureTransform", (), {})
    ft.register_audio_feature_transform = lambda name: (lambda cls: cls)

    def imp(name, *exports):
        m = importlib.import_module(name)
        for e in exports:
            setattr(mods, e, getattr(m, e))
        return m

    for placeholder in ("RelPositionMultiheadAttention", "RelativeMultiheadAttention", "LocalMultiheadAttention",
                        "ReducedMultiheadAttention", "MultiHeadSelfAttentionModule"):
        setattr(mods, placeholder, type(placeholder, (nn.Module,), {}))
    imp("fairseq.modules.layer_norm", "LayerNorm", "Fp32LayerNorm")
    imp("fairseq.modules.activations", "swish", "Swish")
    imp("fairseq.modules.fairseq_dropout", "FairseqDropout")
    imp("fairseq.modules.learned_positional_embedding", "LearnedPositionalEmbedding")
    imp("fairseq.modules.sinusoidal_positional_embedding", "SinusoidalPositionalEmbedding")
    imp("fairseq.modules.positional_embedding", "PositionalEmbedding")
    imp("fairseq.modules.positional_encoding", "PositionalEncoding", "LegacyRelPositionalEncoding",
        "RelPositionalEncoding")
    imp("fairseq.modules.rotary_positional_embedding", "RotaryPositionalEmbedding")
    imp("fairseq.modules.espnet_multihead_attention", "ESPNETMultiHeadedAttention",
        "RelPositionMultiHeadedAttention", "ReducedRelPositionMultiHeadedAttention",
        "LegacyRelPositionMultiHeadedAttention", "RotaryPositionMultiHeadedAttention")
    importlib.import_module("fairseq.incremental_decoding_utils")
    imp("fairseq.modules.quant_noise")
    imp("fairseq.modules.multihead_attention", "MultiheadAttention")
    imp("fairseq.modules.convolution", "ConvolutionModule")
    imp("fairseq.modules.downsample_convolution", "DownSampleConvolutionModule")
    imp("fairseq.modules.dlcl", "DynamicLinearCombination")
    imp("fairseq.modules.s2t_transformer_layer", "S2TTransformerEncoderLayer")
    imp("fairseq.modules.pds_layer", "PDSTransformerEncoderLayer")
    stt = importlib.import_module("fairseq.modules.speech_to_text")
    s2t = importlib.import_module("fairseq.models.speech_to_text.s2t_transformer")
    pds = importlib.import_module("fairseq.models.speech_to_text.pdss2t_transformer")
    ucmvn = importlib.import_module("fairseq.data.audio.feature_transforms.utterance_cmvn")
    gcmvn = importlib.import_module("fairseq.data.audio.feature_transforms.global_cmvn")
    _loaded = Namespace(s2t=s2t, pds=pds, modules=mods, speech_to_text=stt, utterance_cmvn=ucmvn,
                        global_cmvn=gcmvn, FakeTask=FakeTask, FakeDictionary=FakeDictionary)
    return _loaded
