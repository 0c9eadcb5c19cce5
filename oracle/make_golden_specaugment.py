"""Golden vectors for SpecAugment from the reference class itself (TEST INFRASTRUCTURE; needs /root/reference).

Loads fairseq/data/audio/feature_transforms/specaugment.py UNMODIFIED — its only fairseq import is the transform
registry (`AudioFeatureTransform`, `register_audio_feature_transform`), faked here as a plain base class and an
identity decorator — runs `SpecAugmentTransform.__call__` under seeded numpy global RNG states, checks the restatement
in oracle/specaugment.py against it bit for bit, and writes tests/golden/specaugment.npz.

    python -m oracle.make_golden_specaugment
"""
import importlib.util
import os
import sys
from pathlib import Path
from types import ModuleType

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
REF = os.environ.get("S2T_REFERENCE", "/root/reference")

# (name, T, F, config, seed); policies from examples/speech_to_text/data_utils.py gen_config_yaml ("lb", "ld", "sm")
CASES = [
    ("lb", 998, 80, dict(freq_mask_n=1, freq_mask_f=27, time_mask_n=1, time_mask_t=100, time_mask_p=1.0, mask_value=None), 11),
    ("ld", 731, 80, dict(freq_mask_n=2, freq_mask_f=27, time_mask_n=2, time_mask_t=100, time_mask_p=1.0, mask_value=None), 12),
    ("sm", 412, 80, dict(freq_mask_n=2, freq_mask_f=15, time_mask_n=2, time_mask_t=70, time_mask_p=0.2, mask_value=0.0), 13),
    ("short", 3, 80, dict(freq_mask_n=2, freq_mask_f=27, time_mask_n=2, time_mask_t=100, time_mask_p=0.2, mask_value=0.0), 14),
    ("narrow", 50, 20, dict(freq_mask_n=1, freq_mask_f=27, time_mask_n=1, time_mask_t=10, time_mask_p=1.0, mask_value=1.5), 15),
    ("empty", 0, 80, dict(freq_mask_n=1, freq_mask_f=27, time_mask_n=1, time_mask_t=100, time_mask_p=1.0, mask_value=0.0), 16),
]


def load_reference_class():
    for name in ("fairseq", "fairseq.data", "fairseq.data.audio"):
        sys.modules.setdefault(name, ModuleType(name))
    reg = ModuleType("fairseq.data.audio.feature_transforms")

    class AudioFeatureTransform:            # registry glue only (feature_transforms/__init__.py)
        pass

    reg.AudioFeatureTransform = AudioFeatureTransform
    reg.register_audio_feature_transform = lambda name: (lambda cls: cls)
    sys.modules["fairseq.data.audio.feature_transforms"] = reg
    path = os.path.join(REF, "fairseq", "data", "audio", "feature_transforms", "specaugment.py")
    spec = importlib.util.spec_from_file_location("fairseq.data.audio.feature_transforms.specaugment", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.SpecAugmentTransform


def main():
    sys.path.insert(0, str(ROOT))
    from oracle.specaugment import specaugment
    cls = load_reference_class()
    out = {}
    for name, T, F, cfg, seed in CASES:
        x = np.random.RandomState(100 + seed).randn(T, F).astype(np.float32) * 1.3 + 0.05
        tr = cls(0, cfg["freq_mask_n"], cfg["freq_mask_f"], cfg["time_mask_n"], cfg["time_mask_t"], cfg["time_mask_p"],
                 cfg["mask_value"])
        np.random.seed(seed)
        with np.errstate(all="ignore"):
            y_ref = tr(x)
        np.random.seed(seed)
        with np.errstate(all="ignore"):
            y = specaugment(x, **cfg)
        assert y.shape == y_ref.shape and np.array_equal(y, y_ref), name      # bit-exact restatement
        out[f"{name}_x"] = x
        out[f"{name}_y"] = y_ref
        out[f"{name}_seed"] = np.int64(seed)
        out[f"{name}_cfg"] = np.array([cfg["freq_mask_n"], cfg["freq_mask_f"], cfg["time_mask_n"], cfg["time_mask_t"]],
                                      dtype=np.int64)
        out[f"{name}_p"] = np.float64(cfg["time_mask_p"])
        out[f"{name}_value"] = np.float64(np.nan if cfg["mask_value"] is None else cfg["mask_value"])
        print(f"{name}: T={T} F={F} changed {int((y_ref != x).sum())} elements, oracle == reference")
    np.savez_compressed(ROOT / "tests" / "golden" / "specaugment.npz", **out)


if __name__ == "__main__":
    main()
