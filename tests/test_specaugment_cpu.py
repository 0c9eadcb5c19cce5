"""SpecAugment: the CPU restatement (oracle/specaugment.py) and the product's host-side mask draws
(s2t_b200/specaugment.py) against outputs of the reference class itself (tests/golden/specaugment.npz, written by
oracle/make_golden_specaugment.py from fairseq/data/audio/feature_transforms/specaugment.py).  Bit-exact."""
import numpy as np
import pytest

from pathlib import Path

G = np.load(Path(__file__).parent / "golden" / "specaugment.npz")
NAMES = sorted({k.rsplit("_", 1)[0] for k in G.files if k.endswith("_x")})


def _cfg(name):
    n_f, F, n_t, T = (int(v) for v in G[f"{name}_cfg"])
    v = float(G[f"{name}_value"])
    return dict(freq_mask_n=n_f, freq_mask_f=F, time_mask_n=n_t, time_mask_t=T, time_mask_p=float(G[f"{name}_p"]),
                mask_value=None if np.isnan(v) else v)


@pytest.mark.parametrize("name", NAMES)
def test_oracle_matches_reference_golden(name):
    from oracle.specaugment import specaugment
    np.random.seed(int(G[f"{name}_seed"]))
    with np.errstate(all="ignore"):
        y = specaugment(G[f"{name}_x"], **_cfg(name))
    assert np.array_equal(y, G[f"{name}_y"])


@pytest.mark.parametrize("name", NAMES)
def test_host_draws_replay_the_reference_order(name):
    """Masks drawn by the product module, applied with plain numpy slicing, reproduce the reference output."""
    from s2t_b200.specaugment import SpecAugmentConfig, draw_masks
    c = _cfg(name)
    cfg = SpecAugmentConfig(0, c["freq_mask_n"], c["freq_mask_f"], c["time_mask_n"], c["time_mask_t"], c["time_mask_p"],
                            c["mask_value"])
    x = G[f"{name}_x"]
    np.random.seed(int(G[f"{name}_seed"]))
    fm, tm = draw_masks(cfg, [x.shape[0]], x.shape[1])
    y = x.copy()
    with np.errstate(all="ignore"):
        value = x.mean() if cfg.mask_value is None else cfg.mask_value
    for f0, f in fm[0].tolist():
        if f:
            y[:, f0:f0 + f] = value
    for t0, t in tm[0].tolist():
        if t:
            y[t0:t0 + t, :] = value
    assert np.array_equal(y, G[f"{name}_y"])


def test_config_checks_follow_the_reference():
    from s2t_b200.specaugment import SpecAugmentConfig
    with pytest.raises(ValueError):
        SpecAugmentConfig(freq_mask_N=1, freq_mask_F=0)
    with pytest.raises(ValueError):
        SpecAugmentConfig(time_mask_N=1, time_mask_T=0)
    with pytest.raises(NotImplementedError):
        SpecAugmentConfig(time_warp_W=5)
    c = SpecAugmentConfig.from_config_dict({"freq_mask_N": 2, "freq_mask_F": 27, "time_mask_N": 2, "time_mask_T": 100,
                                            "time_mask_p": 1.0})
    assert c.mask_value is None and c.freq_mask_N == 2
