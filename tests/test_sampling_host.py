"""Host-side pieces of the sampler mirror (no GPU): schedule construction."""
import torch

from owl_audio_exps_b200.sampling import get_deltas, get_sampler_cls, get_sd3_euler, AudioCachingSampler


def test_get_deltas_appends_final_zero():
    assert get_deltas([1.0, 0.75, 0.4]) == [0.25, 0.35, 0.4]
    assert get_deltas([1.0, 0.5, 0.0]) == [0.5, 0.5]
    sched = [1.0, 0.6]
    get_deltas(sched)
    assert sched == [1.0, 0.6]          # the caller's list is left alone


def test_sd3_euler_closed_form_properties():
    for n in (1, 4, 16):
        dt = get_sd3_euler(n)
        assert dt.shape == (n,) and (dt > 0).all()
        assert abs(dt.sum().item() - 1.0) < 1e-6            # sigma_0 = 3*1/(1+2*1) = 1 -> the steps walk t from 1 to 0
    dt = get_sd3_euler(2)                                    # sigma = [1, .5] -> shifted [1, .75]
    assert torch.allclose(dt, torch.tensor([0.25, 0.75]))


def test_sampler_registry():
    assert get_sampler_cls("audio_caching") is AudioCachingSampler


def test_cfg_dropout_is_branch_free_and_matches_reference_semantics():
    """gamerft.py:69-91 without the data-dependent branch: enough control-free samples -> untouched; otherwise controls
    are dropped only where present, bringing the control-free share to ~cfg_prob."""
    from owl_audio_exps_b200.models.gamerft import cfg_dropout
    torch.manual_seed(0)
    full = torch.ones(2000, dtype=torch.bool)
    out = cfg_dropout(full, 0.3)
    assert abs((1 - out.float().mean().item()) - 0.3) < 0.04
    half = full.clone()
    half[:1000] = False
    assert torch.equal(cfg_dropout(half, 0.3), half)
    some = full.clone()
    some[:200] = False
    out = cfg_dropout(some, 0.3)
    assert not (out & ~some).any() and abs((1 - out.float().mean().item()) - 0.3) < 0.04
    assert not cfg_dropout(torch.zeros(4, dtype=torch.bool), 0.3).any()
