"""owl_wms/sampling/schedulers.py:5-13 without the diffusers dependency.

The reference builds `FlowMatchEulerDiscreteScheduler(shift=3, num_train_timesteps=n)` (diffusers, not vendored, not
installed here) and takes dt between consecutive `timesteps / n`.  That scheduler's published construction is
sigma_i = (n - i) / n for i = 0..n-1, shifted to s*sigma / (1 + (s - 1)*sigma), timesteps = sigma * n; restated here in
closed form.  PARITY UNPINNED against diffusers itself (absent from this image); pinned only by its own properties
(tests/test_sampling_host.py: positive steps summing to the first shifted sigma, sigma_0 = 1)."""
import torch


def get_sd3_euler(n_steps: int, shift: float = 3.0) -> torch.Tensor:
    sigma = torch.linspace(1, n_steps, n_steps).flip(0) / n_steps
    ts = shift * sigma / (1 + (shift - 1) * sigma)
    ts = torch.cat([ts, torch.zeros(1, dtype=ts.dtype)])
    return ts[:-1] - ts[1:]
