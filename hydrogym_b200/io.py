"""Host-side I/O of the direct-solver surface (reference: hydrogym/jax/utils/io.py; docs kolmogorov.md:56-102):
the text-log callback ``RungeKuttaCrankNicolson.solve`` invokes after a rollout and the on-disk trajectory format
(``.npy`` of shape (T, N, N//2+1) complex, exactly what ``jnp.save`` writes in the reference notebook)."""
from __future__ import annotations

from typing import Callable, Optional

import numpy as np
import torch


class CallbackBase:
    """hydrogym/core.py:245-269: something that happens every ``interval`` iterations."""

    def __init__(self, interval: int = 1):
        self.interval = interval

    def __call__(self, iter: int, t: float, flow) -> bool:  # noqa: A002 (reference argument name)
        return iter % self.interval == 0


def _to_numpy(a):
    return a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)


class LogCallback(CallbackBase):
    """io.py:13-28: ``solve`` calls it once with the flow (whose ``vorticity`` holds the trajectory); the
    post-processed array is written with ``np.savetxt(filename, ..., fmt="%.5f")``."""

    def __init__(self, postprocess: Callable, interval: Optional[int] = 1, filename: Optional[str] = None,
                 print_fmt: Optional[str] = None):
        super().__init__(interval=interval)
        self.filename, self.postprocess, self.print_fmt = filename, postprocess, print_fmt

    def __call__(self, flow):
        values = _to_numpy(self.postprocess(flow))
        if self.print_fmt is not None:
            print(self.print_fmt.format(*np.atleast_1d(values).ravel()[:8]))
        np.savetxt(self.filename, np.atleast_1d(values), fmt="%.5f")
        return values


def save_trajectory(path: str, trajectory) -> None:
    """(T, N, N//2+1) complex -> ``.npy`` (kolmogorov.md:88)."""
    np.save(path, _to_numpy(trajectory))


def load_trajectory(path: str, device=None) -> torch.Tensor:
    """``.npy`` trajectory -> complex tensor (on ``device`` if given)."""
    t = torch.from_numpy(np.load(path))
    return t.to(device) if device is not None else t


def create_animation(trajectory, gif_name, frame_interval_factor):
    """io.py:31-66 renders a GIF with matplotlib; matplotlib is not installed in this image and nothing on the
    hot path needs it, so this only explains itself."""
    raise RuntimeError("create_animation needs matplotlib (host-side plotting, unavailable on this box); "
                       "use save_trajectory() and plot elsewhere")
