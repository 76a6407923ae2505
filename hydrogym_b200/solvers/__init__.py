from .base import PseudoSpectralNavierStokes2D, RungeKuttaCrankNicolson  # noqa: F401
