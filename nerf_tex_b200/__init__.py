"""nerf_tex_b200 — B200-native (sm_100a) NeRF-Tex volume-rendering hot path.

The package mirrors the reference's plugin surface (``network.renderer``, ``network.model``,
``network.ray_sampler`` ...) on top of a C-ABI CUDA library; see include/nerftex_b200.h.
"""

__version__ = "0.1.0"
