"""Import-name shim: the reference does
``from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer``
(/root/reference/gaussian_renderer/__init__.py:15).  Putting this repository on PYTHONPATH makes
that line resolve to the B200 implementation."""
from langsplatv2_b200.rasterizer import GaussianRasterizationSettings, GaussianRasterizer  # noqa: F401
