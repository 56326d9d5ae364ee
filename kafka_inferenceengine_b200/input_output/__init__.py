"""The pieces of ``kafka.input_output`` that sit next to the assimilation path and
need no GDAL: the spatial tiling helper, the raster output stage and the markers
for on-device observation ingest.  Readers and GeoTIFF writers stay with the
reference (out of scope)."""
from ..engine import BackscatterSigma0, BHRKernelWeights, ScaledReflectance, pinned_raster      # noqa: F401
from ..output import RasterOutput                             # noqa: F401
from .utils import get_chunks                                 # noqa: F401
