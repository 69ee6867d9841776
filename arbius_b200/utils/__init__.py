"""Cross-cutting helpers: stage timing on the device (max over ranks), latency histograms."""
from .timing import LatencyHistogram, StageTimer  # noqa: F401
