"""shapely.ops.split as the reference's loader uses it for turn prohibitions (queue_class.py:84, 91): a link's polyline
split by `point.buffer(1)`.  TEST INFRASTRUCTURE: delegates to the planar helper of spatial_queue_b200.geo (cuts at arc
length +-1 m around the point instead of intersecting a 64-gon; the same to within a millimetre)."""
from spatial_queue_b200.geo import split  # noqa: F401
