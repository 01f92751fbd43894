"""Stand-in for the slice of shapely the reference tick path touches.
TEST INFRASTRUCTURE: lets /root/reference's queue_class.py import unmodified in a container without shapely."""
from . import geometry, ops, wkt  # noqa: F401
