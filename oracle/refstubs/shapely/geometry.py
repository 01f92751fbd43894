from spatial_queue_b200.geo import LineString, Point  # noqa: F401
