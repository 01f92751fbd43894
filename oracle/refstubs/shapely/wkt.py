from spatial_queue_b200.geo import loads  # noqa: F401
