def split(*a, **k):  # only reached for turn prohibitions (queue_class.py:92,99), unused by every shipped config
    raise NotImplementedError("shapely.ops.split is not stubbed (turn prohibitions are out of scope)")
