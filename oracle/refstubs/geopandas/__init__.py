"""Stand-in for geopandas.GeoDataFrame(df, crs='epsg:4326', geometry=[...]).to_crs('epsg:26910'), the only
geopandas surface the reference loader uses (queue_class.py:39, 64).  TEST INFRASTRUCTURE."""
import pandas as pd

from spatial_queue_b200.geo import project


class GeoDataFrame:
    def __init__(self, df, crs=None, geometry=None):
        self._df = df.copy()
        self._df["geometry"] = list(geometry)
        self._crs = crs

    def to_crs(self, crs):
        if str(crs).lower() not in ("epsg:26910", "26910"):
            raise NotImplementedError(crs)
        out = pd.DataFrame(self._df)
        out["geometry"] = [project(g) for g in out["geometry"]]
        return out
