"""Minimal planar geometry for the network loader (the reference leans on shapely/geopandas, which
are not a dependency here): WKT LINESTRING parsing, length / interpolate, and the EPSG:4326 ->
EPSG:26910 (NAD83 / UTM zone 10N) projection the loader applies (queue_class.py:39, 64).

Only what the tick path consumes is provided: `Link.__init__` derives in/out angles from points at 1 %
and 99 % of the projected geometry (queue_class.py:423-428)."""
import math
import re


class Point:
    __slots__ = ("x", "y")

    def __init__(self, x, y=None):
        if y is None:
            x, y = x
        self.x = float(x)
        self.y = float(y)

    @property
    def coords(self):
        return [(self.x, self.y)]

    def buffer(self, radius):
        """The disc around the point (the loader splits a link's polyline with `point.buffer(1)`, queue_class.py:84)."""
        return Disc(self, float(radius))

    def __repr__(self):
        return f"POINT ({self.x!r} {self.y!r})"


class Disc:
    __slots__ = ("center", "radius")

    def __init__(self, center, radius):
        self.center, self.radius = center, radius


class LineString:
    __slots__ = ("coords", "_cum")

    def __init__(self, coords):
        self.coords = [(float(p[0]), float(p[1])) for p in coords]
        cum = [0.0]
        for (x0, y0), (x1, y1) in zip(self.coords[:-1], self.coords[1:]):
            cum.append(cum[-1] + math.hypot(x1 - x0, y1 - y0))
        self._cum = cum

    @property
    def length(self):
        return self._cum[-1]

    def interpolate(self, distance, normalized=False):
        total = self._cum[-1]
        d = distance * total if normalized else distance
        if d <= 0.0 or total == 0.0:
            return Point(*self.coords[0])
        if d >= total:
            return Point(*self.coords[-1])
        for i in range(1, len(self._cum)):
            if d <= self._cum[i]:
                seg = self._cum[i] - self._cum[i - 1]
                f = (d - self._cum[i - 1]) / seg if seg > 0 else 0.0
                (x0, y0), (x1, y1) = self.coords[i - 1], self.coords[i]
                return Point(x0 + f * (x1 - x0), y0 + f * (y1 - y0))
        return Point(*self.coords[-1])

    def project(self, point):
        """Arc length of the point of the line closest to `point`."""
        best, best_d = 0.0, float("inf")
        for i in range(1, len(self.coords)):
            (x0, y0), (x1, y1) = self.coords[i - 1], self.coords[i]
            dx, dy = x1 - x0, y1 - y0
            seg2 = dx * dx + dy * dy
            f = 0.0 if seg2 == 0.0 else min(1.0, max(0.0, ((point.x - x0) * dx + (point.y - y0) * dy) / seg2))
            d = math.hypot(point.x - (x0 + f * dx), point.y - (y0 + f * dy))
            if d < best_d:
                best_d, best = d, self._cum[i - 1] + f * (self._cum[i] - self._cum[i - 1])
        return best

    def cut(self, d0, d1):
        """The part of the line between arc lengths d0 <= d1 (end points interpolated, inner vertices kept)."""
        a, b = self.interpolate(d0), self.interpolate(d1)
        inner = [c for c, s in zip(self.coords, self._cum) if d0 < s < d1]
        return LineString([(a.x, a.y)] + inner + [(b.x, b.y)])

    def __repr__(self):
        return "LINESTRING (" + ", ".join(f"{x!r} {y!r}" for x, y in self.coords) + ")"


def split(line, disc):
    """shapely.ops.split(line, point.buffer(r)) for a disc centred ON the line, as the loader's turn-prohibition code
    uses it (queue_class.py:84, 91): the pieces before, inside and after the disc, in line order.  The cuts are made
    at arc length +-r from the centre (shapely intersects with a 64-gon: the same to within a millimetre)."""
    s = line.project(disc.center)
    a, b = s - disc.radius, s + disc.radius
    if not (0.0 < a and b < line.length):
        raise ValueError("the split disc reaches an end of the line (link shorter than ~10 x the disc radius)")
    return [line.cut(0.0, a), line.cut(a, b), line.cut(b, line.length)]


_NUM = re.compile(r"[-+]?(?:\d+\.?\d*|\.\d+)(?:[eE][-+]?\d+)?")


def loads(wkt):
    """WKT reader for the two geometry types the inputs use (LINESTRING, POINT)."""
    head = wkt.lstrip().upper()
    nums = [float(t) for t in _NUM.findall(wkt[wkt.index("("):])]
    pts = list(zip(nums[0::2], nums[1::2]))
    if head.startswith("LINESTRING"):
        return LineString(pts)
    if head.startswith("POINT"):
        return Point(*pts[0])
    raise ValueError("unsupported WKT geometry: " + wkt[:30])


# ---- EPSG:26910: transverse Mercator, GRS80, central meridian -123, k0 0.9996, FE 500 km ----------
_A = 6378137.0
_F = 1.0 / 298.257222101
_K0 = 0.9996
_LON0 = math.radians(-123.0)
_FE = 500000.0
_N = _F / (2.0 - _F)
_AHAT = _A / (1.0 + _N) * (1.0 + _N ** 2 / 4.0 + _N ** 4 / 64.0)
_ALPHA = (
    _N / 2.0 - 2.0 * _N ** 2 / 3.0 + 5.0 * _N ** 3 / 16.0,
    13.0 * _N ** 2 / 48.0 - 3.0 * _N ** 3 / 5.0,
    61.0 * _N ** 3 / 240.0,
)


def utm10n(lon, lat):
    """(lon, lat) degrees -> (easting, northing) metres, Krueger series to n^3 (sub-millimetre)."""
    phi = math.radians(lat)
    lam = math.radians(lon) - _LON0
    e = 2.0 * math.sqrt(_N) / (1.0 + _N)
    t = math.sinh(math.atanh(math.sin(phi)) - e * math.atanh(e * math.sin(phi)))
    xi = math.atan2(t, math.cos(lam))
    eta = math.atanh(math.sin(lam) / math.sqrt(1.0 + t * t))
    x, y = eta, xi
    for j, a in enumerate(_ALPHA, start=1):
        x += a * math.cos(2 * j * xi) * math.sinh(2 * j * eta)
        y += a * math.sin(2 * j * xi) * math.cosh(2 * j * eta)
    return _FE + _K0 * _AHAT * x, _K0 * _AHAT * y


def project(geom):
    if isinstance(geom, Point):
        return Point(*utm10n(geom.x, geom.y))
    return LineString([utm10n(x, y) for x, y in geom.coords])
