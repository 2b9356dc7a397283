"""Pixel sizes and areas from the RTI georeferencing
(topoflow/utils/pixels.py:72-349)."""
import numpy as np

_A = np.float64(6378137.0)        # WGS-84 equatorial / polar radii [m]
_B = np.float64(6356752.3)


def _eccentricity(a, b):
    f = (a - b) / a
    return np.sqrt((np.float64(2) * f) - f ** np.float64(2))


def meters_per_degree_lon(lat_deg, a_radius=_A, b_radius=_B, mean_elev=np.float64(0)):
    e = _eccentricity(a_radius, b_radius)
    dtor = np.pi / np.float64(180)
    lat = lat_deg * dtor
    p1 = mean_elev * np.cos(lat)
    p2 = a_radius * np.cos(lat) / np.sqrt(np.float64(1) - (e * np.sin(lat)) ** np.float64(2))
    return dtor * (p1 + p2)


def meters_per_degree_lat(lat_deg, a_radius=_A, b_radius=_B, mean_elev=np.float64(0)):
    e = _eccentricity(a_radius, b_radius)
    dtor = np.pi / np.float64(180)
    lat = lat_deg * dtor
    p = (a_radius * (np.float64(1) - e ** np.float64(2))
         / (np.float64(1) - (e * np.sin(lat)) ** np.float64(2)) ** np.float64(1.5))
    return dtor * (mean_elev + p)


def get_sizes_by_row(rti, REPORT=False, METERS=False):
    """(dx, dy, dd) per row as float64 [nrows]; km unless METERS."""
    ufactor = np.float64(1) if METERS else np.float64(1000)
    if rti.pixel_geom == 0:                          # fixed-angle (lat/lon) pixels
        rows = np.arange(rti.nrows, dtype='int16')
        yresdeg = rti.yres / np.float64(3600)
        lats = rti.y_north_edge - (yresdeg * (rows + np.float64(1)))
        dx = rti.xres / np.float64(3600) * meters_per_degree_lon(lats) / ufactor
        dy = rti.yres / np.float64(3600) * meters_per_degree_lat(lats) / ufactor
        dd = np.sqrt(dx ** 2 + dy ** 2)
    else:
        dx = rti.xres / ufactor
        dy = rti.yres / ufactor
        dd = np.sqrt(dx ** 2 + dy ** 2)
        zeros = np.zeros([rti.nrows], dtype='float64')
        dx, dy, dd = zeros + dx, zeros + dy, zeros + dd
    return (dx, dy, dd)


def get_da(rti, METERS=False, REPORT=False, VERBOSE=False):
    """Cell area [m^2]: a scalar for fixed-length pixels, an [ny, nx] grid for
    fixed-angle pixels (pixels.py:174-240)."""
    dx, dy, dd = get_sizes_by_row(rti, METERS=True)
    if rti.pixel_geom == 1:
        return dx[0] * dy[0]
    return np.reshape(np.repeat(dx * dy, rti.ncols), (rti.nrows, rti.ncols))
