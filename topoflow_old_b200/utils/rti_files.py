"""RiverTools Info (.rti) files: the text header that describes every binary
grid of a data set.  Same on-disk format and field order as the reference
(topoflow/utils/rti_files.py:267-382 read_info, :505-614 write_info).
"""
import glob
import os
import sys

import numpy as np

RTI_TO_NUMPY = {'BYTE': 'uint8', 'INTEGER': 'int16', 'LONG': 'int32',
                'LONG64': 'int64', 'FLOAT': 'float32', 'DOUBLE': 'float64'}
NUMPY_TO_RTI = dict((v, k) for k, v in RTI_TO_NUMPY.items())
BYTES_PER_ELEMENT = {'BYTE': 1, 'INTEGER': 2, 'LONG': 4, 'LONG64': 8, 'FLOAT': 4, 'DOUBLE': 8}

# (attribute, reader type, upper-case?) in file order (rti_files.py:311-330)
_FIELDS = (('grid_file', 'STRING', False), ('data_source', 'STRING', False),
           ('ncols', 'LONG', False), ('nrows', 'LONG', False),
           ('data_type', 'STRING', True), ('byte_order', 'STRING', True),
           ('pixel_geom', 'BYTE', False), ('xres', 'DOUBLE', False),
           ('yres', 'DOUBLE', False), ('zres', 'FLOAT', False),
           ('z_units', 'STRING', True), ('y_south_edge', 'DOUBLE', False),
           ('y_north_edge', 'DOUBLE', False), ('x_east_edge', 'DOUBLE', False),
           ('x_west_edge', 'DOUBLE', False), ('box_units', 'STRING', True),
           ('gmin', 'FLOAT', False), ('gmax', 'FLOAT', False),
           ('UTM_zone', 'STRING', True))

_CASTS = {'BYTE': np.uint8, 'INTEGER': np.int16, 'FLOAT': np.float32,
          'DOUBLE': np.float64, 'LONG': np.int32, 'LONG64': np.int64}


class GridInfo(object):
    """Attribute bag returned by read_info / make_info."""

    def __init__(self, **kw):
        self.__dict__.update(kw)

    def copy(self):
        return GridInfo(**self.__dict__)


def get_file_prefix(file_name):
    return file_name[:file_name.rfind('.')]


def get_rti_file_name(file_name):
    if file_name[-4:].lower() == '.rti':
        return file_name
    return get_file_prefix(file_name) + '.rti'


def exists(rti_file, SILENT=True):
    return len(glob.glob(rti_file)) > 0


def try_to_find_rti_file(file_name, SILENT=False):
    """<prefix>.rti, else drop '_suffix' parts of the prefix one at a time
    (rti_files.py:87-126)."""
    rti_file = get_rti_file_name(file_name)
    if exists(rti_file):
        return rti_file
    prefix = get_file_prefix(file_name)
    while True:
        pos = prefix.rfind('_')
        if pos == -1:
            break
        prefix = prefix[:pos]
        if exists(prefix + '.rti'):
            return prefix + '.rti'
    if not SILENT:
        print('ERROR: Could not find RTI file for:')
        print('      ' + file_name)
    return None


def get_rti_data_type(dtype):
    return NUMPY_TO_RTI[str(dtype).lower()]


def get_numpy_data_type(data_type):
    return RTI_TO_NUMPY[data_type.upper()]


def get_rti_byte_order(python_byte_order=sys.byteorder):
    return {'big': 'MSB', 'little': 'LSB'}[python_byte_order]


def get_bpe(data_type):
    return BYTES_PER_ELEMENT[data_type.upper()]


def byte_swap_needed(file_name, info=None):
    if info is None:
        info = read_info(get_rti_file_name(file_name))
    return get_rti_byte_order() != info.byte_order


def _next_value(unit, kind, upcase):
    """Value after the first ':' of the next line that is not a ';' comment
    (rti_files.py:210-263)."""
    text = ''
    while True:
        line = unit.readline()
        if line == '':
            break
        if line[0] != ';' and ':' in line:
            text = line[line.find(':') + 1:].strip()
            break
    if kind == 'STRING':
        return text.upper() if upcase else text
    return _CASTS[kind](float(text) if kind in ('FLOAT', 'DOUBLE') else int(float(text)))


def read_info(file_name, SILENT=False, REPORT=False):
    rti_file = try_to_find_rti_file(file_name, SILENT=SILENT)
    if rti_file is None:
        return None
    with open(rti_file, 'r') as unit:
        if unit.readline().upper().strip() != 'RIVERTOOLS INFO FILE':
            if not SILENT:
                print('ERROR:  Invalid RTI file: ' + rti_file)
            return None
        info = GridInfo()
        for name, kind, upcase in _FIELDS:
            setattr(info, name, _next_value(unit, kind, upcase))
    if info.UTM_zone == '-1':
        info.UTM_zone = 'unknown'
    info.n_pixels = info.ncols * info.nrows
    info.bpe = get_bpe(info.data_type)
    info.grid_size = info.bpe * info.n_pixels
    info.SWAP_ENDIAN = byte_swap_needed(rti_file, info)
    if REPORT:
        for name, _, _ in _FIELDS:
            print('%-12s = %s' % (name, getattr(info, name)))
    return info


def make_info(grid_file=None, ncols=None, nrows=None, xres=None, yres=None,
              data_source='Unknown', data_type='FLOAT', byte_order=None,
              pixel_geom=1, zres=1.0, z_units='METERS', y_south_edge=0.0,
              y_north_edge=None, x_west_edge=0.0, x_east_edge=None,
              box_units='METERS', gmin=-9999.0, gmax=-9999.0,
              UTM_zone='UNKNOWN', dtype=None, SILENT=True):
    if byte_order is None:
        byte_order = get_rti_byte_order()
    if dtype is not None:
        data_type = get_rti_data_type(dtype)
    if x_east_edge is None:
        dx = xres if pixel_geom == 1 else xres / np.float64(3600)
        x_east_edge = x_west_edge + ncols * dx
    if y_north_edge is None:
        dy = yres if pixel_geom == 1 else yres / np.float64(3600)
        y_north_edge = y_south_edge + nrows * dy
    bpe = get_bpe(data_type)
    return GridInfo(grid_file=grid_file, data_source=data_source, ncols=ncols, nrows=nrows,
                    data_type=data_type, byte_order=byte_order, pixel_geom=pixel_geom,
                    xres=xres, yres=yres, zres=zres, z_units=z_units,
                    y_south_edge=y_south_edge, y_north_edge=y_north_edge,
                    x_west_edge=x_west_edge, x_east_edge=x_east_edge, box_units=box_units,
                    gmin=gmin, gmax=gmax, UTM_zone=UTM_zone, n_pixels=ncols * nrows, bpe=bpe,
                    grid_size=bpe * ncols * nrows,
                    SWAP_ENDIAN=(get_rti_byte_order() != byte_order))


_TRAILER = """;----------------------------------
;Notes About RiverTools Info Files
;----------------------------------
;(1)  RiverTools grid info filenames end with '.rti'.
;(2)  The first line should be:  RiverTools Info File
;(3)  Lines starting with a semi-colon are ignored.
;(4)  Colons are used to delimit labels from values.
;(5)  The order of the numbers above is important.
;(6)  Number of rows and columns are required.
;(7)  Pixel geometry codes are: 0=Fixed-Angle, 1=Fixed-Length.
;(8)  Pixel x-resolution and y-resolution are required.
;(9)  Measurement units must be METERS or DEGREES.
;(10) Elevation data type is required.
;     Allowed types are BYTE, INTEGER, LONG, LONG64, FLOAT, DOUBLE.
;(11) Byte order is required; either LSB or MSB.
;     (LSB = Least Significant Byte = little-endian)
;     (MSB = Most Significant Byte  = big-endian)
;(12) For 'fixed-angle' pixels, bounding lats and lons
;     are required to compute lengths and areas correctly.
;(13) Latitudes south of equator are negative and
;     longitudes west of prime meridian are negative.
;(14) This file is best modified with the View DEM Info
;     dialog in the File menu.
"""


def write_info(file_name, info, SILENT=True):
    """Same labels, order and '%22.12f' number format as rti_files.py:505-614,
    so the reference (and RiverTools) can read it back."""
    f = lambda v: ('%22.12f' % v).strip()
    sections = (
        ('Description of Grid', (('Grid filename:  ', info.grid_file),
                                 ('Grid source:    ', info.data_source))),
        ('Grid dimensions, etc.', (('Number of columns:  ', str(info.ncols)),
                                   ('Number of rows:     ', str(info.nrows)),
                                   ('Data type:          ', info.data_type),
                                   ('Byte order:         ', info.byte_order))),
        ('Pixel geometry info', (('Pixel geometry code:  ', str(np.int16(info.pixel_geom))),
                                 ('Pixel x-resolution:   ', f(info.xres)),
                                 ('Pixel y-resolution:   ', f(info.yres)),
                                 ('Pixel z-resolution:   ', f(info.zres)),
                                 ('Z-resolution units:   ', info.z_units))),
        ('Bounding box coordinates (degrees lat/lon or UTM)',
         (('South edge y-coordinate:  ', f(info.y_south_edge)),
          ('North edge y-coordinate:  ', f(info.y_north_edge)),
          ('East  edge x-coordinate:  ', f(info.x_east_edge)),
          ('West  edge x-coordinate:  ', f(info.x_west_edge)),
          ('Measurement units:        ', info.box_units))),
        ('Min and max value', (('Minimum value:  ', f(info.gmin)),
                               ('Maximum value:  ', f(info.gmax)))),
        ('UTM zone', (('UTM zone: ', str(info.UTM_zone)),)))
    with open(get_rti_file_name(file_name), 'w') as out:
        out.write('RiverTools Info File\n\n')
        for title, rows in sections:
            bar = ';' + '-' * (len(title) + 1)
            out.write(bar + '\n;' + title + '\n' + bar + '\n')
            for label, value in rows:
                out.write(label + value + '\n')
            out.write('\n')
        out.write('\n' + _TRAILER + '\n\n')
