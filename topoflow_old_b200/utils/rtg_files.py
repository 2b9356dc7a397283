"""Flat binary grids (.rtg): row-major, no header, dtype named by RTG_type,
byte order from the RTI (topoflow/utils/rtg_files.py:427-507)."""
import numpy as np

from . import rti_files


def read_grid(RTG_file, rti, RTG_type='FLOAT', REPORT=False, SILENT=True):
    dtype = rti_files.get_numpy_data_type(RTG_type)
    with open(RTG_file, 'rb') as unit:
        grid = np.fromfile(unit, count=rti.n_pixels, dtype=dtype)
    grid = grid.reshape(rti.nrows, rti.ncols)
    if rti.SWAP_ENDIAN:
        grid.byteswap(True)
    if REPORT:
        print('    min(grid) =', grid.min())
        print('    max(grid) =', grid.max())
    return grid


def write_grid(grid, RTG_file, rti, RTG_type='FLOAT', REPORT=False, SILENT=True):
    if RTG_type == 'BYTE':
        grid = np.uint8(np.int16(grid))
    elif RTG_type in rti_files.RTI_TO_NUMPY:
        grid = np.asarray(grid).astype(rti_files.get_numpy_data_type(RTG_type))
    else:
        raise RuntimeError('No match found for expression')
    if rti.SWAP_ENDIAN:
        grid = grid.byteswap()
    with open(RTG_file, 'wb') as unit:
        grid.tofile(unit)
