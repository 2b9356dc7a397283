"""Input "data streams" of a component: a scalar from the cfg file, a text
time series, or float32 grids in a flat binary file
(topoflow/utils/model_input.py:21-240)."""
import os

import numpy as np


def open_file(var_type, input_file):
    if var_type.lower() == 'scalar':
        return None
    if input_file == '':
        print('ERROR in model_input.open_file():')
        print('    Input file is null string.')
        return None
    if not os.path.exists(input_file):
        print('ERROR in model_input.open_file():')
        print('    Could not find input file =')
        print('    ' + input_file)
        return None
    return open(input_file, 'r' if var_type.lower() == 'time_series' else 'rb')


def read_scalar(file_unit, dtype='float32'):
    line = file_unit.readline().strip()
    if line == '':
        return None
    return np.float32(float(line.split()[0]))


def read_grid(file_unit, rti, dtype='float32'):
    if file_unit.tell() == os.path.getsize(file_unit.name):
        return None                                   # end of file: keep the last grid
    grid = np.fromfile(file_unit, count=rti.n_pixels, dtype=str(dtype).lower())
    grid = np.reshape(grid, (rti.nrows, rti.ncols))
    if rti.SWAP_ENDIAN:
        grid.byteswap(True)
    return grid


def read_next(file_unit, var_type, rti, dtype='float32', factor=1.0):
    """Next value of an input variable, widened to float64; None for a Scalar
    (already set from the cfg) or at end of file (model_input.py:127-172)."""
    kind = var_type.lower()
    if kind == 'scalar':
        data = None
    elif kind == 'time_series':
        data = read_scalar(file_unit, dtype)
    elif kind in ('grid', 'grid_sequence'):
        data = read_grid(file_unit, rti, dtype)
    else:
        raise RuntimeError('No match found for ' + var_type + '.')
    if data is None:
        return None
    if factor != 1:
        # "data *= factor" under NumPy 1.x casting: a float32 grid stays float32
        # (the factor is cast down), a float32 scalar times a float64 widens
        data = data * np.float32(factor) if np.ndim(data) else np.float64(data) * factor
    return np.float64(data)


def close_file(file_unit):
    file_unit.close()
