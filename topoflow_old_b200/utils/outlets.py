"""Monitored-cell ("outlet") file: 6 header lines, then
column, row, area [km^2], relief [m] (topoflow/utils/outlets.py:45-173)."""
import os

import numpy as np


def read_outlet_file(self):
    """Sets outlet_file, n_outlets, outlet_IDs, outlet_ID, basin_area and
    basin_relief on the BMI object `self`."""
    if hasattr(self, 'pixel_file'):
        outlet_file = self.pixel_file
    else:
        outlet_file = self.case_prefix + '_outlets.txt'
    self.outlet_file = self.in_directory + outlet_file
    if not os.path.exists(self.outlet_file):
        bar = '#' * 70
        print(bar)
        print(' ERROR: Could not find monitored pixel file:')
        print(' ' + self.outlet_file)
        print(bar)
        print(' ')
        raise IOError('Could not find outlet_file.')
    with open(self.outlet_file, 'r') as unit:
        lines = unit.readlines()[6:]
    cols, rows, areas, reliefs = [], [], [], []
    for line in lines:
        words = line.split()
        if len(words) >= 4:
            cols.append(np.int64(np.float64(words[0])))
            rows.append(np.int64(np.float64(words[1])))
            areas.append(np.float64(words[2]))
            reliefs.append(np.float64(words[3]))
    outlet_cols = np.array(cols, dtype='int64')
    outlet_rows = np.array(rows, dtype='int64')
    self.basin_area = np.float64(areas[0])
    self.basin_relief = np.float64(reliefs[0])
    ids = (outlet_rows * self.nx) + outlet_cols
    if np.any((ids < 0) | (ids > (self.rti.n_pixels - 1))):
        print('ERROR: Some outlet_IDs lie outside of DEM.')
        print(' ')
        return
    self.n_outlets = len(cols)
    self.outlet_IDs = (outlet_rows, outlet_cols)
    self.outlet_ID = (outlet_rows[0], outlet_cols[0])
