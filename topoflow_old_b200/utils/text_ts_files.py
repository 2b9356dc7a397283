"""Multi-column text time series: 15-character columns, float32 values written
with '%15.7f' (topoflow/utils/text_ts_files.py:84-200)."""
import numpy as np

from . import file_utils

_UNIT_ABBREV = {'minutes': 'min', 'seconds': 'sec', 'hours': 'hrs', 'days': 'days',
                'months': 'mon', 'years': 'yrs'}
COL_WIDTH = 15


class ts_file(object):

    def open_new_file(self, file_name, var_names=('X',), dtype='float64', time_units='minutes'):
        self.file_name = file_utils.check_overwrite(file_name)
        self.var_names, self.time_units = list(var_names), time_units
        try:
            self.ts_unit = open(self.file_name, 'w')
        except IOError:
            print('SORRY, Could not open new multi-column text file:')
            print('   ' + self.file_name)
            self.ts_unit = False
            return False
        self.write_header(self.var_names)
        return True

    def open_file(self, file_name):
        try:
            self.ts_unit = open(file_name, 'r')
            return True
        except IOError:
            return False

    def write_header(self, var_names, SILENT=True):
        head = ('Time [%s]' % _UNIT_ABBREV[self.time_units.lower()]).rjust(COL_WIDTH)
        head += ''.join(v.rjust(COL_WIDTH) for v in var_names)
        self.ts_unit.write(head + '\n')
        self.ts_unit.write('-' * ((len(var_names) + 1) * COL_WIDTH) + '\n')

    def add_values(self, time, values):
        row = ('%15.7f' % time).rjust(COL_WIDTH)
        for v in np.ravel(values):
            row += ('%15.7f' % v).rjust(COL_WIDTH)
        self.ts_unit.write(row + '\n')

    def values_at_IDs(self, var, IDs):
        if np.ndim(var) > 0:
            return np.float32(var[IDs])
        return np.zeros(np.size(IDs[0]), dtype='float32') + np.float32(var)

    def add_values_at_IDs(self, time, var, IDs):
        self.add_values(time, self.values_at_IDs(var, IDs))

    def close_file(self):
        self.ts_unit.close()

    close = close_file
