"""Output files of a component: grid stacks (.rts + .rti + .bov) and
multi-column text time series (topoflow/utils/model_output.py:111-430).

The reference also writes netCDF when the third-party `netCDF4` package is
importable and always writes RTS / text as well (MAKE_RTS = True, :206;
MAKE_TTS = True, :343).  Only the always-written formats are produced here.
"""
import numpy as np

from . import file_utils, rts_files, text_ts_files


def check_netcdf():
    """The reference prints the netCDF4 version or a warning; nothing to do."""
    return False


def open_new_gs_file(self, file_name, info=None, var_name='X', long_name='Unknown',
                     units_name='None', dtype='float32', time_units='minutes',
                     nx=None, ny=None, dx=None, dy=None):
    rts_name = file_utils.replace_extension(getattr(self, var_name + '_gs_file'), '.rts')
    unit = rts_files.rts_file()
    ok = unit.open_new_file(rts_name, info if info is not None else self.rti, var_name,
                            dtype=dtype, MAKE_BOV=True)
    if not ok:
        print('ERROR: Unable to open new RTS file:')
        print('      ' + rts_name)
        return
    setattr(self, var_name + '_rts_file', unit.file_name)
    setattr(self, var_name + '_rts_unit', unit)


def add_grid(self, var, var_name, time=None, SILENT=True):
    unit = getattr(self, var_name + '_rts_unit', None)
    if unit is not None:
        unit.add_grid(var)


def close_gs_file(self, var_name):
    unit = getattr(self, var_name + '_rts_unit', None)
    if unit is not None:
        unit.close()


def open_new_ts_file(self, file_name, IDs, var_name='X', long_name='Unknown',
                     units_name='None', dtype='float32', time_units='minutes'):
    rows, cols = IDs[0], IDs[1]
    names = ['%s_%s_%s' % (var_name, rows[k], cols[k]) for k in range(np.size(rows))]
    txt_name = file_utils.replace_extension(getattr(self, var_name + '_ts_file'), '.txt')
    unit = text_ts_files.ts_file()
    if unit.open_new_file(txt_name, names, dtype=dtype, time_units=time_units):
        setattr(self, var_name + '_tts_file', unit.file_name)
        setattr(self, var_name + '_tts_unit', unit)
    else:
        print('ERROR: Unable to open new text file:')
        print('      ' + txt_name)


def add_values_at_IDs(self, time_min, var, var_name, IDs):
    unit = getattr(self, var_name + '_tts_unit', None)
    if unit is not None:
        unit.add_values_at_IDs(time_min, var, IDs)


def close_ts_file(self, var_name):
    unit = getattr(self, var_name + '_tts_unit', None)
    if unit is not None:
        unit.close()
