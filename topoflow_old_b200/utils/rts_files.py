"""Grid stacks (.rts): grids of one variable appended frame after frame, same
element type and byte order as the data set (topoflow/utils/rts_files.py:
313-396 write side, :236-310 and :397-420 read side)."""
import os

import numpy as np

from . import bov_files, file_utils, rti_files


class rts_file(object):

    def open_file(self, file_name, UPDATE=False):
        info = rti_files.read_info(file_name)
        if info is None:
            return False
        self.info, self.nx, self.ny = info, info.ncols, info.nrows
        self.grid_size = self.nx * self.ny * rti_files.get_bpe(info.data_type)
        self.dtype = rti_files.get_numpy_data_type(info.data_type)
        self.file_name, self.time_index = file_name, 0
        try:
            self.rts_unit = open(file_name, 'rb+' if UPDATE else 'rb')
            return True
        except IOError:
            print('ERROR during rts.open_file().')
            return False

    def open_new_file(self, file_name, info=None, var_name='UNKNOWN', dtype='float32',
                      VERBOSE=False, MAKE_RTI=True, MAKE_BOV=False):
        file_name = file_utils.check_overwrite(file_name)
        self.file_name, self.dtype, self.time_index = file_name, dtype, 0
        if info is None:
            info = rti_files.read_info(rti_files.try_to_find_rti_file(file_name))
        info = info.copy()            # the component's own rti stays as it was read
        info.grid_file = file_name
        info.data_type = rti_files.get_rti_data_type(dtype)
        info.data_source = 'TopoFlow 3.0'
        info.gmin = info.gmax = -9999.0
        self.info, self.nx, self.ny = info, info.ncols, info.nrows
        self.grid_size = self.nx * self.ny * rti_files.get_bpe(info.data_type)
        if MAKE_RTI:
            rti_files.write_info(rti_files.get_file_prefix(file_name) + '.rti', info)
        if MAKE_BOV:
            bov_files.write_info_as_bov(file_name, info, var_name)
        try:
            self.rts_unit = open(file_name, 'wb')
            return True
        except IOError:
            return False

    def add_grid(self, grid, time_index=-1):
        if time_index >= 0:
            self.rts_unit.seek(time_index * self.grid_size)
        out = np.asarray(grid).astype(self.dtype)
        if out.ndim == 0:             # a scalar variable is saved as a constant grid
            out = out + np.zeros([self.ny, self.nx], dtype=self.dtype)
        if self.info.SWAP_ENDIAN:
            out = out.byteswap()
        out.tofile(self.rts_unit)
        self.time_index += 1

    def add_frame(self, frame):
        """Append a frame that already is in the file's dtype and byte order
        (converted on the device, engine.Strip.frame)."""
        assert frame.dtype == np.dtype(self.dtype) and frame.size == self.nx * self.ny
        frame.tofile(self.rts_unit)
        self.time_index += 1

    def get_grid(self, time_index, dtype=None):
        self.rts_unit.seek(time_index * self.grid_size)
        grid = np.fromfile(self.rts_unit, count=self.nx * self.ny, dtype=dtype or self.dtype)
        grid = grid.reshape(self.ny, self.nx)
        if self.info.SWAP_ENDIAN:
            grid.byteswap(True)
        return grid

    def number_of_grids(self):
        return os.path.getsize(self.file_name) // self.grid_size

    def close_file(self):
        self.rts_unit.close()

    close = close_file
