"""BMI component base: the interface the EMELI framework drives
(topoflow/utils/BMI_base.py:187-2159, restated for Python 3).

Only behaviour is mirrored -- same method names, arguments, status strings,
cfg-file grammar and 0-d "mutable scalar" conventions -- so a component built
on this class can be dropped into the reference's framework.
"""
import os
import time

import numpy as np

from . import outlets, pixels, rti_files


class BMI_component(object):

    _DEFAULT_ATTS = {
        'model_name': 'Model_Name', 'version': '1.0', 'author_name': 'Scott D. Peckham',
        'grid_type': 'uniform', 'time_step_type': 'fixed', 'step_method': 'explicit',
        'comp_name': 'Model_Name', 'model_family': 'TopoFlow',
        'cfg_template_file': 'Model_Name.cfg.in', 'cfg_extension': '_Model_Name.cfg',
        'cmt_var_prefix': '/ModelName/Input/Var/', 'gui_xml_file': '',
        'dialog_title': 'ProcessName: Component Parameters', 'time_units': 'seconds'}

    def __init__(self):                                   # BMI_base.py:189-210
        self.DEBUG = False
        self.SKIP_ERRORS = False
        self.SILENT = True
        self.REPORT = False
        self.DONE = False
        self.status = 'created'
        self.in_directory = None
        self.out_directory = None
        self.site_prefix = None
        self.case_prefix = None
        self.cfg_prefix = None
        self.comp_status = 'Enabled'

    # -- status / attributes ------------------------------------------------
    def get_status(self):
        """created, initializing, initialized, updating, updated, finalizing,
        finalized or failed (:220-230)."""
        return self.status

    def get_attribute(self, att_name):
        atts = getattr(self, '_att_map', self._DEFAULT_ATTS)
        try:
            return atts[att_name.lower()]
        except KeyError:
            print('###################################################')
            print(' ERROR: Could not find attribute: ' + att_name)
            print('###################################################')
            print(' ')

    # -- variable introspection (:456-586) ------------------------------------
    def get_input_var_names(self):
        return getattr(self, '_input_var_names', ['None'])

    def get_output_var_names(self):
        return getattr(self, '_output_var_names', ['None'])

    def get_var_name(self, long_var_name):
        return self._var_name_map[long_var_name]

    def get_var_units(self, long_var_name):
        return self._var_units_map[long_var_name]

    def _var(self, long_var_name):
        return getattr(self, self.get_var_name(long_var_name))

    def get_var_rank(self, long_var_name):
        try:
            return self._var(long_var_name).ndim
        except Exception:
            return 0

    def get_var_type(self, long_var_name):
        try:
            return str(self._var(long_var_name).dtype)
        except Exception:
            return 'unknown'

    def get_var_itemsize(self, long_var_name):
        try:
            return self._var(long_var_name).itemsize
        except Exception:
            return -1

    def get_var_nbytes(self, long_var_name):
        try:
            return self._var(long_var_name).nbytes
        except Exception:
            return -1

    # -- grid description (:278-354) --------------------------------------------
    def get_grid_shape(self, long_var_name):
        try:
            return np.array(self._var(long_var_name).shape)
        except Exception:
            return np.array((1,))

    def get_grid_spacing(self, long_var_name):
        if self.get_var_rank(long_var_name) != 2:
            return None
        if not hasattr(self, 'grid_info'):
            self.read_grid_info()
        return np.array([0, self.grid_info.yres, self.grid_info.xres])

    def get_grid_origin(self, long_var_name):
        if self.get_var_rank(long_var_name) != 2:
            return None
        if not hasattr(self, 'grid_info'):
            self.read_grid_info()
        return np.array([0, self.grid_info.y_south_edge, self.grid_info.x_west_edge])

    def get_grid_attribute(self, long_var_name, att_name):
        amap = {'x_units': 'meters', 'y_units': 'meters', 'z_units': 'meters',
                'ellipsoid': 'None', 'datum': 'None', 'projection': 'None'}
        return amap.get(att_name.lower())

    # -- values: references, never copies (:590-692) ----------------------------
    def get_values(self, long_var_name):
        var_name = self.get_var_name(long_var_name)
        try:
            return getattr(self, var_name)
        except AttributeError:
            print('ERROR in BMI_base.get_values()')
            print('    for var_name =', var_name)
            print('    Returning 0.')
            value = np.array(0, dtype='float64')
            setattr(self, var_name, value)
            return value

    def set_values(self, long_var_name, value):
        setattr(self, self.get_var_name(long_var_name), value)

    def get_values_at_indices(self, long_var_name, IDs):
        return np.array(self._var(long_var_name).flat[IDs], dtype='float64')

    def set_values_at_indices(self, long_var_name, IDs, values):
        self._var(long_var_name).flat[IDs] = values

    # -- time (:698-765) -----------------------------------------------------
    def get_time_step(self):
        return np.float64(self.dt)

    def get_time_units(self):
        return self.get_attribute('time_units')

    def get_start_time(self):
        return np.float64(0)

    def get_current_time(self):
        return self.time

    def get_end_time(self):
        if hasattr(self, 'stop_time'):
            return np.float64(self.stop_time)
        computable = (self.get_attribute('time_step_type') == 'fixed') and hasattr(self, 'n_steps')
        if getattr(self, 'stop_code', 0) != 0:
            computable = False
        if computable:
            return np.float64(self.n_steps * self.dt)
        print(' ERROR: Unable to compute model stop_time.')
        return np.float64(-1)

    def initialize_time_vars(self, units='seconds'):          # :1007-1063
        self.start_time = time.time()
        self.time_units = units.lower()
        self.time_index = np.int32(0)
        self.time = self.initialize_scalar(0, dtype='float64')
        self.DONE = False
        if not hasattr(self, 'stop_code'):
            self.stop_code = 0
        if not hasattr(self, 'n_steps'):
            self.n_steps = 1
        self.sec_per_year = np.float64(365) * 24 * 3600
        self.min_per_year = np.float64(365) * 24 * 60
        self.time_sec = self.initialize_scalar(0, dtype='float64')
        self.time_min = self.initialize_scalar(0, dtype='float64')
        self.last_print_time = time.time() - 100.0

    _SEC_PER = {'seconds': 1.0, 'minutes': 60.0, 'hours': 3600.0, 'days': 86400.0}

    def update_time(self, dt=-1):                             # :1067-1114
        self.time_index += 1
        if dt == -1:
            self.time += self.dt
        else:
            self.time += dt
        if self.time_units == 'seconds':
            self.time_sec = self.time
        elif self.time_units == 'years':
            self.time_sec = self.time * self.sec_per_year
        else:
            self.time_sec = self.time * np.float64(self._SEC_PER[self.time_units])
        self.time_min = self.time_sec / np.float64(60)
        self.check_finished()

    def check_finished(self):                                 # :968-1003
        if self.mode != 'driver':
            return
        if self.stop_code == 0:
            times_up = (self.time_index >= self.n_steps)
        elif self.stop_code == 1:
            times_up = (self.time >= self.stop_time)
        else:
            times_up = (self.dz_max < self.dz_tolerance)
        self.DONE = (self.DONE or times_up)

    def print_time_and_value(self, var, var_name='Q_out', units_name='[m^3/s]',
                             interval=2.0, PRINT_INDEX=False):   # :1118-1157
        if self.time_index == 0:
            print('Will print values every', interval, 'seconds.')
        if (time.time() - self.last_print_time) > interval:
            if self.time_units == 'seconds':
                cur, ustr = self.time_min, ' [min]'
            else:
                cur, ustr = self.time, ' [' + self.time_units + ']'
            print('Time = ' + ('%10.2f' % cur) + ustr + ',  ' + var_name + ' = '
                  + ('%10.5f' % var) + ' ' + units_name)
            if PRINT_INDEX:
                print('n =', self.time_index + 1, 'of', self.n_steps)
            self.last_print_time = time.time()

    def get_run_time_string(self, proc_name='component', sec_digits=4, seconds=None, SILENT=None):
        if seconds is None:
            seconds = time.time() - self.start_time
        whole = int(seconds)
        frac = ('%.*f' % (sec_digits, seconds - whole))[1:]
        days, rem = divmod(whole, 86400)
        hours, rem = divmod(rem, 3600)
        minutes, secs = divmod(rem, 60)
        out = ''
        for n, word in ((days, 'day'), (hours, 'hour'), (minutes, 'minute')):
            if n > 0:
                out += '%d %s%s, ' % (n, word, 's' if n > 1 else '')
        return out + str(secs) + frac + ' seconds.'

    def print_final_report(self, comp_name='BMI component', mode='nondriver'):   # :1261-1311
        if mode == 'nondriver':
            print(comp_name + ': Finished.')
            return
        units = ' [' + self.time_units + ']'
        print(''.ljust(60, '-'))
        print(comp_name)
        print(time.asctime())
        print(' ')
        print('Input directory:      ' + self.in_directory)
        print('Output directory:     ' + self.out_directory)
        print('Site prefix:          ' + self.site_prefix)
        print('Case prefix:          ' + self.case_prefix)
        print(' ')
        print('Simulated time:      ' + str(self.time) + units)
        print('Program run time:    ' + self.get_run_time_string())
        print(' ')
        print('Number of timesteps: ' + str(self.time_index))
        print('Process timestep:    ' + str(self.dt) + units)
        print('Number of columns:   ' + str(self.nx))
        print('Number of rows:      ' + str(self.ny))
        print(' ')
        print('Finished. (' + self.case_prefix + ')')
        print(' ')

    # -- generic lifecycle (:814-966) -------------------------------------------
    def run_model(self, cfg_directory=None, cfg_prefix=None, n_steps=5):
        if cfg_directory is not None:
            os.chdir(cfg_directory)
        self.cfg_prefix = cfg_prefix
        cfg_file = None
        if cfg_prefix is not None:
            cfg_file = os.path.join(os.getcwd(), cfg_prefix + self.get_attribute('cfg_extension'))
        self.initialize(cfg_file=cfg_file, mode='driver')
        if not hasattr(self, 'stop_code'):
            self.stop_code = 0
        while not self.DONE:
            self.update()
            self.check_finished()
        self.finalize()

    def finalize(self):
        self.status = 'finalizing'
        self.close_input_files()
        self.close_output_files()
        self.status = 'finalized'

    def open_input_files(self):
        pass

    def read_input_files(self):
        pass

    def close_input_files(self):
        pass

    def open_output_files(self):
        pass

    def close_output_files(self):
        pass

    def set_computed_input_vars(self):
        pass

    def initialize_layer_vars(self):
        pass

    # -- grid info, basins (:373-441, :1851-1859) ----------------------------------
    def read_grid_info(self):
        self.grid_info_file = self.in_directory + self.site_prefix + '.rti'
        info = rti_files.read_info(self.grid_info_file)
        if info is None:
            self.grid_info_file = self.out_directory + self.case_prefix + '.rti'
            info = rti_files.read_info(self.grid_info_file)
        self.rti = info
        self.nx = info.ncols
        self.ny = info.nrows
        self.grid_info = self.rti
        self.grid_info.nx = info.ncols
        self.grid_info.ny = info.nrows
        self.da = pixels.get_da(info)
        if info.pixel_geom == 1:
            self.dx, self.dy = info.xres, info.yres
        else:
            self.dx, self.dy = info.xres / 3600, info.yres / 3600

    def initialize_basin_vars(self):
        outlets.read_outlet_file(self)

    # -- cfg files (:1324-1693) ---------------------------------------------------
    @staticmethod
    def _cfg_records(path):
        """(name, value, type) of every 'name | value | type | help' line."""
        with open(path, 'r') as unit:
            for line in unit:
                if line[:1] == '#':
                    continue
                words = line.split('|')
                if len(words) == 4:
                    yield words[0].strip(), words[1].strip(), words[2].strip()

    def read_path_info(self):
        ext = self.get_attribute('cfg_extension')
        path = self.cfg_file.replace(ext, '') + '_path_info.cfg'
        if not os.path.exists(path):
            print('WARNING: file info CFG file not found:')
            print('         ' + path)
            return
        for name, value, _ in self._cfg_records(path):
            setattr(self, name, value)

    def _get_tf(self, name, index):
        var = getattr(self, name)
        return var[index] if index != -1 else var

    def _set_tf(self, name, value, index):
        if index != -1:
            getattr(self, name)[index] = value
        else:
            setattr(self, name, value)

    def read_config_file(self):
        """name | value | type | help.  'X_type' on the line before 'X' decides
        whether X is a Scalar (value kept) or comes from a file (value is the
        file name, stored as X_file; X starts as a 0-d placeholder).  double /
        float -> 0-d float64, long / int -> 0-d int32, SAVE_* Yes/No -> bool,
        [case_prefix] / [site_prefix] are expanded (:1439-1693)."""
        if not os.path.exists(self.cfg_file):
            print('WARNING: cfg_file not found:')
            print('         ' + self.cfg_file)
            return
        last = ''
        for var_name, value, var_type in self._cfg_records(self.cfg_file):
            if var_name == 'n_layers':
                self.n_layers = np.int32(value)
                self.initialize_layer_vars()
            p1, p2 = var_name.rfind('['), var_name.rfind(']')
            if p1 > 0 and p2 > p1:
                base, index = var_name[:p1], np.int32(var_name[p1 + 1:p2])
            else:
                base, index = var_name, -1
            read_filename = False
            if last.startswith(base + '_type'):
                choice = self._get_tf(base + '_type', index).lower()
                if choice == 'scalar':
                    self._set_tf(base + '_file', '', index)
                else:
                    self._set_tf(base, self.initialize_scalar(0), index)
                    read_filename = True
            if var_type in ('double', 'float'):
                self._set_tf(base, self.initialize_scalar(np.float64(value)), index)
            elif var_type in ('long', 'int'):
                self._set_tf(base, self.initialize_scalar(np.int32(value), dtype='int32'), index)
            elif var_type == 'string':
                if value[:13] == '[case_prefix]':
                    text = self.case_prefix + value[13:]
                elif value[:13] == '[site_prefix]':
                    text = self.site_prefix + value[13:]
                else:
                    text = value
                if var_name[:5] == 'SAVE_':
                    setattr(self, var_name, value.lower() == 'yes')
                elif read_filename:
                    self._set_tf(base + '_file', text, index)
                else:
                    self._set_tf(base, text, index)
            else:
                print('ERROR in BMI_base.read_config_file().')
                print('   Unsupported data type = ' + var_type + '.')
            last = var_name

    def initialize_config_vars(self):                          # :1697-1823
        if self.cfg_prefix is None:
            self.cfg_prefix = self.case_prefix
        self.read_path_info()
        self.read_config_file()
        self.check_directories()
        self.set_computed_input_vars()

    def check_directories(self):                               # :1862-1937
        if self.in_directory is None:
            self.in_directory = os.getcwd() + os.sep
        if self.out_directory is None:
            self.out_directory = os.getcwd() + os.sep
        if self.site_prefix is None:
            self.site_prefix = self.cfg_prefix
        if self.case_prefix is None:
            self.case_prefix = self.cfg_prefix
        if self.cfg_file is not None:
            self.cfg_directory = os.path.dirname(self.cfg_file) + os.sep
            if self.in_directory[0] == '.':
                self.in_directory = self.cfg_directory
            if self.out_directory[0] == '.':
                self.out_directory = self.cfg_directory
        self.in_directory = os.path.expanduser(self.in_directory)
        self.out_directory = os.path.expanduser(self.out_directory)
        if self.in_directory[-1] != os.sep:
            self.in_directory += os.sep
        if self.out_directory[-1] != os.sep:
            self.out_directory += os.sep

    def prepend_directory(self, file_list, INPUT=True):
        prefix = self.in_directory if INPUT else self.out_directory
        for name in file_list:
            value = getattr(self, name, '')
            if value != '':
                setattr(self, name, prefix + value)

    # -- mutable scalars and grids (:1941-2082) ---------------------------------
    def initialize_scalar(self, value=0.0, dtype='float64'):
        """0-d ndarray: other components keep a reference and see .fill() /
        in-place updates."""
        return np.array(value, dtype=dtype)

    def initialize_grid(self, value=0.0, dtype='float64'):
        grid = np.zeros([self.ny, self.nx], dtype=dtype)
        if value != 0.0:
            grid += value
        return grid

    def initialize_var(self, var_type, value=0.0, dtype='float64'):
        if var_type.lower() == 'time_series':
            return self.initialize_scalar(value=value, dtype=dtype)
        return self.initialize_grid(value=value, dtype=dtype)

    def update_var(self, var_name, value):
        if value is None:
            return
        var = getattr(self, var_name)
        if self.is_scalar(var_name):
            var[()] = value
        else:
            var[:] = value

    def _ndim(self, var_name):
        try:
            return getattr(self, var_name).ndim
        except AttributeError:
            return 0

    def is_scalar(self, var_name):
        return self._ndim(var_name) == 0

    def is_vector(self, var_name):
        return self._ndim(var_name) == 1

    def is_grid(self, var_name):
        return self._ndim(var_name) == 2
