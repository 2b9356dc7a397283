"""D8 component, grid-independent part: configuration, static tables, pixel
geometry, DEM input and output files.  Mirrors the interface of
topoflow/components/d8_base.py (class d8_component, :108-1235); the flow-grid
algorithms themselves live in d8_global.py and run on the GPU.
"""
import os

import numpy as np

from ..utils import BMI_base, model_output, pixels, rtg_files

_OUTPUTS = (('code', 'D8 flow direction codes', 'none', 'uint8', 'd8_grid'),
            ('area', 'D8 contributing area', 'km^2', 'float32', 'A'),
            ('ds', 'flow length in D8 direction', 'm', 'float32', 'ds'),
            ('dw', 'flow width orthogonal to D8 direction', 'm', 'float32', 'dw'))


class d8_component(BMI_base.BMI_component):

    _input_var_names = ['land_surface__elevation']            # (DEM)
    _output_var_names = []
    _var_name_map = {'land_surface__elevation': 'DEM'}
    _var_units_map = {'land_surface__elevation': 'm'}

    # -- configuration ----------------------------------------------------------
    def set_constants(self):                                  # d8_base.py:178-192
        self.dt = 1.0
        self.nodata = np.float32(-9999)
        self.BREAK_TIES = True

    def set_default_config_vars(self):                        # :196-249
        """Used when the D8 component is embedded and has no cfg file of its
        own; values already set on the object win."""
        keep = ('in_directory', 'site_prefix', 'DEM_file', 'A_units', 'FILL_PITS_IN_Z0')
        defaults = dict(
            comp_status='Enabled', in_directory='~/CMT_Output/', out_directory='~/CMT_Output/',
            site_prefix='NOT_SET', case_prefix='NOT_SET', n_steps=1, dt=1.0,
            DEM_file='NOT_SET', A_units='km^2', LINK_FLATS=1, FILL_PITS_IN_Z0=0,
            LR_PERIODIC=0, TB_PERIODIC=0, save_grid_dt=1.0, save_pixels_dt=1.0,
            pixel_file='NOT_SET')
        for stem in ('AREA', 'CODE', 'DS', 'DW'):
            defaults['SAVE_%s_GRIDS' % stem] = False
            defaults['SAVE_%s_PIXELS' % stem] = False
            defaults['%s_gs_file' % stem.lower()] = ''
            defaults['%s_ts_file' % stem.lower()] = ''
        for name, value in defaults.items():
            if name in keep and getattr(self, name, None) is not None:
                continue
            if name == 'case_prefix' and getattr(self, 'case_prefix', None) is not None:
                continue       # the embedding component sets it before initialize()
            setattr(self, name, value)

    def set_computed_input_vars(self):                        # :481-501
        self.LR_PERIODIC = (self.LR_PERIODIC != 0)
        self.TB_PERIODIC = (self.TB_PERIODIC != 0)
        self.ALL_PERIODIC = (self.LR_PERIODIC and self.TB_PERIODIC)

    # -- lifecycle ----------------------------------------------------------------
    def initialize(self, cfg_file=None, mode='nondriver', SILENT=False, REPORT=True):
        """:253-349"""
        if not SILENT:
            print('D8 component: Initializing...')
        self.status = 'initializing'
        self.mode = mode
        if cfg_file is None:
            cfg_file = self.in_directory + self.case_prefix + self.get_attribute('cfg_extension')
        self.cfg_file = cfg_file
        self.set_constants()
        self.set_default_config_vars()
        self.initialize_config_vars()
        self.read_grid_info()
        self.initialize_basin_vars()
        self.initialize_time_vars()
        self.get_pixel_dimensions(SILENT=SILENT, REPORT=REPORT)
        self.get_flow_code_list()
        self.get_flow_code_list_opps()
        self.get_ID_grid(SILENT=SILENT)
        self.get_parent_inc_map()
        self.get_edge_IDs(SILENT=SILENT)
        self.get_not_edge_grid(SILENT=SILENT)
        self.get_resolve_array(SILENT=SILENT)
        self.get_valid_code_map(SILENT=SILENT)
        self.initialize_computed_vars(SILENT=SILENT, REPORT=REPORT)
        self.open_output_files()
        self.status = 'initialized'

    def _update(self, with_area, time, DEM, SILENT, REPORT):
        self.status = 'updating'
        if not SILENT:
            print('D8 component: Updating...')
        self.update_flow_grid(DEM=DEM, SILENT=SILENT, REPORT=REPORT)
        self.update_parent_ID_grid()
        self.update_parent_IDs()
        self.update_flow_from_IDs()
        self.update_flow_to_IDs()
        self.update_flow_width_grid(SILENT=SILENT, REPORT=REPORT)
        self.update_flow_length_grid(SILENT=SILENT, REPORT=REPORT)
        if with_area:
            self.update_area_grid(SILENT=SILENT, REPORT=REPORT)
        self.write_output_files(time)
        self.update_time()
        self.status = 'updated'

    def update(self, time=None, DEM=None, SILENT=True, REPORT=False):
        """Everything including the contributing area (:351-411)."""
        self._update(True, time, DEM, SILENT, REPORT)

    def update_most(self, time=None, DEM=None, SILENT=True, REPORT=False):
        """All but the area grid, which channels do not need (:413-479)."""
        self._update(False, time, DEM, SILENT, REPORT)

    # -- DEM and pixel geometry -----------------------------------------------------
    def initialize_computed_vars(self, DOUBLE=False, SILENT=True, REPORT=False):
        """:505-555.  dw, ds and A are float32 grids; the DEM is read as float32."""
        self.RT3_TEST = False
        dtype = 'float64' if DOUBLE else 'float32'
        self.dw = np.zeros([self.ny, self.nx], dtype=dtype)
        self.ds = np.zeros([self.ny, self.nx], dtype=dtype)
        self.A = np.zeros([self.ny, self.nx], dtype=dtype)
        self.DEM_file = self.in_directory + self.DEM_file
        if not os.path.exists(self.DEM_file):
            print('ERROR: Could not find DEM file:')
            print(self.DEM_file)
            return
        DEM = rtg_files.read_grid(self.DEM_file, self.rti, SILENT=False)
        print('  min(DEM), max(DEM) =', DEM.min(), ',', DEM.max(), '[m]')
        self.DEM = DEM
        if self.FILL_PITS_IN_Z0:
            raise NotImplementedError(
                'FILL_PITS_IN_Z0 = 1 is not supported: the reference fills pits with a serial '
                'priority flood (utils/fill_pits.py) and overwrites the DEM file; fill the DEM '
                'beforehand and set FILL_PITS_IN_Z0 to 0 (as all shipped cfg files do).')

    def get_pixel_dimensions(self, DOUBLE=False, SILENT=True, REPORT=False):   # :559-612
        dx, dy, dd = pixels.get_sizes_by_row(self.rti, METERS=True)
        if not DOUBLE:
            dx, dy, dd = np.float32(dx), np.float32(dy), np.float32(dd)
        self.dx, self.dy, self.dd = dx, dy, dd
        if REPORT:
            for name, v, unit in (('dx', dx, '[m]'), ('dy', dy, '[m]'), ('dd', dd, '[m]'),
                                  ('da', np.asarray(self.da), '[m^2]')):
                print('    min(%s), max(%s) = %s, %s %s' % (name, name, str(v.min()).ljust(8),
                                                             str(v.max()).ljust(8), unit))

    # -- static tables (:616-816) ------------------------------------------------------
    def get_flow_code_list(self, ARC=False):
        self.code_list = np.int16([128, 1, 2, 4, 8, 16, 32, 64] if ARC else
                                  [1, 2, 4, 8, 16, 32, 64, 128])

    def get_flow_code_list_opps(self, ARC=False):
        if ARC:
            self.code_opps = np.int16([8, 16, 32, 64, 128, 1, 2, 4])
            self.code_opps_3x3 = np.int16([[2, 4, 8], [1, 0, 16], [128, 64, 32]])
        else:
            self.code_opps = np.int16([16, 32, 64, 128, 1, 2, 4, 8])
            self.code_opps_3x3 = np.int16([[4, 8, 16], [2, 0, 32], [1, 128, 64]])

    def get_valid_code_map(self, SILENT=True):
        self.valid_code_map = np.zeros([256], dtype='uint8')
        self.valid_code_map[self.code_list] = np.uint8(self.code_list)

    def get_ID_grid(self, SILENT=True):
        """Lazy: the int32 calendar-ID grid is built on first access."""
        self.__dict__.pop('ID_grid', None)

    def get_parent_inc_map(self):
        nx = self.nx
        incs = np.array([-nx + 1, 1, nx + 1, nx, nx - 1, -1, -nx - 1, -nx], dtype='int32')
        self.inc_map = np.zeros([129], dtype='int32')
        self.inc_map[self.code_list] = incs

    def get_edge_IDs(self, SILENT=True):
        nx, ny = self.nx, self.ny
        top = np.arange(nx, dtype='int32')
        left = (1 + np.arange(ny - 2, dtype='int32')) * nx
        ids = np.concatenate([top, top + (ny - 1) * nx, left, left + (nx - 1)])
        self.edge_IDs = divmod(ids, nx)

    def get_not_edge_grid(self, SILENT=True):
        self.__dict__.pop('not_edge_grid', None)       # lazy, see __getattr__

    def resolve_array_cycle(self, w, t):
        v = np.asarray(w, dtype='int64')
        for _ in range(t):
            v = v * 2
            v = v + (v > 255)
            v = np.bitwise_and(v, 255)
        return v

    def get_resolve_array(self, SILENT=True, REPORT=False):
        """The RT3 tie-break table.  The GPU holds its own copy
        (csrc/tf_d8.cu:make_resolve); this one is for callers that read
        d8.resolve."""
        r = np.zeros([256], dtype='uint8')
        for value, members in ((1, (17, 85)), (2, (34, 102, 119, 238, 170)),
                               (8, (136, 153, 187, 221, 255)), (4, (68,)), (32, (51,)),
                               (128, (204,)), (1, (1, 5, 69)), (4, (4, 20, 21)),
                               (16, (16, 80, 84)), (64, (64, 65, 81))):
            r[list(members)] = value
        fam = np.array([2, 3, 6, 7, 10, 11, 14, 15, 18, 19, 22, 23, 27, 31, 35, 38, 39, 43, 47,
                        54, 55, 66, 67, 70, 71, 74, 75, 78, 79, 82, 83, 86, 87, 91, 95, 103, 107,
                        111, 134, 135, 138, 139, 142, 143, 150, 151, 155, 159, 166, 167, 171, 175,
                        183, 191, 206, 207, 223])
        for turns, value in ((0, 2), (2, 8), (4, 32), (6, 128)):
            r[self.resolve_array_cycle(fam, turns)] = value
        self.resolve = r
        if REPORT:
            print('resolve =')
            print(np.reshape(r, (16, 16)))
            print('sum(resolve) =', np.sum(r))

    # -- output files (:1004-1233) --------------------------------------------------
    def change_extension_to_rtg(self, filename):
        p = filename.find('.')
        return filename if p == -1 else filename[:p] + '.rtg'

    def update_outfile_names(self):
        for stem, _, _, _, _ in _OUTPUTS:
            gs = self.change_extension_to_rtg(getattr(self, stem + '_gs_file'))
            setattr(self, stem + '_gs_file', self.out_directory + gs)
            setattr(self, stem + '_ts_file', self.out_directory + getattr(self, stem + '_ts_file'))

    def open_output_files(self):
        model_output.check_netcdf()
        self.update_outfile_names()
        for stem, long_name, units, dtype, _ in _OUTPUTS:
            if getattr(self, 'SAVE_%s_GRIDS' % stem.upper()):
                model_output.open_new_gs_file(self, getattr(self, stem + '_gs_file'), self.rti,
                                              dtype=dtype, var_name=stem, long_name=long_name,
                                              units_name=units)
            if getattr(self, 'SAVE_%s_PIXELS' % stem.upper()):
                model_output.open_new_ts_file(self, getattr(self, stem + '_ts_file'),
                                              self.outlet_IDs, dtype=dtype, var_name=stem,
                                              long_name=long_name, units_name=units)

    def write_output_files(self, time_seconds=None):
        if time_seconds is None:
            time_seconds = self.time_sec
        model_time = int(time_seconds)
        if model_time % int(self.save_grid_dt) == 0:
            self.save_grids()
        if model_time % int(self.save_pixels_dt) == 0:
            self.save_pixel_values()

    def save_grids(self, SILENT=True, REPORT=False):
        for stem, _, _, _, attr in _OUTPUTS:
            if getattr(self, 'SAVE_%s_GRIDS' % stem.upper()):
                model_output.add_grid(self, getattr(self, attr), stem, self.time)

    def save_pixel_values(self):
        for stem, _, _, _, attr in _OUTPUTS:
            if getattr(self, 'SAVE_%s_PIXELS' % stem.upper()):
                model_output.add_values_at_IDs(self, self.time, getattr(self, attr), stem,
                                               self.outlet_IDs)

    def close_output_files(self):
        for stem, _, _, _, _ in _OUTPUTS:
            if getattr(self, 'SAVE_%s_GRIDS' % stem.upper(), False):
                model_output.close_gs_file(self, stem)
            if getattr(self, 'SAVE_%s_PIXELS' % stem.upper(), False):
                model_output.close_ts_file(self, stem)
