"""D8 component, whole-grid ("global") flow algorithms on the GPU.

Same class and method names as topoflow/components/d8_global.py
(d8_component, :57-1316).  The work is done by libtopoflow_b200.so; this class
keeps the host views other components read (d8_grid, ds, dw, A,
parent_ID_grid, parent_IDs, noflow_IDs, w1..w8, p1..p8, ...).  The index lists
are built on first access only: the device routing path never needs them (it
uses the one-byte child mask instead) and at 16384^2 they would be 17 GB.
"""
import numpy as np

from .. import _lib as L
from ..engine import Strip
from . import d8_base

_LAZY = ('ID_grid', 'not_edge_grid', 'parent_IDs', 'noflow_IDs')


class d8_component(d8_base.d8_component):

    _att_map = {
        'model_name': 'D8_Global', 'version': '3.5', 'author_name': 'Scott D. Peckham',
        'grid_type': 'uniform', 'time_step_type': 'fixed', 'step_method': 'explicit',
        'comp_name': 'D8Global', 'model_family': 'TopoFlow',
        'cfg_template_file': 'd8_global.cfg.in', 'cfg_extension': '_d8_global.cfg',
        'cmt_var_prefix': '/D8Global/Input/Var/',
        # (the reference's d8_global does name the channels dialog here, d8_global.py:74)
        'gui_xml_file': '/home/csdms/cca/topoflow/3.1/src/share/cmt/gui/Channels_Kinematic_Wave.xml',
        'dialog_title': 'D8 Global Parameters', 'time_units': 'seconds'}

    def get_component_name(self):
        return 'TopoFlow_D8_Global'

    # -- device strip -------------------------------------------------------------
    @property
    def strip(self):
        """The device strip holding this grid (created on first use; the
        channels component routes on the same strip)."""
        s = self.__dict__.get('_strip')
        if s is None:
            s = Strip(self.nx, self.ny, device=getattr(self, 'device', 0))
            self.__dict__['_strip'] = s
        return s

    def finalize(self):
        d8_base.d8_component.finalize(self)
        s = self.__dict__.pop('_strip', None)
        if s is not None and not getattr(self, 'strip_is_shared', False):
            s.close()

    # -- lazily built host structures ------------------------------------------------
    def __getattr__(self, name):
        # only called when normal lookup fails
        if name == 'ID_grid':                                   # d8_base.py:664-675
            v = np.reshape(np.arange(self.nx * self.ny, dtype='int32'), [self.ny, self.nx])
        elif name == 'not_edge_grid':                           # d8_base.py:728-735
            v = np.ones([self.ny, self.nx], dtype='uint8')
            v[self.edge_IDs] = 0
        elif name == 'parent_IDs':                              # d8_global.py:648-675
            v = divmod(self.parent_ID_grid, self.nx)
        elif name == 'noflow_IDs':                              # :839-859
            v = np.where(self.d8_grid <= 0)
            if v[0].size == 0:
                v = self.edge_IDs
        elif len(name) == 2 and name[0] in 'wpn' and name[1] in '12345678' and 'd8_grid' in self.__dict__:
            self._build_flow_lists()
            return self.__dict__[name]
        elif name.endswith('_OK') and name[:2] in self._p_names() and 'd8_grid' in self.__dict__:
            self._build_flow_lists()
            return self.__dict__[name]
        else:
            raise AttributeError(name)
        self.__dict__[name] = v
        return v

    @staticmethod
    def _p_names():
        return ['p%d' % k for k in range(1, 9)]

    def _build_flow_lists(self):
        """w_k = where(code == g_k) in row-major order, p_k their parents,
        n_k, p_k_OK (d8_global.py:719-835)."""
        for k in range(8):
            w = np.where(self.d8_grid == self.code_list[k])
            n = w[0].size
            p = divmod(self.parent_ID_grid[w], self.nx) if n else None
            self.__dict__['w%d' % (k + 1)] = w
            self.__dict__['n%d' % (k + 1)] = n
            self.__dict__['p%d' % (k + 1)] = p
            self.__dict__['p%d_OK' % (k + 1)] = (p is not None)

    def _forget_derived(self):
        for name in _LAZY[2:]:
            self.__dict__.pop(name, None)
        for k in range(1, 9):
            for stem in ('w%d', 'n%d', 'p%d', 'p%d_OK'):
                self.__dict__.pop(stem % k, None)

    # -- flow grid (d8_global.py:97-567) ----------------------------------------------
    def update_flow_grid(self, DEM=None, SILENT=True, REPORT=False):
        """start_new_d8_codes + break_flow_grid_ties + link_flats + the uint8
        cast, in one call to the library."""
        if not SILENT:
            print('Updating D8 flow grid...')
        if DEM is None:
            DEM = self.DEM
        stats = self.strip.d8_build(np.asarray(DEM, dtype=np.float32), self.dx, self.dy, self.dd,
                                    nodata=float(self.nodata), link_flats=bool(self.LINK_FLATS),
                                    lr_periodic=bool(self.LR_PERIODIC),
                                    tb_periodic=bool(self.TB_PERIODIC))
        self.d8_stats = stats
        self.total_flats = stats['n_flats']
        self.d8_grid = self.strip.d8_get(L.D8_CODE)
        self._forget_derived()
        if REPORT:
            print('   --------------------------------------------')
            print('   Data type of flow grid at start = int16')
            if self.LINK_FLATS:
                print('   Number of flats         =', stats['n_flats'])
                print('   Number of 1-pixel pits  =', stats['n_pits'])
            print('   Number of nodata/NaN    =', stats['n_bad'])
            print('   --------------------------------------------')
        if self.LINK_FLATS and not SILENT:
            print('   Number of iterations =', stats['n_reps'], ' (in link_flats())')

    # the reference exposes the three stages separately; here they are one pass
    def start_new_d8_codes(self, DEM=None, SILENT=True, REPORT=False):
        self.update_flow_grid(DEM=DEM, SILENT=SILENT, REPORT=REPORT)

    def break_flow_grid_ties(self, SILENT=True):
        pass

    def link_flats(self, SILENT=True):
        pass

    def read_flow_grid(self, code_file, SILENT=True):
        """Use a stored flow grid (e.g. <site>_flow.rtg) instead of deriving one."""
        from ..utils import rtg_files
        codes = rtg_files.read_grid(code_file, self.rti, RTG_type='BYTE')
        self.strip.d8_set_codes(codes, self.dx, self.dy, self.dd,
                                lr_periodic=bool(self.LR_PERIODIC),
                                tb_periodic=bool(self.TB_PERIODIC))
        self.d8_grid = self.strip.d8_get(L.D8_CODE)
        self._forget_derived()

    # -- derived grids: fetched from the device ------------------------------------------
    def update_parent_ID_grid(self, SILENT=True):             # :574-644
        self.parent_ID_grid = self.strip.d8_get(L.D8_PARENT_ID)

    def update_parent_IDs(self):                              # :648-675 (lazy)
        self.__dict__.pop('parent_IDs', None)

    def update_flow_from_IDs(self):                           # :719-754 (lazy)
        pass

    def update_flow_to_IDs(self):                             # :762-835 (lazy)
        pass

    def update_noflow_IDs(self):                              # :839-859 (lazy)
        self.__dict__.pop('noflow_IDs', None)

    def update_flow_width_grid(self, DOUBLE=False, METHOD2=False, SILENT=True, REPORT=False):
        if METHOD2:
            raise NotImplementedError('METHOD2 flow widths are not used by TopoFlow 3.5')
        if not SILENT:
            print('Updating flow width grid...')
        self.dw = self.strip.d8_get(L.D8_DW)                  # :863-935
        if REPORT:
            print('    min(dw), max(dw) = ' + str(self.dw.min()) + ', ' + str(self.dw.max()) + ' [m]')

    def update_flow_length_grid(self, DOUBLE=False, SILENT=True, REPORT=False):
        if not SILENT:
            print('Updating flow length grid...')
        self.ds = self.strip.d8_get(L.D8_DS)                  # :949-1004
        if REPORT:
            print('    min(ds), max(ds) = ' + str(self.ds.min()) + ', ' + str(self.ds.max()) + ' [m]')

    def update_area_grid(self, SILENT=True, REPORT=False):    # :1018-1261
        if not SILENT:
            print('Updating upstream area grid...')
        da = np.asarray(self.da, dtype='float64')
        da_rows = da[:, 0] if da.ndim == 2 else da
        n_bad = self.strip.d8_area(da_rows, units_km=('km' in self.A_units.lower()))
        self.A = self.strip.d8_get(L.D8_AREA)
        if n_bad:
            print('Upstream area not defined for all pixels.')
        if REPORT:
            unit = ' [km^2]' if 'km' in self.A_units.lower() else ' [m^2]'
            print('    min(A), max(A) = ' + str(self.A.min()) + ', ' + str(self.A.max()) + unit)
