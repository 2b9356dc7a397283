"""Host-side mirror of the reference's file formats and BMI plumbing (no GPU)."""
import os
import shutil

import numpy as np
import pytest

from topoflow_old_b200.utils import (BMI_base, model_input, outlets, pixels, rtg_files,
                                     rti_files, rts_files, text_ts_files)


def test_rti_read_matches_reference_file(golden_dir):
    info = rti_files.read_info(os.path.join(golden_dir, 'treynor', 'Treynor.rti'))
    assert (info.ncols, info.nrows) == (29, 44)
    assert info.data_type == 'FLOAT' and info.byte_order == 'MSB' and info.pixel_geom == 1
    assert info.xres == 30.0 and info.yres == 30.0 and info.UTM_zone == '15'
    assert info.n_pixels == 1276 and info.SWAP_ENDIAN
    assert float(pixels.get_da(info)) == 900.0
    dx, dy, dd = pixels.get_sizes_by_row(info, METERS=True)
    assert dx.shape == (44,) and np.float32(dd[0]) == np.float32(np.sqrt(1800.0))
    # try_to_find strips "_suffix" parts: Treynor_chan-n.rtg -> Treynor.rti
    assert rti_files.try_to_find_rti_file(
        os.path.join(golden_dir, 'treynor', 'Treynor_DEM.rtg')).endswith('Treynor.rti')


def test_rti_write_read_round_trip(tmp_path):
    info = rti_files.make_info('x.rtg', ncols=7, nrows=5, xres=10.0, yres=20.0,
                               byte_order='MSB', x_west_edge=100.5, y_south_edge=-3.25)
    p = str(tmp_path / 'x.rti')
    rti_files.write_info(p, info)
    back = rti_files.read_info(p)
    for k in ('ncols', 'nrows', 'xres', 'yres', 'byte_order', 'data_type', 'x_west_edge',
              'x_east_edge', 'y_south_edge', 'y_north_edge', 'pixel_geom'):
        assert getattr(back, k) == getattr(info, k), k
    assert open(p).readline().strip() == 'RiverTools Info File'


def test_rtg_and_rts_byte_order(tmp_path, golden_dir):
    info = rti_files.read_info(os.path.join(golden_dir, 'treynor', 'Treynor.rti'))
    g = rtg_files.read_grid(os.path.join(golden_dir, 'treynor', 'Treynor_DEM.rtg'), info)
    assert g.dtype == np.float32 and abs(g.min() - 335.0505) < 1e-3
    out = str(tmp_path / 'copy.rtg')
    rtg_files.write_grid(g, out, info)
    assert open(out, 'rb').read() == open(os.path.join(golden_dir, 'treynor', 'Treynor_DEM.rtg'), 'rb').read()
    shutil.copy(os.path.join(golden_dir, 'treynor', 'Treynor.rti'), str(tmp_path / 'stack.rti'))
    s = rts_files.rts_file()
    assert s.open_new_file(str(tmp_path / 'stack.rts'), info, 'Q', MAKE_BOV=True)
    s.add_grid(np.float64(g))
    s.add_grid(np.array(2.5))                       # scalar -> constant grid
    s.close()
    r = rts_files.rts_file()
    assert r.open_file(str(tmp_path / 'stack.rts'))
    assert r.number_of_grids() == 2
    assert np.array_equal(r.get_grid(0), g) and np.all(r.get_grid(1) == np.float32(2.5))
    r.close()
    assert os.path.exists(str(tmp_path / 'stack.bov'))
    # a second file of the same name gets a version suffix, never overwritten
    s2 = rts_files.rts_file()
    s2.open_new_file(str(tmp_path / 'stack.rts'), info, 'Q')
    s2.close()
    assert s2.file_name.endswith('stack_1.rts')


def test_text_time_series_format_matches_golden(tmp_path, golden_dir):
    gold = open(os.path.join(golden_dir, 'plane_canal', 'out', 'Case1_0D-Q.txt')).read().splitlines()
    t = text_ts_files.ts_file()
    names = gold[0].split()[2:]
    t.open_new_file(str(tmp_path / 'q.txt'), names, time_units='minutes')
    for row in gold[2:6]:
        vals = [float(x) for x in row.split()]
        t.add_values(vals[0], np.float32(vals[1:]))
    t.close()
    mine = open(str(tmp_path / 'q.txt')).read().splitlines()
    assert mine[:6] == gold[:6]


class _Comp(BMI_base.BMI_component):
    _att_map = dict(BMI_base.BMI_component._DEFAULT_ATTS,
                    cfg_extension='_channels_kinematic_wave.cfg')


def test_cfg_grammar(golden_dir):
    c = _Comp()
    c.cfg_file = os.path.join(golden_dir, 'treynor', 'June_20_67_channels_kinematic_wave.cfg')
    c.mode = 'nondriver'
    c.initialize_config_vars()
    assert c.site_prefix == 'Treynor' and c.case_prefix == 'June_20_67'
    assert c.in_directory == os.path.join(golden_dir, 'treynor') + os.sep     # "." -> cfg dir
    assert c.out_directory.endswith('TF_Output/Treynor/') and '~' not in c.out_directory
    assert c.dt.ndim == 0 and c.dt.dtype == np.float64 and float(c.dt) == 2.0
    assert c.MANNING.dtype == np.int32 and int(c.MANNING) == 1 and int(c.LAW_OF_WALL) == 0
    assert c.nval_type == 'Grid' and c.nval_file == 'Treynor_chan-n.rtg' and c.nval.ndim == 0
    assert c.z0val_type == 'Scalar' and c.z0val_file == '' and float(c.z0val) == 0.0001
    assert c.SAVE_Q_GRIDS is True and c.SAVE_U_GRIDS is False
    assert c.Q_gs_file == 'June_20_67_2D-Q.nc' and c.pixel_file == 'June_20_67_outlets.txt'
    # 0-d scalars are mutable references (BMI_base.py:1941-1972)
    ref = c.dt
    c.update_var('dt', 3.0)
    assert float(ref) == 3.0
    g = np.zeros((2, 2))
    c.grid = g
    c.update_var('grid', np.ones((2, 2)))
    assert g.sum() == 4


def test_outlets_and_missing_file(golden_dir, tmp_path):
    c = _Comp()
    c.in_directory = os.path.join(golden_dir, 'treynor') + os.sep
    c.case_prefix = 'June_20_67'
    c.nx = 29
    c.rti = rti_files.read_info(os.path.join(golden_dir, 'treynor', 'Treynor.rti'))
    outlets.read_outlet_file(c)
    assert c.n_outlets == 5 and c.outlet_ID == (34, 16)
    assert list(c.outlet_IDs[0]) == [34, 32, 37, 26, 30]
    assert c.basin_area == 0.396
    c.in_directory = str(tmp_path) + os.sep
    with pytest.raises(IOError, match='Could not find outlet_file.'):
        outlets.read_outlet_file(c)


def test_model_input_streams(golden_dir, tmp_path):
    info = rti_files.read_info(os.path.join(golden_dir, 'treynor', 'Treynor.rti'))
    assert model_input.open_file('Scalar', 'whatever') is None
    u = model_input.open_file('Grid', os.path.join(golden_dir, 'treynor', 'Treynor_chan-n.rtg'))
    g = model_input.read_next(u, 'Grid', info)
    assert g.dtype == np.float64 and g.shape == (44, 29)
    assert g.min() == np.float64(np.float32(0.03))
    assert model_input.read_next(u, 'Grid', info) is None        # end of file
    u.close()
    p = tmp_path / 'ts.txt'
    p.write_text('5.588\n0.0\n')
    u = model_input.open_file('Time_Series', str(p))
    v = model_input.read_next(u, 'Time_Series', info, factor=np.float64(1) / np.float64(3600000))
    assert v == np.float64(np.float32(5.588)) * (np.float64(1) / np.float64(3600000))
    u.close()


def test_time_bookkeeping():
    c = _Comp()
    c.mode = 'nondriver'
    c.dt = c.initialize_scalar(2.0)
    c.initialize_time_vars()
    for _ in range(45):
        c.update_time(-1.0)
    assert c.time_index == 45 and float(c.time) == 90.0 and float(c.time_min) == 1.5
    assert c.get_status() == 'created'
    assert c.get_run_time_string(seconds=3725.5).startswith('1 hour, 2 minutes, 5.5')
