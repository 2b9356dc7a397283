#!/usr/bin/env python
"""Copy the reference's own fixtures for the D8 + channel-routing path into
tests/golden/ (run once, in the container that has /root/reference).

Nothing here is reference SOURCE: these are the example input grids/configs
and the golden OUTPUT files shipped with TopoFlow 3.5
(docs/00_TF_3.5_Release_Notes.txt:69-72).  The GPU box has no /root/reference,
so the tests read these copies.
"""
import json
import os
import re
import shutil

REF = '/root/reference/topoflow/examples'
HERE = os.path.dirname(os.path.abspath(__file__))

TREYNOR_IN = ['Treynor.rti', 'Treynor_DEM.rtg', 'Treynor_flow.rtg',
              'Treynor_area.rtg', 'Treynor_slope.rtg', 'Treynor_chan-n.rtg',
              'Treynor_chan-w.rtg', 'Treynor_chan-a.rtg', 'Treynor_chan-z0.rtg',
              'June_20_67_channels_kinematic_wave.cfg',
              'June_20_67_channels_diffusive_wave.cfg',
              'June_20_67_channels_dynamic_wave.cfg',
              'June_20_67_d8_global.cfg', 'June_20_67_path_info.cfg',
              'June_20_67_outlets.txt', 'June_20_67_rain_rates.txt',
              'June_20_67_meteorology.cfg', 'June_20_67_infil_green_ampt.cfg',
              'June_20_67_snow_degree_day.cfg']
TREYNOR_OUT = {'Test1_kin_wave': ['June_20_67_0D-Q.txt', 'June_20_67_2D-Q.rti', 'June_20_67_2D-Q.rts'],
               'Test2_diff_wave': ['June_20_67_0D-Q.txt', 'June_20_67_2D-Q.rts'],
               'Test2_diff_wave_dt=6s': ['June_20_67_0D-Q.txt', 'June_20_67_2D-Q.rts'],
               'Test12_dynam_wave': ['June_20_67_0D-Q.txt'],
               'Test8_snow_degree_day': ['June_20_67_0D-Q.txt']}
PLANE_IN = ['plane.rti', 'plane_DEM.rtg', 'plane_slope.rtg', 'plane_flow.rtg',
            'plane_chan-a.rtg', 'plane_chan-n.rtg', 'plane_chan-w.rtg',
            'Rain_100mmph_on_right.rtg', 'Rain_100mmph_on_right.rti',
            'plane_canals.txt', 'Case1_channels_kinematic_wave.cfg',
            'Case1_channels_diffusive_wave.cfg',
            'Case1_channels_dynamic_wave.cfg', 'Case1_d8_global.cfg',
            'Case1_path_info.cfg', 'Case1_outlets.txt',
            'Case1_diversions_fraction_method.cfg', 'Case1_meteorology.cfg',
            'Case1_topoflow.cfg']
PLANE_OUT = ['Case1_0D-Q.txt', 'Case1_0D-d.txt', 'Case1_0D-u.txt',
             'Case1_2D-Q.rts', 'Case1_2D-Q.rti']


def cp(src, dst):
    os.makedirs(os.path.dirname(dst), exist_ok=True)
    shutil.copyfile(src, dst)
    os.chmod(dst, 0o644)


def log_scalars(path):
    """Pull the numbers the parity tests assert on out of a console log."""
    txt = open(path, errors='replace').read()
    pats = {'n_flats': r'Number of flats\s+=\s+(\d+)',
            'n_pits': r'Number of 1-pixel pits\s+=\s+(\d+)',
            'link_iters': r'Number of iterations =\s+(\d+)',
            'n_zero_slope': r'(\d+) grid cells have slope = 0',
            'S_min': r'min\(S\) = ([0-9.eE+-]+)',
            'S_max': r'max\(S\) = ([0-9.eE+-]+)',
            'Q_peak': r'Q_peak:\s+([0-9.eE+-]+)',
            'T_peak': r'Q_peak_time:\s+([0-9.eE+-]+)',
            'u_peak': r'u_peak:\s+([0-9.eE+-]+)',
            'd_peak': r'd_peak:\s+([0-9.eE+-]+)',
            'Tu_peak': r'u_peak_time:\s+([0-9.eE+-]+)',
            'Td_peak': r'd_peak_time:\s+([0-9.eE+-]+)',
            'Q_final': r'Q_final:\s+([0-9.eE+-]+)',
            'vol_P': r'vol_P\s+\(rainfall\):\s+([0-9.eE+-]+)',
            'vol_IN': r'vol_IN\s+\(infiltration\):\s+([0-9.eE+-]+)',
            'vol_Q': r'vol_Q\s+\(discharge\):\s+([0-9.eE+-]+)',
            'vol_R': r'vol_R\s+\(runoff\):\s+([0-9.eE+-]+)',
            'vol_SM': r'vol_SM\s+\(snowmelt\):\s+([0-9.eE+-]+)',
            'n_steps': r'Number of timesteps:\s+(\d+)'}
    out = {}
    for k, pat in pats.items():
        m = re.search(pat, txt)
        if m:
            out[k] = float(m.group(1))
    return out


def bmi_tables():
    """The introspection tables of the reference's channels components, read
    out of its source as literals (the files are Python 2 and cannot be
    imported): what get_input_var_names / get_var_name / get_var_units /
    get_attribute must answer."""
    import ast
    comp = '/root/reference/topoflow/components/'

    def literal(path, name):
        lines = open(comp + path).read().splitlines()
        start = next(i for i, l in enumerate(lines) if l.startswith('    %s = ' % name))
        body, depth = '', 0
        for l in lines[start:]:
            l = l.split('#')[0]
            if not body:
                l = l.split('=', 1)[1]
            body += l + '\n'
            depth += l.count('[') + l.count('{') - l.count(']') - l.count('}')
            if depth == 0 and body.strip():
                break
        return ast.literal_eval(body.strip())

    out = {'input_var_names': literal('channels_base.py', '_input_var_names'),
           'output_var_names': literal('channels_base.py', '_output_var_names'),
           'var_name_map': literal('channels_base.py', '_var_name_map'),
           'var_units_map': literal('channels_base.py', '_var_units_map'),
           'att_map': {w: literal('channels_%s_wave.py' % w, '_att_map')
                       for w in ('kinematic', 'diffusive', 'dynamic')},
           'd8_att_map': literal('d8_global.py', '_att_map')}
    return out


def main():
    t = os.path.join(REF, 'Treynor_Iowa_30m')
    for f in TREYNOR_IN:
        cp(os.path.join(t, f), os.path.join(HERE, 'treynor', f))
    scal = {}
    for case, files in TREYNOR_OUT.items():
        for f in files:
            cp(os.path.join(t, '00_Output_Files', case, f),
               os.path.join(HERE, 'treynor', 'out_' + case.replace('=', ''), f))
        scal[case] = log_scalars(os.path.join(t, '00_Output_Files', case,
                                              '00_Full_Output.txt'))
    json.dump(scal, open(os.path.join(HERE, 'treynor', 'log_scalars.json'), 'w'),
              indent=1, sort_keys=True)
    json.dump(bmi_tables(), open(os.path.join(HERE, 'bmi_tables.json'), 'w'), indent=1, sort_keys=True)
    p = os.path.join(REF, 'Test_Plane_Canal')
    for f in PLANE_IN:
        cp(os.path.join(p, f), os.path.join(HERE, 'plane_canal', f))
    for f in PLANE_OUT:
        cp(os.path.join(p, '00_Output_Files', f),
           os.path.join(HERE, 'plane_canal', 'out', f))


if __name__ == '__main__':
    main()
