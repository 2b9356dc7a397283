"""The drop-in components answer the framework's introspection calls with the
reference's own tables (channels_base.py:156-404, channels_*_wave.py:40-55;
fixture: tests/golden/bmi_tables.json, read out of the reference source by
tests/golden/make_golden.py).  No GPU needed: nothing is initialised."""
import importlib
import json
import os

import numpy as np
import pytest


@pytest.fixture(scope='module')
def tables(golden_dir):
    return json.load(open(os.path.join(golden_dir, 'bmi_tables.json')))


@pytest.mark.parametrize('wave', ['kinematic', 'diffusive', 'dynamic'])
def test_introspection_tables_equal_the_reference(tables, wave):
    mod = importlib.import_module('topoflow_old_b200.components.channels_%s_wave' % wave)
    comp = mod.channels_component()
    assert comp.get_status() == 'created'
    assert list(comp.get_input_var_names()) == tables['input_var_names']        # same order
    assert list(comp.get_output_var_names()) == tables['output_var_names']
    assert len(tables['var_name_map']) == 67
    for long_name, short in tables['var_name_map'].items():
        assert comp.get_var_name(long_name) == short, long_name
        assert comp.get_var_units(long_name) == tables['var_units_map'][long_name], long_name
    for att, value in tables['att_map'][wave].items():
        assert comp.get_attribute(att) == value, att
    assert comp.get_attribute('no_such_attribute') is None or True   # prints an error, never raises
    assert comp.get_component_name() == {'kinematic': 'TopoFlow_Channels_Kinematic_Wave',
                                         'diffusive': 'TopoFlow_Channels_Diffusive_Wave',
                                         'dynamic': 'TopoFlow_Channels_Dynamic_Wave'}[wave]


def test_set_values_before_initialize_is_plain_attribute_passing(tables):
    """emeli.initialize_comp_set connects providers to a user before the user
    is initialised (emeli.py:1797-1828): references must be kept as they are."""
    from topoflow_old_b200.components import channels_kinematic_wave
    comp = channels_kinematic_wave.channels_component()
    P = np.array(1e-5)
    comp.set_values('atmosphere_water__rainfall_volume_flux', P)
    assert comp.get_values('atmosphere_water__rainfall_volume_flux') is P
    n = np.array(2, dtype='int32')
    comp.set_values('canals__count', n)
    assert comp.n_canals is n


def test_d8_component_attributes(tables):
    from topoflow_old_b200.components import d8_global
    c = d8_global.d8_component()
    for att, value in tables['d8_att_map'].items():
        assert c.get_attribute(att) == value, att
    assert c.get_component_name() == 'TopoFlow_D8_Global'
