"""Kinematic-wave channel component: channels_base on the GPU with the
kinematic-wave velocity update (topoflow/components/channels_kinematic_wave.py).
The wave type is picked from 'cfg_extension', exactly as the reference does in
set_computed_input_vars (channels_base.py:732-736)."""
from . import channels_base


class channels_component(channels_base.channels_component):

    _att_map = {
        'model_name': 'Channels_Kinematic_Wave', 'version': '3.1',
        'author_name': 'Scott D. Peckham', 'grid_type': 'uniform',
        'time_step_type': 'fixed', 'step_method': 'explicit',
        'comp_name': 'ChannelsKinWave', 'model_family': 'TopoFlow',
        'cfg_template_file': 'Channels_Kinematic_Wave.cfg.in',
        'cfg_extension': '_channels_kinematic_wave.cfg',
        'cmt_var_prefix': '/ChannelsKinWave/Input/Var/',
        'gui_xml_file': '/home/csdms/cca/topoflow/3.1/src/share/cmt/gui/Channels_Kinematic_Wave.xml',
        'dialog_title': 'Channels: Kinematic Wave Parameters', 'time_units': 'seconds'}

    def get_component_name(self):
        return 'TopoFlow_Channels_Kinematic_Wave'
