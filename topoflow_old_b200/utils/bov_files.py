"""Brick-of-values header next to an RTS file (topoflow/utils/bov_files.py)."""
import sys

from . import rti_files


def write_info_as_bov(file_name, info, var_name='X', time=0.0):
    bov_file = rti_files.get_file_prefix(file_name) + '.bov'
    dims = '%s %s 1' % (info.ncols, info.nrows)
    rows = (('TIME:', str(time)), ('DATA_FILE:', file_name), ('DATA_SIZE:', dims),
            ('DATA_FORMAT:', info.data_type), ('VARIABLE:', var_name),
            ('DATA_ENDIAN:', sys.byteorder.upper()), ('CENTERING:', 'zonal'),
            ('BRICK_ORIGIN:', '%s %s 0.0' % (info.x_west_edge, info.y_south_edge)),
            ('BRICK_SIZE:', dims))
    with open(bov_file, 'w') as out:
        for key, val in rows:
            out.write(key.ljust(15) + val + '\n')
