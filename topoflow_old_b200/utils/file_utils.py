"""File-name helpers (topoflow/utils/file_utils.py)."""
import os


def replace_extension(file_name, extension='.nc'):
    dot = file_name.rfind('.')
    return (file_name[:dot] if dot != -1 else file_name) + extension


def check_overwrite(file_name):
    """The reference appends '_1', '_2', ... rather than overwrite
    (file_utils.py:63-...); keep that: runs never clobber earlier output."""
    if not os.path.exists(file_name):
        return file_name
    dot = file_name.rfind('.')
    prefix, ext = (file_name[:dot], file_name[dot:]) if dot != -1 else (file_name, '')
    n = 1
    while os.path.exists('%s_%d%s' % (prefix, n, ext)):
        n += 1
    return '%s_%d%s' % (prefix, n, ext)
