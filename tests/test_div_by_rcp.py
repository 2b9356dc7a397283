"""div_by_rcp (csrc/tf_chan_math.cuh) replaces the two divisions of the trapezoid
depth by a multiply and two fused multiply-adds with a precomputed reciprocal.
It is only allowed because it returns the IEEE quotient bit for bit; this runs
the same three operations on the host (gcc, fma()) against a / b."""
import os
import shutil
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.skipif(shutil.which('gcc') is None, reason='needs gcc')
def test_reciprocal_division_is_the_ieee_quotient(tmp_path):
    exe = str(tmp_path / 'div_by_rcp_check')
    subprocess.check_call(['gcc', '-O2', '-ffp-contract=off', '-o', exe,
                           os.path.join(HERE, 'host', 'div_by_rcp_check.c'), '-lm'])
    out = subprocess.run([exe, '4000000'], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout
    assert out.stdout.strip() == 'bad 0 of 12000000'
