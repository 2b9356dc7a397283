"""Row-strip decomposition of a grid over several GPUs / processes.

The routing step needs, from the strip to the north and to the south, one row
of the rasters it reads with a one-cell halo.  Strips are stacked north to
south in rank order; a strip keeps its neighbours' edge rows in apron rows
(row -1 and row ny of every device raster).  The reference is single-process
and has no counterpart of this module.
"""


def strip_rows(global_ny, nranks, rank):
    """(first_row, n_rows) of strip `rank`: rows are dealt out as evenly as
    possible, the first `global_ny % nranks` strips get one more."""
    if not (0 <= rank < nranks):
        raise ValueError('rank %d of %d' % (rank, nranks))
    if global_ny < 3 * nranks:
        raise ValueError('%d rows cannot be cut into %d strips of >= 3 rows' % (global_ny, nranks))
    base, extra = divmod(global_ny, nranks)
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def neighbours(nranks, rank):
    """(north_rank or None, south_rank or None)."""
    return (rank - 1 if rank > 0 else None, rank + 1 if rank < nranks - 1 else None)


def halo_plan(wave):
    """Which rasters cross a strip boundary in one time step, in order -- read
    from the library's own table (tf_halo_plan), the one its launch sequence
    follows:

    kinematic : the next step's Q (u * A_wet) once per step;
    diffusive : the new depth d between the two phases (the free-surface slope
                reads the parent's depth), then next Q;
    dynamic   : d between the phases, then next Q and the new u (the momentum
                flux reads the children's u and Q).
    Returns a list of (point_in_step, [raster names])."""
    import ctypes as C
    from . import _lib as L
    waves = {'kinematic': L.KINEMATIC, 'diffusive': L.DIFFUSIVE, 'dynamic': L.DYNAMIC}
    if wave not in waves:
        raise ValueError(wave)
    names = {L.F_D: 'd', L.F_QNEXT: 'Qnext', L.F_U: 'u'}
    plan = []
    for point, label in ((0, 'after_depth'), (1, 'end')):
        buf = (C.c_int * 3)()
        L.check(L.lib.tf_halo_plan(waves[wave], point, buf))
        fields = [names[f] for f in buf if f >= 0]
        if fields:
            plan.append((label, fields))
    return plan
